"""ctypes binding of ``librnb_b200.so`` (the C ABI declared in ``include/rnb_b200.h``).

There is no fallback: if the library has not been built, or the process has no sm_100 GPU, every
entry point that computes raises.  Build with ``python __graft_entry__.py`` or
``make -C rnb_b200/csrc``.
"""
import ctypes
import os
import threading

OK = 0
DTYPE_F32, DTYPE_BF16 = 0, 1
FMT_U8_NTCHW, FMT_F32_NCDHW, FMT_BF16_NDHWC, FMT_F32_LOGITS = 0, 1, 2, 3

LIB_PATH = os.path.join(os.path.dirname(os.path.abspath(__file__)), 'csrc', 'librnb_b200.so')

_lock = threading.Lock()
_lib = None

_c_int_p = ctypes.POINTER(ctypes.c_int)
_c_i64_p = ctypes.POINTER(ctypes.c_int64)
_c_f32_p = ctypes.POINTER(ctypes.c_float)

# name -> (restype, argtypes); must list every symbol include/rnb_b200.h declares
PROTOTYPES = {
    'rnb_init': (ctypes.c_int, [ctypes.c_int]),
    'rnb_r2p1d_create': (ctypes.c_int, [_c_int_p, ctypes.c_int, ctypes.c_int, ctypes.c_int,
                                        ctypes.c_int, ctypes.c_int, ctypes.c_int,
                                        ctypes.POINTER(ctypes.c_void_p)]),
    'rnb_r2p1d_load_param': (ctypes.c_int, [ctypes.c_void_p, ctypes.c_char_p, ctypes.c_void_p,
                                            _c_i64_p, ctypes.c_int]),
    'rnb_r2p1d_finalize': (ctypes.c_int, [ctypes.c_void_p]),
    'rnb_r2p1d_forward': (ctypes.c_int, [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int,
                                         ctypes.c_int, ctypes.c_void_p, ctypes.c_int,
                                         ctypes.c_void_p]),
    'rnb_r2p1d_prepare': (ctypes.c_int, [ctypes.c_void_p, ctypes.c_int, ctypes.c_void_p]),
    'rnb_r2p1d_force_kernel': (ctypes.c_int, [ctypes.c_void_p] + [ctypes.c_int] * 8),
    'rnb_r2p1d_plan_entry': (ctypes.c_int, [ctypes.c_void_p, ctypes.c_int, ctypes.c_int, _c_int_p]),
    'rnb_r2p1d_classify': (ctypes.c_int, [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int,
                                          ctypes.c_int, ctypes.c_void_p, ctypes.c_void_p,
                                          ctypes.c_void_p, ctypes.c_void_p]),
    'rnb_pool_fc_classify_scratch_bytes': (ctypes.c_size_t, [ctypes.c_int, ctypes.c_int]),
    'rnb_pool_fc_classify': (ctypes.c_int, [ctypes.c_void_p, ctypes.c_int, ctypes.c_int,
                                            ctypes.c_int, ctypes.c_void_p, ctypes.c_void_p,
                                            ctypes.c_int, ctypes.c_void_p, ctypes.c_void_p,
                                            ctypes.c_void_p, ctypes.c_void_p, ctypes.c_size_t,
                                            ctypes.c_void_p]),
    'rnb_pool_fc_classify_static': (ctypes.c_int, [ctypes.c_void_p, ctypes.c_int, ctypes.c_int,
                                                   ctypes.c_int, ctypes.c_void_p, ctypes.c_void_p,
                                                   ctypes.c_int, ctypes.c_void_p, ctypes.c_void_p,
                                                   ctypes.c_void_p, ctypes.c_void_p]),
    'rnb_r2p1d_profile': (ctypes.c_int, [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int,
                                         ctypes.c_int, ctypes.c_void_p, ctypes.c_int,
                                         ctypes.c_void_p, _c_f32_p, ctypes.c_int]),
    'rnb_r2p1d_op_info': (ctypes.c_int, [ctypes.c_void_p, ctypes.c_int, ctypes.c_char_p,
                                         ctypes.c_int, ctypes.POINTER(ctypes.c_double),
                                         ctypes.POINTER(ctypes.c_double),
                                         ctypes.POINTER(ctypes.c_double), _c_int_p]),
    'rnb_r2p1d_io_shape': (ctypes.c_int, [ctypes.c_void_p, ctypes.c_int, _c_i64_p]),
    'rnb_r2p1d_flops': (ctypes.c_double, [ctypes.c_void_p, ctypes.c_int]),
    'rnb_r2p1d_num_launches': (ctypes.c_int, [ctypes.c_void_p]),
    'rnb_r2p1d_num_convs': (ctypes.c_int, [ctypes.c_void_p]),
    'rnb_r2p1d_destroy': (None, [ctypes.c_void_p]),
    'rnb_preprocess_u8': (ctypes.c_int, [ctypes.c_void_p, ctypes.c_int, _c_f32_p, _c_f32_p,
                                         ctypes.c_void_p, ctypes.c_int, ctypes.c_void_p]),
    'rnb_batch_gather': (ctypes.c_int, [ctypes.POINTER(ctypes.c_void_p), _c_int_p, ctypes.c_int,
                                        ctypes.c_void_p, ctypes.c_size_t, ctypes.c_void_p]),
    'rnb_logits_sum_argmax': (ctypes.c_int, [ctypes.c_void_p, ctypes.c_int, ctypes.c_int,
                                             ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p]),
    'rnb_peer_copy': (ctypes.c_int, [ctypes.c_void_p, ctypes.c_int, ctypes.c_void_p, ctypes.c_int,
                                     ctypes.c_size_t, ctypes.c_void_p]),
    'rnb_peer_copy_fenced': (ctypes.c_int, [ctypes.c_void_p, ctypes.c_int, ctypes.c_void_p,
                                            ctypes.c_int, ctypes.c_size_t, ctypes.c_void_p,
                                            ctypes.c_void_p]),
    'rnb_last_error': (ctypes.c_char_p, []),
}


class RnbError(RuntimeError):
    """A C-ABI call returned a negative code."""

    def __init__(self, code, message):
        super().__init__('rnb_b200 error %d: %s' % (code, message))
        self.code = code


def load():
    """dlopen the library once and attach prototypes.  Raises if it is missing."""
    global _lib
    with _lock:
        if _lib is not None:
            return _lib
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(
                'rnb_b200: %s is missing -- build it with `python __graft_entry__.py` or '
                '`make -C rnb_b200/csrc`; there is no CPU or PyTorch fallback' % LIB_PATH)
        lib = ctypes.CDLL(LIB_PATH)
        for name, (restype, argtypes) in PROTOTYPES.items():
            fn = getattr(lib, name)  # AttributeError here means header and library disagree
            fn.restype = restype
            fn.argtypes = argtypes
        _lib = lib
        return _lib


def check(code):
    if code != OK:
        raise RnbError(code, load().rnb_last_error().decode('utf-8', 'replace'))
    return code
