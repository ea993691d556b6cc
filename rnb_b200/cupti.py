"""Kernel timeline of this process: ``initialize()``, ``flush()``, ``report()``.

The same three calls the reference's ``utils/cupti`` extension offers (reference
``utils/cupti.cpp:142-166``, used by ``test_cupti.py:14-21``), on top of ``librnb_cupti.so``
(``include/rnb_cupti.h``).  ``report()`` returns ``[(kernel name, start ns, end ns), ...]`` and
empties the list, like the reference; ``report(details=True)`` adds device, stream, grid and block.
"""
import ctypes
import os

LIB_PATH = os.path.join(os.path.dirname(os.path.abspath(__file__)), 'csrc', 'librnb_cupti.so')
_lib = None

PROTOTYPES = {
    'rnb_cupti_initialize': (ctypes.c_int, []),
    'rnb_cupti_flush': (ctypes.c_int, []),
    'rnb_cupti_count': (ctypes.c_int, []),
    'rnb_cupti_get': (ctypes.c_int, [ctypes.c_int, ctypes.c_char_p, ctypes.c_int,
                                     ctypes.POINTER(ctypes.c_uint64), ctypes.POINTER(ctypes.c_uint64),
                                     ctypes.POINTER(ctypes.c_uint32), ctypes.POINTER(ctypes.c_uint32),
                                     ctypes.POINTER(ctypes.c_int32), ctypes.POINTER(ctypes.c_int32)]),
    'rnb_cupti_clear': (ctypes.c_int, []),
    'rnb_cupti_finalize': (ctypes.c_int, []),
    'rnb_cupti_last_error': (ctypes.c_char_p, []),
}


def load():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError('rnb_b200: %s is missing -- build it with `python __graft_entry__.py` '
                               'or `make -C rnb_b200/csrc librnb_cupti.so`' % LIB_PATH)
        lib = ctypes.CDLL(LIB_PATH)
        for name, (restype, argtypes) in PROTOTYPES.items():
            fn = getattr(lib, name)
            fn.restype, fn.argtypes = restype, argtypes
        _lib = lib
    return _lib


def _check(rc):
    if rc != 0:
        raise RuntimeError('rnb_cupti error %d: %s' % (rc, load().rnb_cupti_last_error().decode()))


def initialize():
    _check(load().rnb_cupti_initialize())


def flush():
    _check(load().rnb_cupti_flush())


def finalize():
    _check(load().rnb_cupti_finalize())


def report(details=False):
    lib = load()
    out = []
    name = ctypes.create_string_buffer(512)
    t0, t1 = ctypes.c_uint64(), ctypes.c_uint64()
    dev, stream = ctypes.c_uint32(), ctypes.c_uint32()
    grid, block = (ctypes.c_int32 * 3)(), (ctypes.c_int32 * 3)()
    for i in range(lib.rnb_cupti_count()):
        _check(lib.rnb_cupti_get(i, name, 512, ctypes.byref(t0), ctypes.byref(t1), ctypes.byref(dev),
                                 ctypes.byref(stream), grid, block))
        rec = (name.value.decode('utf-8', 'replace'), t0.value, t1.value)
        if details:
            rec += (dev.value, stream.value, tuple(grid), tuple(block))
        out.append(rec)
    _check(lib.rnb_cupti_clear())
    return out
