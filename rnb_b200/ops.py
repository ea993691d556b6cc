"""Torch-facing wrappers of the stand-alone C-ABI kernels (preprocess, gather, reduce, copy)."""
import ctypes

import torch

from . import _lib

CLIP_FRAMES = (8, 3, 112, 112)  # loader output per clip: (T, C, H, W), uint8


def _stream(device):
    return ctypes.c_void_p(torch.cuda.current_stream(device).cuda_stream)


def _vec3(v):
    if v is None:
        return None
    vals = [float(e) for e in v]
    if len(vals) != 3:
        raise ValueError('per-channel scale/bias need 3 entries')
    return (ctypes.c_float * 3)(*vals)


def preprocess_u8(frames, out=None, scale=None, bias=None):
    """uint8 (B, 8, 3, 112, 112) -> fp32 (B, 3, 8, 112, 112), value * scale[c] + bias[c].

    With the default identity scale/bias this equals the reference loader's
    ``frames.float().permute(0, 2, 1, 3, 4)`` made contiguous (models/r2p1d/model.py:149-150).
    """
    lib = _lib.load()
    if frames.dtype != torch.uint8 or tuple(frames.shape[1:]) != CLIP_FRAMES or not frames.is_cuda:
        raise ValueError('frames must be a CUDA uint8 tensor of shape (B, 8, 3, 112, 112)')
    frames = frames.contiguous()
    batch = frames.shape[0]
    if out is None:
        out = torch.empty((batch, 3, 8, 112, 112), dtype=torch.float32, device=frames.device)
    elif out.dtype != torch.float32 or tuple(out.shape) != (batch, 3, 8, 112, 112) \
            or not out.is_contiguous():
        raise ValueError('out must be a contiguous fp32 (B, 3, 8, 112, 112) tensor')
    with torch.cuda.device(frames.device):
        _lib.check(lib.rnb_preprocess_u8(ctypes.c_void_p(frames.data_ptr()), batch, _vec3(scale),
                                         _vec3(bias), ctypes.c_void_p(out.data_ptr()),
                                         _lib.FMT_F32_NCDHW, _stream(frames.device)))
    return out


def batch_gather(tensors, out):
    """Concatenate ``tensors`` along dim 0 into the leading rows of ``out`` (batcher.py:28)."""
    lib = _lib.load()
    k = len(tensors)
    if k == 0:
        return out[:0]
    row_shape = tuple(out.shape[1:])
    row_bytes = out[0].numel() * out.element_size() if out.shape[0] else 0
    total = 0
    for t in tensors:
        if tuple(t.shape[1:]) != row_shape or t.dtype != out.dtype or t.device != out.device:
            raise ValueError('all tensors must share row shape, dtype and device with `out`')
        if not t.is_contiguous():
            raise ValueError('batch_gather needs contiguous sources')
        total += t.shape[0]
    if total > out.shape[0]:
        raise ValueError('%d rows do not fit the %d-row batch buffer' % (total, out.shape[0]))
    srcs = (ctypes.c_void_p * k)(*[t.data_ptr() for t in tensors])
    rows = (ctypes.c_int * k)(*[int(t.shape[0]) for t in tensors])
    with torch.cuda.device(out.device):
        _lib.check(lib.rnb_batch_gather(srcs, rows, k, ctypes.c_void_p(out.data_ptr()), row_bytes,
                                        _stream(out.device)))
    return out[:total]


def logits_sum_argmax(logits, sum_out=None, cls_out=None):
    """Per-class sum over clips and its arg-max, on the GPU (models/r2p1d/model.py:254,259).

    Returns ``(sum, cls)`` device tensors (fp32 (classes,), int32 scalar); asynchronous.
    """
    lib = _lib.load()
    if logits.dtype != torch.float32 or logits.dim() != 2 or not logits.is_cuda:
        raise ValueError('logits must be a CUDA fp32 (rows, classes) tensor')
    logits = logits.contiguous()
    rows, classes = logits.shape
    if sum_out is None:
        sum_out = torch.empty(classes, dtype=torch.float32, device=logits.device)
    if cls_out is None:
        cls_out = torch.empty((), dtype=torch.int32, device=logits.device)
    with torch.cuda.device(logits.device):
        _lib.check(lib.rnb_logits_sum_argmax(ctypes.c_void_p(logits.data_ptr()), rows, classes,
                                             ctypes.c_void_p(sum_out.data_ptr()),
                                             ctypes.c_void_p(cls_out.data_ptr()),
                                             _stream(logits.device)))
    return sum_out, cls_out


def pool_fc_classify(x, weight, bias, scratch=None, want_logits=True):
    """Classifier tail in one launch: mean over positions, Linear, sum over clips, arg-max.

    ``x``: bf16 (B, S, C) channels-last activations of the last stage; ``weight`` fp32 (classes, C);
    ``bias`` fp32 (classes).  Returns ``(logits or None, total, cls)`` device tensors.  ``scratch`` is
    a uint8 CUDA tensor of at least ``pool_fc_classify_scratch_bytes(B, classes)`` bytes (allocated
    here when omitted).
    """
    lib = _lib.load()
    if x.dtype != torch.bfloat16 or x.dim() != 3 or not x.is_cuda:
        raise ValueError('x must be a CUDA bf16 (B, S, C) tensor')
    x = x.contiguous()
    weight = weight.contiguous()
    bias = bias.contiguous()
    batch, positions, channels = x.shape
    classes = weight.shape[0]
    need = int(lib.rnb_pool_fc_classify_scratch_bytes(batch, classes))
    if scratch is None:
        scratch = torch.empty(need, dtype=torch.uint8, device=x.device)
    logits = torch.empty((batch, classes), dtype=torch.float32, device=x.device) \
        if want_logits else None
    total = torch.empty(classes, dtype=torch.float32, device=x.device)
    cls = torch.empty((), dtype=torch.int32, device=x.device)
    with torch.cuda.device(x.device):
        _lib.check(lib.rnb_pool_fc_classify(
            ctypes.c_void_p(x.data_ptr()), batch, positions, channels,
            ctypes.c_void_p(weight.data_ptr()), ctypes.c_void_p(bias.data_ptr()), classes,
            ctypes.c_void_p(logits.data_ptr()) if logits is not None else None,
            ctypes.c_void_p(total.data_ptr()), ctypes.c_void_p(cls.data_ptr()),
            ctypes.c_void_p(scratch.data_ptr()), scratch.numel(), _stream(x.device)))
    return logits, total, cls


def peer_copy(dst, src, src_stream=None, on=None):
    """``dst.copy_(src)`` for same-shape contiguous CUDA tensors, possibly on different GPUs.

    Runs on stream ``on`` (default: the current stream of the *destination* device -- the consumer
    pulls, as in runner.py:112-113) and is ordered, on the GPU, after everything queued so far on
    ``src_stream`` (default: the current stream of the source device, i.e. after the kernels that
    produced ``src``; it must be a stream of the source device).  That is the ordering
    ``Tensor.copy_`` gives the reference, without its host wait.
    """
    lib = _lib.load()
    if dst.shape != src.shape or dst.dtype != src.dtype:
        raise ValueError('peer_copy needs tensors of identical shape and dtype')
    if not (dst.is_contiguous() and src.is_contiguous() and dst.is_cuda and src.is_cuda):
        raise ValueError('peer_copy needs contiguous CUDA tensors')
    nbytes = src.numel() * src.element_size()
    if src_stream is None:
        src_stream = torch.cuda.current_stream(src.device)
    if on is None:
        on = torch.cuda.current_stream(dst.device)
    with torch.cuda.device(on.device):
        _lib.check(lib.rnb_peer_copy_fenced(ctypes.c_void_p(dst.data_ptr()), dst.device.index,
                                            ctypes.c_void_p(src.data_ptr()), src.device.index,
                                            nbytes, ctypes.c_void_p(src_stream.cuda_stream),
                                            ctypes.c_void_p(on.cuda_stream)))
    return dst
