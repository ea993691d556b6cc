"""The no-allocation classifier entry point, rnb_pool_fc_classify_static.

It exists for callers that cannot call cudaMalloc (a server whose device memory was carved up at
start-up and locked, a latency-critical loop that must stay clear of the allocator, a host that only
lends raw pointers): the scratch the classifier needs is a
fixed array inside the library instead of a buffer the caller provides.  This test drives it through
the C ABI exactly as such a caller would -- raw device pointers, no scratch argument -- and checks
every output against rnb_pool_fc_classify (caller-provided scratch) on the same inputs, twice in a
row, so that whatever the first call leaves in the shared array is shown not to leak into the second.

Collected last on purpose (file name): a run that stops at the first failure has then already
executed every other test.  It must pass on its own in a fresh process as well as after the rest.
"""
import ctypes

import pytest
import torch

pytestmark = pytest.mark.gpu

DEV = 'cuda:0'
TOL = dict(rtol=1e-5, atol=1e-5)  # the tolerance test_pool_fc_classify_standalone uses


def _ptr(t):
    return ctypes.c_void_p(t.data_ptr())


def _run_both(lib, x, w, b):
    """(logits, sum, cls) from the caller-scratch entry point and from the static one."""
    from rnb_b200 import _lib
    batch, positions, channels = x.shape
    classes = w.shape[0]
    stream = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    outs = []
    for static in (False, True):
        logits = torch.full((batch, classes), float('nan'), device=DEV)
        total = torch.full((classes,), float('nan'), device=DEV)
        cls = torch.full((), -1, dtype=torch.int32, device=DEV)
        if static:
            rc = lib.rnb_pool_fc_classify_static(_ptr(x), batch, positions, channels, _ptr(w),
                                                 _ptr(b), classes, _ptr(logits), _ptr(total),
                                                 _ptr(cls), stream)
        else:
            need = int(lib.rnb_pool_fc_classify_scratch_bytes(batch, classes))
            scratch = torch.empty(need, dtype=torch.uint8, device=DEV)
            rc = lib.rnb_pool_fc_classify(_ptr(x), batch, positions, channels, _ptr(w), _ptr(b),
                                          classes, _ptr(logits), _ptr(total), _ptr(cls),
                                          _ptr(scratch), need, stream)
        _lib.check(rc)
        torch.cuda.synchronize()
        outs.append((logits.cpu(), total.cpu(), int(cls.item())))
    return outs


def test_static_scratch_matches_caller_scratch_twice_in_a_row(lib):
    g = torch.Generator().manual_seed(2024)
    w = (torch.randn(400, 512, generator=g) * 0.05).to(DEV)
    b = (torch.randn(400, generator=g) * 0.1).to(DEV)
    # first a full 15-clip video, then a single clip: the second call sees the first one's leftovers
    for batch in (15, 1):
        x = torch.randn(batch, 49, 512, generator=g).relu().to(torch.bfloat16).to(DEV)
        (ref_logits, ref_sum, ref_cls), (logits, total, cls) = _run_both(lib, x, w, b)
        torch.testing.assert_close(logits, ref_logits, **TOL)
        torch.testing.assert_close(total, ref_sum, **TOL)
        assert cls == ref_cls
        assert cls == int(ref_sum.numpy().argmax())


def test_static_scratch_rejects_sizes_beyond_its_limit(lib):
    x = torch.zeros(1, 49, 512, dtype=torch.bfloat16, device=DEV)
    w = torch.zeros(400, 512, device=DEV)
    b = torch.zeros(400, device=DEV)
    out = torch.zeros(65 * 400, device=DEV)
    cls = torch.zeros((), dtype=torch.int32, device=DEV)
    rc = lib.rnb_pool_fc_classify_static(_ptr(x), 65, 49, 512, _ptr(w), _ptr(b), 400, _ptr(out),
                                         _ptr(out), _ptr(cls), None)
    assert rc == -1
    msg = lib.rnb_last_error().decode()
    assert '64' in msg and '1024' in msg and 'rnb_pool_fc_classify' in msg
    rc = lib.rnb_pool_fc_classify_static(_ptr(x), 1, 49, 512, _ptr(w), _ptr(b), 1025, _ptr(out),
                                         _ptr(out), _ptr(cls), None)
    assert rc == -1
