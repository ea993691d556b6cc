"""Every convolution kernel variant, pinned and checked against the CPU oracle.

`rnb_r2p1d_prepare` keeps whichever candidate times fastest, so which kernels an ordinary parity
test exercises depends on the box.  Here the plan is pinned instead (`rnb_r2p1d_force_kernel`): one
layer range of the network (reference `R2P1DRunner(start_index, end_index)`,
models/r2p1d/model.py:29-33, network.py:18-41) runs with the named kernel forced on every layer it
applies to, `op_info()` must report that kernel, and the activations must match the oracle's within
the bf16 tolerance of BASELINE.json (2e-2 of max|ref| per clip).

Also here: the batch sizes and models the BASELINE configs use beyond batch 16 (R18 at 1, 15, 32,
64; R(2+1)D-34 at 5 and 16) and the stand-alone kernel harness `csrc/test_conv`.
"""
import os
import subprocess

import pytest
import torch

from oracle import r2p1d_ref as R

pytestmark = pytest.mark.gpu

DEV = 'cuda:0'
TOL = 2e-2
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def rel_err_per_clip(got, want):
    got, want = got.double().flatten(1), want.double().flatten(1)
    return (got - want).abs().max(dim=1).values / want.abs().max(dim=1).values.clamp_min(1e-30)


_range_cache = {}


def _range_case(start, end, batch):
    """(oracle model, input, oracle output) of one layer range, hardened init, cached."""
    key = (start, end, batch)
    if key not in _range_cache:
        m = R.build(start, end, seed=3, harden=True)
        g = torch.Generator().manual_seed(100 + start)
        x = torch.randn(batch, *R.LAYER_INPUT_SHAPE[start], generator=g).abs()
        torch.set_num_threads(os.cpu_count() or 1)
        _range_cache[key] = (m, x, R.forward(m, x))
    return _range_cache[key]


# (test id, layer range, batch, which layers, kernel, parameters)
VARIANTS = [
    ('tc-1x1', (2, 2), 2, 'conv', 'conv_tc', dict(cm=1, cn=1)),
    ('tc-2x1', (3, 3), 4, 'conv', 'conv_tc', dict(cm=2, cn=1)),
    ('tc-1x2', (3, 3), 4, 'conv', 'conv_tc', dict(cm=1, cn=2)),
    ('tc-2x2', (3, 3), 4, 'conv', 'conv_tc', dict(cm=2, cn=2)),
    ('tc-4x1', (4, 4), 16, 'conv', 'conv_tc', dict(cm=4, cn=1)),
    ('tc-1x4', (4, 4), 16, 'conv', 'conv_tc', dict(cm=1, cn=4)),
    ('tc-narrow-tile', (4, 4), 16, 'temporal_conv', 'conv_tc', dict(cm=1, cn=1, block_n=96)),
    ('patch-resident', (2, 2), 2, 'conv', 'conv_patch', dict(pair=0)),
    ('patch-resident-pair', (2, 2), 2, 'conv', 'conv_patch', dict(pair=1)),
    ('patch-stream', (3, 3), 4, 'conv', 'conv_patch', dict(pair=0)),
    ('patch-stream-pair', (3, 3), 4, 'conv', 'conv_patch', dict(pair=1)),
    ('patch-stream-conv4', (4, 4), 16, 'conv', 'conv_patch', dict(pair=0)),
    ('pair-256', (3, 3), 4, 'conv', 'conv_pair', dict(block_n=256)),
    ('pair-128', (4, 4), 16, 'conv', 'conv_pair', dict(block_n=128)),
    ('pair-64', (5, 5), 16, 'conv', 'conv_pair', dict(block_n=64)),
    ('tpatch-256', (3, 3), 4, 'temporal_conv', 'conv_tpatch', dict(block_n=256)),
    ('tpatch-128', (4, 4), 16, 'temporal_conv', 'conv_tpatch', dict(block_n=128)),
    ('splitk-2', (5, 5), 16, 'conv', 'conv_splitk', dict(ks=2, block_n=256)),
    ('splitk-3', (5, 5), 16, 'conv', 'conv_splitk', dict(ks=3, block_n=128)),
    ('splitk-4', (4, 4), 16, 'conv', 'conv_splitk', dict(ks=4, block_n=128)),
    ('splitk-2-conv3', (3, 3), 4, 'spatial_conv', 'conv_splitk', dict(ks=2, block_n=256)),
    ('splitk-8', (5, 5), 2, 'conv', 'conv_splitk', dict(ks=8, block_n=128)),
    ('splitk-8-narrow', (5, 5), 1, 'spatial_conv', 'conv_splitk', dict(ks=8, block_n=64)),
    ('splitk-6-conv4', (4, 4), 1, 'conv', 'conv_splitk', dict(ks=6, block_n=128)),
    # one co-scheduled launch per SpatioTemporalConv (spatial CTAs + temporal CTAs, per-clip counters)
    ('dual-conv2', (2, 2), 3, 'spatial_conv', 'conv_patch_dual', dict(cm=100, cn=4)),
    ('dual-conv2-few-spatial', (2, 2), 2, 'spatial_conv', 'conv_patch_dual', dict(cm=20, cn=1)),
    ('dual-stem', (1, 1), 3, 'spatial_conv', 'conv_patch_dual', dict(cm=0, cn=0)),
    # 3x1x1 layers with the filter taps stacked along the output columns and summed in the epilogue
    ('tsum-conv2', (2, 2), 3, 'temporal_conv', 'conv_tsum', dict()),
    ('tsum-stem', (1, 1), 2, 'temporal_conv', 'conv_tsum', dict()),
]


@pytest.mark.parametrize('tag,rng,batch,which,kind,kw', VARIANTS, ids=[v[0] for v in VARIANTS])
def test_forced_kernel_matches_oracle(lib, monkeypatch, tag, rng, batch, which, kind, kw):
    from rnb_b200 import _lib
    from rnb_b200.engine import R2P1DEngine
    start, end = rng
    if start == 1:
        # pinning the stem's two convolutions needs the two-launch stem (the default is one fused kernel)
        monkeypatch.setenv('RNB_B200_STEM', 'unfolded')
    m, x, want = _range_case(start, end, batch)
    eng = R2P1DEngine(DEV, m.state_dict(), start, end, dtype='bf16', max_batch=batch)
    forced = []
    for i, (name, *_rest) in enumerate(eng.op_info()):
        if which not in name:
            continue
        try:
            eng.force_kernel(batch, i, kind, **kw)
            forced.append(i)
        except _lib.RnbError as e:
            assert e.code == -1, e          # "does not apply to this layer" is the only excuse
    assert forced, '%s applies to no layer of range %s' % (kind, rng)
    got = eng(x.to(DEV)).cpu()
    torch.cuda.synchronize()
    info = eng.op_info()
    for i in forced:
        assert info[i][4] == kind, (info[i][0], info[i][4])
    assert got.shape == want.shape
    if end == 5:
        err = rel_err_per_clip(got, want)
        assert (got.argmax(1) == want.argmax(1)).all()
    else:
        err = rel_err_per_clip(got, want)
    assert float(err.max()) <= TOL, '%s on %d layers: rel err %s' % (tag, len(forced), err.tolist())
    eng.close()


def test_unprepared_batch_size_does_not_wait_for_the_gpu(lib):
    """A batch size nobody prepared gets its plan on the host: the call returns while the stream is
    still busy with work queued before it (a tuning pass would have synchronised it)."""
    from rnb_b200.engine import R2P1DEngine, random_state_dict
    eng = R2P1DEngine(DEV, random_state_dict(seed=1), max_batch=16).prepare(16)
    x = torch.randint(0, 256, (16, 8, 3, 112, 112), dtype=torch.uint8, device=DEV)
    a = torch.randn(8192, 8192, device=DEV)
    torch.cuda.synchronize()
    stream = torch.cuda.current_stream()
    for n in (7, 3, 11):
        for _ in range(20):
            a @ a                       # ~20 x 1 ms of queued work
        eng(x[:n])
        assert not stream.query(), 'forward of an unprepared batch size waited for the stream'
        torch.cuda.synchronize()
    # derived plans copy the prepared size's kernel choices where they apply
    y16 = eng(x).cpu().clone()
    y7 = eng(x[:7]).cpu()
    assert float(rel_err_per_clip(y7, y16[:7]).max()) <= TOL / 4
    eng.close()


@pytest.fixture(scope='module')
def hardened_r18():
    return R.build(seed=0, harden=True)


@pytest.mark.parametrize('batch', [1, 15, 32, 64])
def test_r18_batch_sizes_of_the_baseline_configs(lib, hardened_r18, batch):
    """BASELINE configs 3 and 5: 1- and 15-clip videos, batch sweep up to 64 (the L2 persistence
    policy and the launch plan change above batch 16)."""
    from rnb_b200.engine import R2P1DEngine
    clips = R.synthetic_clips(batch, seed=900 + batch)
    torch.set_num_threads(os.cpu_count() or 1)
    want = R.forward_u8(hardened_r18, clips)
    eng = R2P1DEngine(DEV, hardened_r18.state_dict(), dtype='bf16', max_batch=batch).prepare(batch)
    got = eng(clips.to(DEV)).cpu()
    err = rel_err_per_clip(got, want)
    assert float(err.max()) <= TOL, err.tolist()
    assert (got.argmax(1) == want.argmax(1)).all()
    assert R.video_class(got) == R.video_class(want)
    eng.close()


@pytest.mark.parametrize('batch', [5, 16])
def test_r34_batches_of_configs_4_and_5(lib, batch):
    """R(2+1)D-34 [0,3,4,6,3]: 5-clip segments (config 4) and batch 16 (config 5)."""
    from rnb_b200.engine import R2P1DEngine
    m = R.build(layer_sizes=[0, 3, 4, 6, 3], seed=0, harden=True)
    clips = R.synthetic_clips(batch, seed=77)
    torch.set_num_threads(os.cpu_count() or 1)
    want = R.forward_u8(m, clips)
    eng = R2P1DEngine(DEV, m.state_dict(), layer_sizes=[0, 3, 4, 6, 3], dtype='bf16',
                      max_batch=batch).prepare(batch)
    got = eng(clips.to(DEV)).cpu()
    err = rel_err_per_clip(got, want)
    assert float(err.max()) <= TOL, err.tolist()
    assert (got.argmax(1) == want.argmax(1)).all()
    eng.close()


def test_single_step_call(lib, hardened_r18):
    """R2P1DSingleStep.__call__ (models/r2p1d/model.py:216-228): loader + whole network in one
    step, driven with the arguments the runner loop passes (runner.py:125-126)."""
    from rnb_b200 import steps
    from rnb_b200.timecard import TimeCard
    dev = torch.device(DEV)
    step = steps.R2P1DSingleStep(dev, state_dict=hardened_r18.state_dict(), seed=5)
    assert steps.R2P1DSingleStep.output_shape() == ((steps.MAX_CLIPS, 400),)
    assert step.input_shape() is None
    seen = set()
    for vid in range(12):
        tensors, non_tensors, tc = step(None, 'synthetic/video_%07d.mp4' % vid, TimeCard(vid))
        torch.cuda.current_stream().synchronize()          # runner.py:127-128
        n = tc.num_clips
        seen.add(n)
        (logits,) = tensors
        assert non_tensors is None and logits.shape == (n, 400)
        want = R.forward_u8(hardened_r18, step.clips.pool[:n].cpu())
        assert float(rel_err_per_clip(logits.cpu(), want).max()) <= TOL
        assert (logits.cpu().argmax(1) == want.argmax(1)).all()
    assert seen <= {1, 15}


def test_standalone_kernel_harness(lib):
    """csrc/test_conv: every kernel mode (cluster multicast shapes, CTA pairs, temporal patch,
    split-K, patch resident / streamed / paired, unfolded stem) against the direct fp32 convolution
    on the GPU; its exit code is the verdict."""
    exe = os.path.join(ROOT, 'rnb_b200', 'csrc', 'test_conv')
    if not os.path.exists(exe):
        import __graft_entry__
        __graft_entry__.build()
    proc = subprocess.run([exe], stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True,
                          timeout=600)
    tail = '\n'.join(proc.stdout.splitlines()[-15:])
    assert proc.returncode == 0 and 'ALL TESTS PASSED' in proc.stdout, tail


def test_cupti_kernel_timeline(lib):
    """initialize / flush / report of the CUPTI bridge (reference utils/cupti.cpp:142-166, driven
    like test_cupti.py:14-21): one record per kernel of a forward pass, with this library's kernel
    names and start < end."""
    from rnb_b200 import cupti
    from rnb_b200.engine import R2P1DEngine, random_state_dict
    eng = R2P1DEngine(DEV, random_state_dict(seed=2), max_batch=4)
    x = torch.randint(0, 256, (4, 8, 3, 112, 112), dtype=torch.uint8, device=DEV)
    eng(x)
    torch.cuda.synchronize()
    cupti.initialize()
    cupti.report()                      # drop anything collected so far
    eng(x)
    torch.cuda.synchronize()
    cupti.flush()
    recs = cupti.report(details=True)
    cupti.finalize()
    ours = [r for r in recs if 'rnb::' in r[0]]
    assert len(ours) == eng.num_launches, [r[0][:60] for r in recs]
    assert all(r[1] < r[2] for r in ours)
    assert any('conv_patch_kernel' in r[0] for r in ours) and any('pool_fc' in r[0] for r in ours)
    assert [r[1] for r in ours] == sorted(r[1] for r in ours)      # one stream: launch order
    assert cupti.report() == []                                    # report() empties the list
    eng.close()
