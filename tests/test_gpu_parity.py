"""GPU parity: the CUDA path, called through the C ABI, against the CPU oracle on the same inputs.

Tolerances are BASELINE.json's: per-clip max|delta| / max|ref| <= 1e-3 in fp32 mode and <= 2e-2 in
bf16 mode, arg-max class identical; byte/index work (preprocess, gather, arg-max) is bit-exact.
"""
import numpy as np
import pytest
import torch

from oracle import r2p1d_ref as R

pytestmark = pytest.mark.gpu

TOL = {'fp32': 1e-3, 'bf16': 2e-2}
DEV = 'cuda:0'


def rel_err_per_clip(got, want):
    got, want = got.double().flatten(1), want.double().flatten(1)
    return ((got - want).abs().max(dim=1).values / want.abs().max(dim=1).values.clamp_min(1e-30))


def check_logits(got, want, dtype):
    err = rel_err_per_clip(got, want)
    assert float(err.max()) <= TOL[dtype], 'rel err %s' % err.tolist()
    assert (got.argmax(1) == want.argmax(1)).all(), 'arg-max class differs'


@pytest.fixture(scope='module')
def models():
    return {False: R.build(seed=0, harden=False), True: R.build(seed=0, harden=True)}


@pytest.fixture(scope='module')
def engines(lib, models):
    from rnb_b200.engine import R2P1DEngine
    made = {}

    def get(dtype, harden):
        key = (dtype, harden)
        if key not in made:
            made[key] = R2P1DEngine(DEV, models[harden].state_dict(), dtype=dtype, max_batch=16)
            made[key].prepare(16, 3)     # timed launch plans; other sizes derive theirs from these
        return made[key]
    return get


def test_extension_is_loaded(lib):
    """The GPU suite must run on the native library, never on a fallback."""
    from rnb_b200 import _lib
    assert lib is _lib.load()
    assert _lib.check(lib.rnb_init(0)) == 0
    maps = open('/proc/self/maps').read()
    assert 'librnb_b200.so' in maps


def test_preprocess_bit_exact(lib):
    from rnb_b200 import ops
    u8 = R.synthetic_clips(3, seed=9)
    got = ops.preprocess_u8(u8.to(DEV)).cpu()
    want = R.preprocess(u8).contiguous()
    assert got.dtype == torch.float32 and got.shape == want.shape
    assert torch.equal(got, want)
    # per-channel affine (identity is the reference behaviour)
    got = ops.preprocess_u8(u8.to(DEV), scale=[0.5, 2.0, 1.0], bias=[1.0, -3.0, 0.25]).cpu()
    s = torch.tensor([0.5, 2.0, 1.0]).view(1, 3, 1, 1, 1)
    b = torch.tensor([1.0, -3.0, 0.25]).view(1, 3, 1, 1, 1)
    assert torch.equal(got, want * s + b)
    # empty batch
    assert ops.preprocess_u8(u8[:0].to(DEV)).shape[0] == 0


@pytest.mark.parametrize('dtype', ['bf16', 'fp32'])
@pytest.mark.parametrize('harden', [False, True])
def test_whole_network_matches_oracle(engines, models, dtype, harden):
    clips = R.synthetic_clips(3, seed=21)
    want = R.forward_u8(models[harden], clips)
    got = engines(dtype, harden)(clips.to(DEV)).cpu()
    check_logits(got, want, dtype)


@pytest.mark.parametrize('dtype', ['bf16', 'fp32'])
def test_goldens(engines, golden, dtype):
    clips = R.synthetic_clips(2, seed=1234).to(DEV)
    for tag, harden in (('default', False), ('hardened', True)):
        got = engines(dtype, harden)(clips).cpu()
        check_logits(got, torch.from_numpy(golden['r18_%s_logits' % tag]), dtype)
    g = torch.Generator().manual_seed(7)
    x = torch.randn(1, 3, 8, 112, 112, generator=g)
    got = engines(dtype, False)(x.to(DEV)).cpu()
    check_logits(got, torch.from_numpy(golden['r18_default_randn_logits']), dtype)


def test_fp32_tensor_input_equals_uint8_input(engines):
    """The runner step receives fp32 (B,3,8,112,112) (control.py:41); same result as raw frames."""
    clips = R.synthetic_clips(2, seed=33)
    eng = engines('bf16', True)
    a = eng(clips.to(DEV)).clone()
    b = eng(R.preprocess(clips).contiguous().to(DEV)).clone()
    assert torch.equal(a, b)
    # and a permuted (non-contiguous) view, which is what the reference loader returns
    c = eng(clips.to(DEV).float().permute(0, 2, 1, 3, 4)).clone()
    assert torch.equal(a, c)


def test_config2_batch16_bf16(engines, models):
    """BASELINE config 2: R(2+1)D-18, bf16, batch 16, at full size."""
    clips = R.synthetic_clips(16, seed=1234)
    torch.set_num_threads(8)
    want = R.forward_u8(models[True], clips)
    eng = engines('bf16', True)
    got = eng(clips.to(DEV)).cpu().clone()
    check_logits(got, want, 'bf16')
    # size-independent property: clips are independent, so any sub-batch gives the same rows
    perm = torch.randperm(16, generator=torch.Generator().manual_seed(1))
    got_perm = eng(clips[perm].to(DEV)).cpu()
    torch.testing.assert_close(got_perm, got[perm], rtol=0, atol=0)
    # Another batch size has its own launch plan, and where that plan splits K across a cluster
    # (conv_splitk.cu) the fp32 summation order differs: same rows within rounding, same classes.
    # (test_batch_invariance_without_splitk checks the bit-exact property.)
    for n in (1, 5, 15):
        sub = eng(clips[:n].to(DEV)).cpu()
        assert float(rel_err_per_clip(sub, got[:n]).max()) <= TOL['bf16'] / 4
        assert (sub.argmax(1) == got[:n].argmax(1)).all()
        check_logits(sub, want[:n], 'bf16')
    assert R.video_class(got) == R.video_class(want)


def test_batch_invariance_without_splitk(lib, tmp_path):
    """With the split-K kernel out of the candidate set (RNB_B200_SPLITK=0) every kernel sums K in
    one fixed order, so a clip's logits do not depend on the batch it travels in: bit-identical
    rows at batch 16, 5 and 1."""
    import os
    import subprocess
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    code = (
        "import sys, torch; sys.path.insert(0, %r)\n"
        "from rnb_b200.engine import R2P1DEngine, random_state_dict\n"
        "eng = R2P1DEngine('cuda:0', random_state_dict(seed=5), max_batch=16).prepare(16, 5, 1)\n"
        "g = torch.Generator().manual_seed(11)\n"
        "x = torch.randint(0, 256, (16, 8, 3, 112, 112), dtype=torch.uint8, generator=g).cuda()\n"
        "y16 = eng(x).cpu().clone(); y5 = eng(x[:5]).cpu().clone(); y1 = eng(x[3:4]).cpu().clone()\n"
        "torch.cuda.synchronize()\n"
        "torch.save((y16, y5, y1), sys.argv[1])\n" % root)
    path = str(tmp_path / 'rows.pt')
    env = dict(os.environ, RNB_B200_SPLITK='0')
    subprocess.run([sys.executable, '-c', code, path], check=True, env=env, timeout=600)
    y16, y5, y1 = torch.load(path)
    assert torch.equal(y5, y16[:5])
    assert torch.equal(y1, y16[3:4])


def test_stem_fused_and_unfolded_agree(lib, monkeypatch):
    """The stem as one fused kernel on a space-to-depth input (default) and as two patch-kernel
    launches on the unfolded layout (RNB_B200_STEM=unfolded): both against the oracle's conv1 output
    (network.py:22), from uint8 frames and from the fp32 tensor the runner step receives."""
    from rnb_b200.engine import R2P1DEngine
    m = R.build(1, 1, seed=3, harden=True)
    clips = R.synthetic_clips(3, seed=17)
    x = R.preprocess(clips).contiguous()
    want = R.forward(m, x)
    outs = {}
    for mode in ('fused', 'unfolded'):
        monkeypatch.setenv('RNB_B200_STEM', mode)
        eng = R2P1DEngine(DEV, m.state_dict(), 1, 1, dtype='bf16', max_batch=3).prepare(3)
        kinds = [k for *_, k in eng.op_info()]
        assert kinds == (['conv_stem_fused'] * 2 if mode == 'fused' else ['conv_patch'] * 2), kinds
        a = eng(clips.to(DEV)).cpu().clone()
        b = eng(x.to(DEV)).cpu().clone()
        assert torch.equal(a, b)
        assert float(rel_err_per_clip(a, want).max()) <= TOL['bf16']
        outs[mode] = a
        eng.close()
    assert float(rel_err_per_clip(outs['fused'], outs['unfolded']).max()) <= TOL['bf16'] / 2


@pytest.mark.parametrize('dtype', ['bf16', 'fp32'])
@pytest.mark.parametrize('start,end', [(1, 1), (2, 2), (3, 3), (4, 4), (3, 4), (5, 5), (2, 5)])
def test_layer_ranges(lib, dtype, start, end):
    """R2P1DRunner(start_index, end_index): fp32 NCDHW in and out (models/r2p1d/model.py:29-33)."""
    from rnb_b200.engine import R2P1DEngine
    m = R.build(start, end, seed=3, harden=True)
    g = torch.Generator().manual_seed(100 + start)
    if start == 1:
        x = R.preprocess(R.synthetic_clips(2, seed=5)).contiguous()
    else:
        x = torch.randn(2, *R.LAYER_INPUT_SHAPE[start], generator=g).abs()
    want = R.forward(m, x)
    eng = R2P1DEngine(DEV, m.state_dict(), start, end, dtype=dtype, max_batch=2).prepare(2)
    got = eng(x.to(DEV)).cpu()
    assert got.shape == want.shape
    if end == 5:
        check_logits(got, want, dtype)
    else:
        assert float(rel_err_per_clip(got, want).max()) <= TOL[dtype]
    eng.close()


def test_r34_runner_form_layer_sizes(lib, golden):
    """R(2+1)D-34: layer_sizes indexed by position in the range, stem slot ignored."""
    from rnb_b200.engine import R2P1DEngine
    m = R.build(layer_sizes=[0, 3, 4, 6, 3], seed=0, harden=True)
    eng = R2P1DEngine(DEV, m.state_dict(), layer_sizes=[0, 3, 4, 6, 3], dtype='bf16', max_batch=2)
    got = eng(R.synthetic_clips(2, seed=1234)[:1].to(DEV)).cpu()
    check_logits(got, torch.from_numpy(golden['r34_hardened_logits']), 'bf16')
    assert abs(eng.flops(1) / 1e9 - 76.853) < 0.01  # SURVEY.md section 8 d
    eng.close()


def test_flops_and_launch_count(engines):
    eng = engines('bf16', False)
    assert abs(eng.flops(1) / 1e9 - 42.172) < 0.01
    # 40 Conv3d modules; the six convolutions of the three downsample shortcuts run fused into the
    # first (N-concatenation) and last (K-concatenation) convolution of their block: 34 launches
    assert eng.num_convs == 34
    eng(R.synthetic_clips(2, seed=1).to(DEV))
    # 34 convolutions + input staging + pool/classifier; a SpatioTemporalConv whose two convolutions
    # the plan runs as one co-scheduled launch counts once
    one_launch = ('conv_patch_dual', 'conv_stem_fused')
    duals = sum(k in one_launch for *_, k in eng.op_info()) // 2
    assert eng.num_launches == 36 - duals
    ms = eng.profile(R.synthetic_clips(16, seed=1).to(DEV))
    duals = sum(k in one_launch for *_, k in eng.op_info()) // 2
    assert len(ms) == 36 and sum(t > 0 for t in ms) == 36 - duals
    assert [k for *_, k in eng.op_info()][:2] == ['conv_stem_fused'] * 2   # the default stem
    assert sum('+shortcut' in name for name, *_ in eng.op_info()) == 6


def test_batch_gather_bit_exact(lib):
    from rnb_b200 import ops
    g = torch.Generator().manual_seed(0)
    parts = [torch.randn(n, 3, 8, 112, 112, generator=g).to(DEV) for n in (1, 0, 2, 1)]
    out = torch.empty(15, 3, 8, 112, 112, device=DEV)
    got = ops.batch_gather(parts, out)
    assert torch.equal(got, torch.cat(parts, dim=0))            # batcher.py:28
    odd = [torch.arange(n * 7, dtype=torch.uint8, device=DEV).view(n, 7) for n in (3, 2)]
    got = ops.batch_gather(odd, torch.empty(8, 7, dtype=torch.uint8, device=DEV))
    assert torch.equal(got, torch.cat(odd, dim=0))
    with pytest.raises(ValueError):
        ops.batch_gather(parts, out[:2])


def test_logits_sum_argmax(lib):
    from rnb_b200 import ops
    g = torch.Generator().manual_seed(4)
    for rows in (1, 5, 15):
        logits = torch.randn(rows, 400, generator=g)
        s, c = ops.logits_sum_argmax(logits.to(DEV))
        want = logits.numpy().sum(axis=0)                          # models/r2p1d/model.py:254
        np.testing.assert_allclose(s.cpu().numpy(), want, rtol=1e-6, atol=1e-6)
        assert int(c.item()) == int(s.cpu().numpy().argmax())
    tie = torch.zeros(2, 400)
    tie[:, 7] = 1.0
    tie[:, 300] = 1.0
    assert int(ops.logits_sum_argmax(tie.to(DEV))[1].item()) == 7  # numpy's lowest-index rule
    # numpy's NaN rule: a NaN sum is the maximum, the first one wins (model.py:259 on a NaN logit)
    nan = torch.randn(3, 400, generator=g)
    nan[1, 250] = float('nan')
    nan[2, 31] = float('nan')
    assert int(ops.logits_sum_argmax(nan.to(DEV))[1].item()) == int(nan.numpy().sum(axis=0).argmax()) == 31
    allnan = torch.full((2, 400), float('nan'))
    assert int(ops.logits_sum_argmax(allnan.to(DEV))[1].item()) == 0


def test_pool_fc_classify_standalone(lib):
    """pool -> view -> linear (network.py:35,41,56,60) -> sum over clips -> argmax (model.py:254,259)."""
    from rnb_b200 import ops
    g = torch.Generator().manual_seed(12)
    lin = torch.nn.Linear(512, 400)
    for batch in (1, 5, 15, 64):
        x = torch.randn(batch, 49, 512, generator=g).relu().to(torch.bfloat16)
        want = torch.nn.functional.linear(x.float().mean(dim=1), lin.weight, lin.bias).detach()
        logits, total, cls = ops.pool_fc_classify(x.to(DEV), lin.weight.detach().to(DEV),
                                                  lin.bias.detach().to(DEV))
        assert float(rel_err_per_clip(logits.cpu(), want).max()) <= 1e-5
        want_sum = want.numpy().sum(axis=0)
        np.testing.assert_allclose(total.cpu().numpy(), want_sum, rtol=1e-5, atol=1e-5)
        assert int(cls.item()) == int(total.cpu().numpy().argmax())
        assert int(cls.item()) == int(want_sum.argmax())
    # the caller's scratch may hold anything on entry, and is reusable back to back
    need = int(lib.rnb_pool_fc_classify_scratch_bytes(15, 400))
    scratch = torch.full((need,), 0xff, dtype=torch.uint8, device=DEV)
    x = torch.randn(15, 49, 512, generator=g).to(torch.bfloat16).to(DEV)
    a = ops.pool_fc_classify(x, lin.weight.detach().to(DEV), lin.bias.detach().to(DEV), scratch)
    b = ops.pool_fc_classify(x, lin.weight.detach().to(DEV), lin.bias.detach().to(DEV), scratch)
    assert torch.equal(a[1], b[1]) and int(a[2]) == int(b[2])
    with pytest.raises(Exception):
        ops.pool_fc_classify(x, lin.weight.detach().to(DEV), lin.bias.detach().to(DEV), scratch[:512])


@pytest.mark.parametrize('dtype', ['bf16', 'fp32'])
def test_engine_classify_matches_oracle_aggregator(engines, models, dtype):
    clips = R.synthetic_clips(15, seed=55)
    want = R.forward_u8(models[True], clips)
    eng = engines(dtype, True)
    logits, total, cls = eng.classify(clips.to(DEV))
    check_logits(logits.cpu(), want, dtype)
    np.testing.assert_allclose(total.cpu().numpy(), logits.cpu().numpy().sum(axis=0), rtol=1e-5,
                               atol=1e-4)
    assert int(cls.item()) == R.video_class(want)
    # and identical logits to the plain forward
    assert torch.equal(logits.clone(), eng(clips.to(DEV)))


def test_video_classifier_step(lib, models):
    from rnb_b200 import steps
    from rnb_b200.timecard import TimeCard
    dev = torch.device(DEV)
    step = steps.R2P1DVideoClassifier(dev, state_dict=models[True].state_dict(), seed=3)
    assert steps.R2P1DVideoClassifier.output_shape() is None
    for vid in range(3):
        tensors, cls_host, tc = step(None, 'v%d.mp4' % vid, TimeCard(vid))
        torch.cuda.current_stream().synchronize()
        n = tc.num_clips
        want = R.forward_u8(models[True], step.clips.pool[:n].cpu())
        assert tensors is None and int(cls_host) == R.video_class(want)


def test_peer_copy_same_device(lib):
    from rnb_b200 import handoff
    src = torch.randn(5, 400, device=DEV)
    shared = torch.zeros(15, 400, device=DEV)
    n = handoff.publish(shared, src)
    place = torch.zeros(15, 400, device=DEV)
    got = handoff.fetch(place, shared, n)
    assert n == 5 and torch.equal(got, src)


def test_errors_are_loud(lib, models):
    from rnb_b200 import _lib
    from rnb_b200.engine import R2P1DEngine
    sd = dict(models[False].state_dict())
    del sd['res2plus1d.conv3.block1.bn1.running_var']
    with pytest.raises(_lib.RnbError) as e:
        R2P1DEngine(DEV, sd)
    assert e.value.code == -3 and 'conv3.block1.bn1.running_var' in str(e.value)
    eng = R2P1DEngine(DEV, models[False].state_dict(), max_batch=2)
    with pytest.raises(ValueError):
        eng(torch.zeros(3, 8, 3, 112, 112, dtype=torch.uint8, device=DEV))
    with pytest.raises(ValueError):
        eng(torch.zeros(1, 3, 8, 56, 56, device=DEV))
    with pytest.raises(TypeError):
        eng(torch.zeros(1, 3, 8, 112, 112, dtype=torch.float16, device=DEV))
    assert eng(torch.zeros(0, 8, 3, 112, 112, dtype=torch.uint8, device=DEV)).shape == (0, 400)
    eng.close()


def test_step_classes_end_to_end(lib, models):
    """Loader -> (hand-off) -> Runner -> Aggregator driven the way runner.py drives steps."""
    from rnb_b200 import handoff, steps
    from rnb_b200.timecard import TimeCard
    dev = torch.device(DEV)
    sd = models[True].state_dict()
    loader = steps.R2P1DLoader(dev, seed=0)
    runner = steps.R2P1DRunner(dev, state_dict=sd, dtype='bf16')
    agg = steps.R2P1DAggregator(dev, aggregate=1)
    shared = torch.empty(steps.R2P1DLoader.output_shape()[0], device=dev)
    placeholder = torch.zeros_like(shared)
    for vid in range(4):
        tc = TimeCard(vid)
        (frames,), _, tc = loader(None, 'video-%d.mp4' % vid, tc)
        n = handoff.publish(shared, frames)                 # runner.py:167-171
        x = handoff.fetch(placeholder, shared, n)           # runner.py:107-114
        (logits,), _, tc = runner((x,), None, tc)
        torch.cuda.current_stream().synchronize()           # runner.py:127-128
        _, cls, tc = agg((logits,), None, tc)
        want = R.forward_u8(models[True], loader.clips.pool[:n].cpu())
        check_logits(logits.cpu(), want, 'bf16')
        assert int(cls) == R.video_class(want)
        assert tc.num_clips == n


def test_segmented_video_matches_whole(lib, models):
    """num_segments=3 scatter + aggregate=3 merge == one 15-clip call (runner.py:140-173)."""
    from rnb_b200 import handoff, steps
    from rnb_b200.timecard import TimeCard
    dev = torch.device(DEV)
    runner = steps.R2P1DRunner(dev, state_dict=models[True].state_dict(), dtype='bf16')
    agg = steps.R2P1DAggregator(dev, aggregate=3)
    clips = R.synthetic_clips(15, seed=77)
    x = R.preprocess(clips).contiguous().to(dev)
    whole = runner.model(x).cpu().clone()
    parent = TimeCard(5)
    parent.record('enqueue_filename')
    out = None
    for s, (lo, hi) in enumerate(handoff.segment_ranges(15, 3)):
        tc = parent.fork(s)
        (logits,), _, tc = runner((x[lo:hi],), None, tc)
        tc.record('inference1_finish')
        out = agg((logits.clone(),), None, tc)
    _, cls, merged = out
    assert int(cls) == int(whole.numpy().sum(axis=0).argmax())
    assert merged.sub_id is None and 'inference1_finish-2' in merged.timings


def test_batcher_copies_instead_of_aliasing(lib):
    from rnb_b200 import steps
    from rnb_b200.timecard import TimeCard, TimeCardList
    dev = torch.device(DEV)
    b = steps.Batcher(dev, batch=3)
    placeholder = torch.zeros(15, 3, 8, 112, 112, device=dev)
    seen = []
    out = None
    for i in range(3):
        placeholder[:1] = float(i + 1)          # the runner overwrites its placeholder per request
        seen.append(placeholder[:1].clone())
        out = b((placeholder[:1],), None, TimeCard(i))
        if i < 2:
            assert out == (None, None, None)
    (batch,), _, tcl = out
    assert isinstance(tcl, TimeCardList) and len(tcl.time_cards) == 3
    assert torch.equal(batch, torch.cat(seen, dim=0))
    # pass-through when batch <= 1 (batcher.py:18-19)
    t = (placeholder[:2],)
    assert steps.Batcher(dev, batch=1)(t, 'x', 'tc') == (t, 'x', 'tc')


def test_peer_copy_across_gpus(lib):
    """Segment hand-off between two GPUs (runner.py:112-113,170) over NVLink peer copy."""
    if torch.cuda.device_count() < 2:
        pytest.skip('needs two GPUs')
    from rnb_b200 import handoff
    src = torch.randn(15, 400, device='cuda:0')
    shared = torch.zeros(15, 400, device='cuda:1')
    n = handoff.publish(shared, src[:5].contiguous())          # producer on GPU 0 pushes
    placeholder = torch.zeros(15, 400, device='cuda:0')
    got = handoff.fetch(placeholder, shared, n)                # consumer pulls it back
    assert n == 5 and torch.equal(got, src[:5])
    assert torch.equal(shared[:5].cpu(), src[:5].cpu())


def test_host_pipeline_end_to_end(engines, models):
    """Pinned host frames in, host logits out, copies overlapped with compute: same numbers."""
    from rnb_b200.engine import HostPipeline
    eng = engines('bf16', True)
    pipe = HostPipeline(eng, batch=4)
    batches = [R.synthetic_clips(4, seed=200 + i).pin_memory() for i in range(5)]
    slots, outs = [], []
    for i, hb in enumerate(batches):
        slot = pipe.submit(hb)
        if i >= 1:                       # keep one batch in flight, like a serving loop
            outs.append(pipe.result(slots[-1]).clone())
        slots.append(slot)
    outs.append(pipe.result(slots[-1]).clone())
    for hb, got in zip(batches, outs):
        want = R.forward_u8(models[True], hb)
        check_logits(got, want, 'bf16')


def test_chunked_schedule_gives_identical_logits(lib, tmp_path):
    """RNB_B200_CHUNK_MB > 0 runs each spatial->temporal pair in batch chunks through one reused
    intermediate buffer; a scheduling change only, so the logits must be bit-identical."""
    import os
    import subprocess
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    code = (
        "import sys, torch; sys.path.insert(0, %r)\n"
        "from rnb_b200.engine import R2P1DEngine, random_state_dict\n"
        "sd = random_state_dict(seed=3)\n"
        "eng = R2P1DEngine('cuda:0', sd, max_batch=16).prepare(16)\n"
        "g = torch.Generator().manual_seed(9)\n"
        "x = torch.randint(0, 256, (16, 8, 3, 112, 112), dtype=torch.uint8, generator=g).cuda()\n"
        "y = eng(x).cpu(); torch.cuda.synchronize()\n"
        "torch.save((y, eng.num_launches), sys.argv[1])\n" % root)
    outs = []
    for chunk in ('0', '20'):
        path = str(tmp_path / ('logits_%s.pt' % chunk))
        # (split-K off: a chunk's own launch plan must not change the K summation order)
        env = dict(os.environ, RNB_B200_CHUNK_MB=chunk, RNB_B200_SPLITK='0')
        subprocess.run([sys.executable, '-c', code, path], check=True, env=env, timeout=600)
        outs.append(torch.load(path))
    (y0, n0), (y1, n1) = outs
    assert n0 <= 36 and n1 > 36
    assert torch.equal(y0, y1)


def test_plan_cache_reproduces_the_plan(lib, tmp_path):
    """RNB_B200_PLAN_CACHE: the first process times the candidates and appends its per-layer choices
    to the file; a second process reuses them without timing and must produce bit-identical logits
    (same kernels, same summation order) and report the same kernel per layer."""
    import os
    import subprocess
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    code = (
        "import sys, torch; sys.path.insert(0, %r)\n"
        "from rnb_b200.engine import R2P1DEngine, random_state_dict\n"
        "eng = R2P1DEngine('cuda:0', random_state_dict(seed=7), max_batch=16).prepare(16)\n"
        "g = torch.Generator().manual_seed(13)\n"
        "x = torch.randint(0, 256, (16, 8, 3, 112, 112), dtype=torch.uint8, generator=g).cuda()\n"
        "y = eng(x).cpu().clone(); torch.cuda.synchronize()\n"
        "torch.save((y, [k for *_, k in eng.op_info()]), sys.argv[1])\n" % root)
    cache = str(tmp_path / 'plan.txt')
    env = dict(os.environ, RNB_B200_PLAN_CACHE=cache)
    outs = []
    for i in range(2):
        path = str(tmp_path / ('run%d.pt' % i))
        subprocess.run([sys.executable, '-c', code, path], check=True, env=env, timeout=600)
        outs.append(torch.load(path))
        if i == 0:
            text = open(cache).read()
            assert text.count('plan ') == 1
    assert open(cache).read() == text      # the second run found its plan and appended nothing
    (y0, k0), (y1, k1) = outs
    assert k0 == k1
    assert torch.equal(y0, y1)
