"""Hand-off of tensors between pipeline steps (reference runner.py:63-72,90-121,138-173 and
control.py:19-69): ordering of the copies against the producer's kernels, shared buffers that know
their dtype, segments and layer-range pipelines across GPUs, and the two-process case RnB really
runs (spawned workers, CUDA-IPC tensors, interprocess events).

Tests that need more than one GPU skip on a single-GPU box; the two-process test runs on one GPU
as well (producer and consumer are then two processes on the same device, like loader and runner
in config/r2p1d-whole.json).
"""
import os

import pytest
import torch

from oracle import r2p1d_ref as R

pytestmark = pytest.mark.gpu

TOL = 2e-2


def _ngpu():
    return torch.cuda.device_count()


def _busy(device, ms=30):
    """Queue roughly `ms` of matrix multiplies on the current stream of `device`."""
    a = torch.randn(4096, 4096, device=device)
    for _ in range(max(1, int(ms / 1.5))):
        a = (a @ a).clamp_(-1, 1)
    return a


def rel_err_per_clip(got, want):
    got, want = got.double().flatten(1), want.double().flatten(1)
    return (got - want).abs().max(dim=1).values / want.abs().max(dim=1).values.clamp_min(1e-30)


@pytest.mark.parametrize('with_events', [True, False])
def test_copy_is_ordered_behind_the_producers_kernels(lib, with_events):
    """The segment is written by a kernel queued right before publish(), behind ~30 ms of other
    work: a copy that did not wait for the producer's stream would ship the old contents."""
    from rnb_b200 import handoff
    src_dev = torch.device('cuda:0')
    dst_dev = torch.device('cuda:%d' % (1 if _ngpu() > 1 else 0))
    shape = ((15, 400),)
    shared = handoff.TensorEvent(shape, src_dev) if with_events \
        else torch.zeros(shape[0], device=src_dev)
    placeholder = torch.zeros(shape[0], device=dst_dev)
    for round_ in range(3):
        seg = torch.zeros(5, 400, device=src_dev)
        torch.cuda.synchronize()
        with torch.cuda.device(src_dev):
            _busy(src_dev)
            seg.add_(float(round_ + 1))                 # the producing kernel, last in line
            n = handoff.publish(shared, seg)
        with torch.cuda.device(dst_dev):
            got = handoff.fetch(placeholder, shared, n)
            # a kernel of the consumer right behind the fetch must see the rows too
            total = got.sum()
        assert n == 5
        assert float(total.item()) == 5 * 400 * (round_ + 1)
        if with_events:
            assert shared.valid_batch_sizes[0] == 5


def test_tensor_event_matches_the_reference_class(lib):
    """Attributes the reference's worker loop uses (control.py:40-46), plus the dtype."""
    from rnb_b200 import handoff
    te = handoff.TensorEvent(((4, 8, 3, 112, 112), (4, 400)), 'cuda:0',
                             dtype=(torch.uint8, torch.float32))
    assert [t.dtype for t in te.tensors] == [torch.uint8, torch.float32]
    assert [tuple(t.shape) for t in te.tensors] == [(4, 8, 3, 112, 112), (4, 400)]
    assert te.event.is_set() and len(te.valid_batch_sizes) == 2
    assert te.tensors[0].numel() * 4 == torch.empty(4, 3, 8, 112, 112).numel() * 4  # 1/4 the bytes
    cpu = handoff.TensorEvent(((2, 400),), 'cpu')
    assert cpu.ready is None and cpu.tensors[0].device.type == 'cpu'


def test_uint8_wire_gives_the_same_logits(lib):
    """Loader ships uint8 frames through a dtype-aware shared buffer; the runner feeds them to the
    network's first kernel (RNB_FMT_U8_NTCHW): identical logits to the fp32 wire of control.py:41."""
    from rnb_b200 import handoff, ops
    from rnb_b200.engine import R2P1DEngine
    m = R.build(seed=0, harden=True)
    dev = torch.device('cuda:0')
    eng = R2P1DEngine(dev, m.state_dict(), max_batch=15).prepare(5)
    frames = R.synthetic_clips(5, seed=3).to(dev)
    wire_u8 = handoff.TensorEvent(((5, 8, 3, 112, 112),), dev, dtype=torch.uint8)
    wire_f32 = handoff.TensorEvent(((5, 3, 8, 112, 112),), dev)
    handoff.publish(wire_u8, frames)
    handoff.publish(wire_f32, ops.preprocess_u8(frames))
    a = eng(handoff.fetch(torch.empty_like(wire_u8.tensors[0]), wire_u8)).clone()
    b = eng(handoff.fetch(torch.empty_like(wire_f32.tensors[0]), wire_f32)).clone()
    torch.cuda.synchronize()
    assert torch.equal(a, b)
    assert wire_u8.tensors[0].numel() * 4 == wire_f32.tensors[0].numel() * 4
    eng.close()


def test_layer_range_pipeline_across_two_gpus(lib):
    """Layers 1-2 on GPU 0, hand-off, layers 3-5 on GPU 1 == the whole network on one GPU
    (cut points of models/r2p1d/model.py:29-33).  Both wire formats: fp32 NCDHW (what RnB's shared
    tensors carry) and the library's bf16 channels-last form (half the bytes, no layout kernels)."""
    if _ngpu() < 2:
        pytest.skip('needs two GPUs')
    from rnb_b200 import handoff
    from rnb_b200.engine import R2P1DEngine
    m = R.build(seed=0, harden=True)
    sd = m.state_dict()
    d0, d1 = torch.device('cuda:0'), torch.device('cuda:1')
    front = R2P1DEngine(d0, sd, 1, 2, layer_sizes=[2, 2], max_batch=4).prepare(4)
    back = R2P1DEngine(d1, sd, 3, 5, layer_sizes=[2, 2, 2], max_batch=4).prepare(4)
    clips = R.synthetic_clips(4, seed=42)
    want = R.forward_u8(m, clips)
    for wire in (False, True):
        with torch.cuda.device(d0):
            act = front.forward(clips.to(d0), wire=wire)
            shared = handoff.TensorEvent((tuple(act.shape),), d0, dtype=act.dtype)
            n = handoff.publish(shared, act)
        with torch.cuda.device(d1):
            x = handoff.fetch(torch.empty(tuple(act.shape), dtype=act.dtype, device=d1), shared, n)
            got = back(x).cpu()
        assert float(rel_err_per_clip(got, want).max()) <= TOL
        assert (got.argmax(1) == want.argmax(1)).all()
    front.close()
    back.close()


def test_three_segments_across_gpus(lib):
    """num_segments = 3 (config/r2p1d-segment.json:10): a 15-clip video is cut 5/5/5
    (runner.py:149-151), each segment crosses to another GPU's runner, the aggregator's sum over all
    clips picks the same class as one 15-clip call."""
    if _ngpu() < 2:
        pytest.skip('needs two GPUs')
    from rnb_b200 import handoff, steps
    from rnb_b200.engine import R2P1DEngine
    from rnb_b200.timecard import TimeCard
    m = R.build(seed=0, harden=True)
    sd = m.state_dict()
    d0 = torch.device('cuda:0')
    consumers = [torch.device('cuda:%d' % (1 + s % (_ngpu() - 1))) for s in range(3)]
    engines = {}
    for d in consumers:
        if d not in engines:
            engines[d] = R2P1DEngine(d, sd, max_batch=15).prepare(5)
    clips = R.synthetic_clips(15, seed=8)
    want = R.forward_u8(m, clips)
    frames = clips.to(d0)
    shapes = handoff.get_segmented_shapes(((15, 8, 3, 112, 112),), 3)
    ring = [handoff.TensorEvent(shapes, d0, dtype=torch.uint8) for _ in range(3)]
    agg = steps.R2P1DAggregator(torch.device('cpu'), aggregate=3)
    parent = TimeCard(1)
    parent.record('enqueue_filename')
    out = None
    for s, (lo, hi) in enumerate(handoff.segment_ranges(15, 3)):
        with torch.cuda.device(d0):
            n = handoff.publish(ring[s], frames[lo:hi])
        d = consumers[s]
        with torch.cuda.device(d):
            x = handoff.fetch(torch.empty(shapes[0], dtype=torch.uint8, device=d), ring[s], n)
            logits = engines[d](x)
            torch.cuda.current_stream().synchronize()
        tc = parent.fork(s)
        tc.record('inference1_finish')
        out = agg((logits.cpu(),), None, tc)
    _, cls, merged = out
    assert int(cls) == R.video_class(want)
    assert merged.sub_id is None
    for e in engines.values():
        e.close()


# ------------------------------------------------------------------------------ two processes
def _producer(ring, n_items, dev_index, done):
    import torch
    from rnb_b200 import handoff
    torch.cuda.set_device(dev_index)
    dev = torch.device('cuda', dev_index)
    stream = torch.cuda.Stream(device=dev)
    with torch.cuda.stream(stream):
        for i in range(n_items):
            te = ring[i % len(ring)]
            te.event.wait()                              # runner.py:163
            a = torch.randn(2048, 2048, device=dev)
            for _ in range(10):
                a = (a @ a).clamp_(-1, 1)                # keep the stream busy ahead of the write
            seg = torch.full((3, 400), float(i + 1), device=dev)
            handoff.publish(te, seg)
            te.event.clear()                             # runner.py:173
            done.put(i)                                  # the Signal of runner.py:221-228
        stream.synchronize()
    done.put(None)


def _consumer(ring, dev_index, done, results):
    import torch
    from rnb_b200 import handoff
    torch.cuda.set_device(dev_index)
    dev = torch.device('cuda', dev_index)
    stream = torch.cuda.Stream(device=dev)
    placeholder = torch.zeros(5, 400, device=dev)
    sums = []
    with torch.cuda.stream(stream):
        while True:
            i = done.get()
            if i is None:
                break
            te = ring[i % len(ring)]
            x = handoff.fetch(placeholder, te)            # runner.py:107-114
            te.event.set()                                # runner.py:117, right away
            sums.append((i, x.sum()))
        stream.synchronize()
    results.put([(i, float(s.item())) for i, s in sums])


def test_two_processes_share_buffers_and_events(lib):
    """Producer and consumer are spawned processes, as RnB's runners are (benchmark.py:132,260-269):
    the buffers cross as CUDA-IPC tensors, `ready` / `free` as interprocess events, and the buffers
    are recycled (2 buffers, 8 items) without any host-side wait for the GPU."""
    import torch.multiprocessing as mp
    from rnb_b200 import handoff
    ctx = mp.get_context('spawn')
    p_dev, c_dev = 0, (1 if _ngpu() > 1 else 0)
    ring = [handoff.TensorEvent(((5, 400),), 'cuda:%d' % p_dev, ctx=ctx) for _ in range(2)]
    done, results = ctx.Queue(), ctx.Queue()
    n_items = 8
    env_root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    # the spawned workers unpickle functions of this module: make it (and the package) importable
    os.environ['PYTHONPATH'] = os.pathsep.join(
        [env_root, os.path.join(env_root, 'tests'), os.environ.get('PYTHONPATH', '')])
    procs = [ctx.Process(target=_producer, args=(ring, n_items, p_dev, done)),
             ctx.Process(target=_consumer, args=(ring, c_dev, done, results))]
    for p in procs:
        p.start()
    got = results.get(timeout=300)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert got == [(i, 3 * 400 * float(i + 1)) for i in range(n_items)]
