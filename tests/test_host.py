"""CPU tests of the host side: C-ABI symbols, key/shape tables, step-class contract, hand-off
arithmetic, TimeCard fork/merge, configs.  No compute call is made (there is no GPU here)."""
import json
import os
import re

import pytest
import torch

from oracle import r2p1d_ref as R

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_library_exports_every_declared_symbol(lib):
    from rnb_b200 import _lib
    header = open(os.path.join(ROOT, 'include', 'rnb_b200.h')).read()
    declared = set(re.findall(r'\b(rnb_[a-z0-9_]+)\s*\(', header))
    assert declared, 'no declarations found in include/rnb_b200.h'
    assert declared == set(_lib.PROTOTYPES), declared ^ set(_lib.PROTOTYPES)
    for name in declared:
        assert hasattr(lib, name), name


def test_cupti_library_exports_every_declared_symbol():
    from rnb_b200 import cupti
    header = open(os.path.join(ROOT, 'include', 'rnb_cupti.h')).read()
    declared = set(re.findall(r'\b(rnb_cupti_[a-z0-9_]+)\s*\(', header))
    assert declared == set(cupti.PROTOTYPES), declared ^ set(cupti.PROTOTYPES)
    lib = cupti.load()
    for name in declared:
        assert hasattr(lib, name), name


@pytest.mark.parametrize('name', ['rnb_b200.h', 'rnb_cupti.h'])
def test_every_entry_point_cites_the_reference(name):
    header = open(os.path.join(ROOT, 'include', name)).read()
    for block in re.findall(r'/\*(?:[^*]|\*(?!/))*\*/\s*(?:int|double|void|const char\*) rnb_\w+', header):
        name = re.search(r'rnb_\w+$', block).group(0)
        assert re.search(r'\w+\.(py|cpp):\d+', block) or 'no counterpart in the reference' in block.lower(), \
            '%s lacks a reference citation' % name


def test_no_gpu_fails_loudly(lib):
    if torch.cuda.is_available():
        pytest.skip('GPU present')
    from rnb_b200 import _lib
    assert lib.rnb_init(0) != 0
    assert lib.rnb_last_error()
    from rnb_b200.engine import R2P1DEngine
    with pytest.raises(Exception):
        R2P1DEngine('cuda:0', {})
    with pytest.raises(RuntimeError):
        R2P1DEngine('cpu', {})


def test_create_validates_like_the_reference_runner(lib):
    import ctypes
    h = ctypes.c_void_p()
    assert lib.rnb_r2p1d_create(None, 0, 0, 5, 400, 1, 8, ctypes.byref(h)) == -1   # model.py:38-40
    assert b'start_index' in lib.rnb_last_error()
    assert lib.rnb_r2p1d_create(None, 0, 1, 6, 400, 1, 8, ctypes.byref(h)) == -1   # model.py:42-44
    assert b'end_index' in lib.rnb_last_error()
    ls = (ctypes.c_int * 4)(3, 4, 6, 3)
    assert lib.rnb_r2p1d_create(ls, 4, 1, 5, 400, 1, 8, ctypes.byref(h)) == -1     # needs 5 entries
    assert lib.rnb_r2p1d_create(None, 0, 1, 5, 400, 1, 8, ctypes.byref(h)) == 0
    assert lib.rnb_r2p1d_forward(h, None, 0, 1, None, 3, None) == -4               # not finalized
    shape = (ctypes.c_int64 * 4)()
    assert lib.rnb_r2p1d_io_shape(h, 0, shape) == 0 and list(shape) == [3, 8, 112, 112]
    assert lib.rnb_r2p1d_io_shape(h, 1, shape) == 0 and list(shape) == [400, 1, 1, 1]
    lib.rnb_r2p1d_destroy(h)


@pytest.mark.parametrize('kw', [dict(), dict(layer_sizes=[0, 3, 4, 6, 3]),
                                dict(start_index=3, end_index=4), dict(start_index=2, end_index=5)])
def test_state_dict_spec_matches_oracle_keys(kw):
    from rnb_b200.engine import state_dict_spec, random_state_dict
    start, end = kw.get('start_index', 1), kw.get('end_index', 5)
    m = R.build(start, end, layer_sizes=kw.get('layer_sizes'))
    sd = {k: v for k, v in m.state_dict().items() if not k.endswith('num_batches_tracked')}
    spec = dict(state_dict_spec(**kw))
    assert set(spec) == set(sd)
    assert all(tuple(sd[k].shape) == tuple(spec[k]) for k in sd)
    rnd = random_state_dict(seed=1, **kw)
    m.load_state_dict(rnd, strict=False)  # shapes fit the oracle's modules


def test_step_classes_contract():
    """What RnB's main process relies on before spawning (control.py:125-131)."""
    from rnb_b200 import steps
    from rnb_b200.runner_model import RunnerModel
    for cls in (steps.R2P1DRunner, steps.R2P1DLoader, steps.R2P1DSingleStep, steps.R2P1DAggregator,
                steps.Batcher):
        assert issubclass(cls, RunnerModel)
    assert steps.R2P1DLoader.output_shape() == ((15, 3, 8, 112, 112),)     # model.py:156-158
    assert steps.Batcher.output_shape() == ((15, 3, 8, 112, 112),)         # batcher.py:13-15
    assert steps.R2P1DSingleStep.output_shape() == ((15, 400),)            # model.py:233-235
    assert steps.R2P1DRunner.output_shape()[0][1] == 400
    assert steps.R2P1DAggregator.output_shape() is None                    # model.py:283-285
    assert steps.R2P1DRunner.input_dict[4] == (10, 128, 4, 28, 28)         # model.py:29-33
    sel = steps.LargeSmallSelector(2)

    class TC:
        pass
    tc = TC()
    assert sel.select(None, None, tc) == 0
    tc.num_clips = 15
    assert sel.select(None, None, tc) == 1
    tc.num_clips = 1
    assert sel.select(None, None, tc) == 0
    it = iter(steps.R2P1DVideoPathIterator())
    assert next(it) != next(it)


def test_runner_rejects_bad_layer_indices_like_the_reference(capsys):
    from rnb_b200 import steps
    with pytest.raises(SystemExit):
        steps.R2P1DRunner(torch.device('cpu'), start_index=0)
    assert 'start_index (0)' in capsys.readouterr().out
    with pytest.raises(SystemExit):
        steps.R2P1DRunner(torch.device('cpu'), end_index=6)


def test_aggregator_on_cpu_tensors_follows_the_reference():
    """The reference places the aggregator on the CPU (config/r2p1d-segment.json, gpus [-1])."""
    from rnb_b200 import steps
    from rnb_b200.timecard import TimeCard
    g = torch.Generator().manual_seed(0)
    parts = [torch.randn(n, 400, generator=g) for n in (5, 5, 5)]
    agg = steps.R2P1DAggregator(torch.device('cpu'), aggregate=3)
    parent = TimeCard(3)
    parent.record('enqueue_filename')
    outs = []
    for s, p in enumerate(parts):
        tc = parent.fork(s)
        tc.record('inference1_finish')
        outs.append(agg((p,), None, tc))
    assert outs[0] == (None, None, None) and outs[1] == (None, None, None)
    _, cls, merged = outs[2]
    assert int(cls) == int(torch.cat(parts).numpy().sum(axis=0).argmax())
    assert list(merged.timings) == ['enqueue_filename', 'inference1_finish-0',
                                    'inference1_finish-1', 'inference1_finish-2']
    one = steps.R2P1DAggregator(torch.device('cpu'), aggregate=1)
    _, cls, tc = one((parts[0],), None, parent)
    assert int(cls) == int(parts[0].numpy().sum(axis=0).argmax()) and tc is parent


def test_timecard_fork_merge_rules():
    from rnb_b200.timecard import TimeCard, TimeCardList
    tc = TimeCard(1)
    tc.record('a')
    tc.add_gpu(0)
    kids = [tc.fork(i) for i in range(2)]
    with pytest.raises(Exception):
        kids[0].fork(0)                                  # rnb_logging.py:60-62
    for i, k in enumerate(kids):
        k.record('b')
        k.add_gpu(i + 1)
    merged = TimeCard.merge(kids[::-1])
    assert list(merged.timings) == ['a', 'b-0', 'b-1']
    assert merged.gpus == [(0,), (1, 2)]
    kids[1].record('c')
    with pytest.raises(Exception):
        TimeCard.merge(kids)
    with pytest.raises(NotImplementedError):
        TimeCardList(kids).fork(0)                       # rnb_logging.py:141-142


def test_handoff_arithmetic():
    from rnb_b200 import handoff
    assert handoff.segment_ranges(11, 3) == R.segment_ranges(11, 3) == [(0, 4), (4, 8), (8, 11)]
    for rows in range(0, 20):
        for s in (1, 2, 3, 7):
            r = handoff.segment_ranges(rows, s)
            assert r[0][0] == 0 and r[-1][1] == rows
            assert all(a[1] == b[0] for a, b in zip(r, r[1:]))
            assert max(hi - lo for lo, hi in r) <= R.segmented_rows(max(rows, 1), s)
    assert handoff.get_segmented_shapes(((10, 400),), 3) == [(4, 400)]   # control.py:63-66
    assert handoff.get_segmented_shapes(None, 3) is None
    with pytest.raises(ValueError):
        handoff.get_segmented_shapes(((2, 400),), 3)


def test_configs_name_importable_classes():
    import importlib
    cfg_dir = os.path.join(ROOT, 'config')
    names = sorted(os.listdir(cfg_dir))
    assert 'r2p1d-whole.json' in names and 'r2p1d-segment.json' in names
    for name in names:
        cfg = json.load(open(os.path.join(cfg_dir, name)))
        paths = [cfg['video_path_iterator']] + [s['model'] for s in cfg['pipeline']]
        for s in cfg['pipeline']:
            paths += [g['queue_selector'] for g in s['queue_groups'] if 'queue_selector' in g]
        for p in paths:
            mod, cls = p.rsplit('.', 1)
            assert hasattr(importlib.import_module(mod), cls), p


def test_tensor_event_on_the_host_has_the_reference_attributes():
    """control.TensorEvent's attributes (control.py:40-46) plus a dtype per tensor; no CUDA needed
    for a CPU buffer (the aggregator of r2p1d-segment.json lives on the host)."""
    from rnb_b200 import handoff
    te = handoff.TensorEvent(((5, 400), (5, 8, 3, 112, 112)), 'cpu', dtype=(torch.float32, torch.uint8))
    assert [t.dtype for t in te.tensors] == [torch.float32, torch.uint8]
    assert [tuple(t.shape) for t in te.tensors] == [(5, 400), (5, 8, 3, 112, 112)]
    assert te.event.is_set() and len(te.valid_batch_sizes) == 2
    assert te.ready is None and te.free is None
    with pytest.raises(ValueError):
        handoff.TensorEvent(((5, 400),), 'cpu', dtype=(torch.float32, torch.uint8))


def test_wire_format_of_the_step_classes(monkeypatch):
    """RNB_B200_WIRE decides, statically, what the loader puts on the wire (control.py:125-131
    evaluates output_shape() in the main process)."""
    import importlib
    from rnb_b200 import steps
    assert steps.R2P1DLoader.output_shape() == ((steps.MAX_CLIPS, 3, 8, 112, 112),)
    assert steps.R2P1DLoader.output_dtype() == (torch.float32,)
    monkeypatch.setenv('RNB_B200_WIRE', 'uint8')
    try:
        importlib.reload(steps)
        assert steps.R2P1DLoader.output_shape() == ((steps.MAX_CLIPS, 8, 3, 112, 112),)
        assert steps.R2P1DLoader.output_dtype() == (torch.uint8,)
        assert steps.Batcher.output_dtype() == (torch.uint8,)
    finally:
        monkeypatch.delenv('RNB_B200_WIRE')
        importlib.reload(steps)
    assert steps._default_prepare_batches(64) == sorted({1, 3, 4, 5, 10, 15, steps.MAX_CLIPS})


def test_latency_percentiles_from_a_timecard_log(tmp_path):
    """scripts/latency_percentiles.py on a log in the reference's format (rnb_logging.py:188-214,
    benchmark.py:297-300): latency = last stamp - enqueue_filename."""
    import sys
    sys.path.insert(0, os.path.join(ROOT, 'scripts'))
    import latency_percentiles
    job = tmp_path / 'logs' / 'job'
    job.mkdir(parents=True)
    rows = ['enqueue_filename runner0_start inference0_start inference0_finish gpu0']
    for i in range(100):
        t0 = 1000.0 + i * 0.01
        rows.append('%f %f %f %f 0' % (t0, t0 + 0.001, t0 + 0.002, t0 + 0.002 + 0.001 * (i + 1)))
    (job / 'g0-group0-0.txt').write_text('\n'.join(rows) + '\n')
    (job / 'log-meta.txt').write_text(
        "Args: Namespace(mean_interval_ms=0, batch_size=1, videos=100, queue_size=50000)\n"
        "1000.000000 1002.000000\nTermination flag: 0\n")
    out = latency_percentiles.parse_job(str(job), skip=0)
    assert out['videos_logged'] == 100 and out['termination_flag'] == 0
    assert abs(out['videos_per_s'] - 50.0) < 1e-6
    assert abs(out['latency_ms']['p50'] - (2 + 51)) < 1e-3      # 51st of 1..100 ms on top of 2 ms
    assert abs(out['latency_ms']['p99'] - (2 + 100)) < 1e-3
    assert out['keys'][0] == 'enqueue_filename'


def test_harness_topologies_are_valid_pipelines():
    """tools/harness_sweep.py cuts the config/*.json mirrors down to N GPUs; the result must still
    satisfy the checks of the reference's sanity_check (benchmark.py:46-91)."""
    import sys
    sys.path.insert(0, os.path.join(ROOT, 'tools'))
    import harness_sweep
    for kind, n in [('whole', 1), ('aggressive', 2), ('aggressive', 8), ('nopipeline', 4),
                    ('fourgroups', 1), ('fourgroups', 4), ('segment', 4), ('rnb', 1), ('rnb', 8)]:
        cfg = harness_sweep.make_config(kind, n)
        out_queues = None
        for i, step in enumerate(cfg['pipeline']):
            assert isinstance(step['model'], str) and step['model'].startswith('rnb_b200.steps.')
            gpus = [g for grp in step['queue_groups'] for g in grp['gpus']]
            assert gpus and all(g < max(n, 1) for g in gpus)
            if i > 0:
                assert set(g['in_queue'] for g in step['queue_groups']) == out_queues
            if i < len(cfg['pipeline']) - 1:
                out_queues = set(q for g in step['queue_groups'] for q in g['out_queues'])


def test_kernel_kind_names_match_the_header():
    """engine.KERNEL_KINDS[k] is the name of kind k of rnb_r2p1d_op_info / force_kernel (include/rnb_b200.h)."""
    import re
    from rnb_b200.engine import R2P1DEngine
    header = open(os.path.join(os.path.dirname(__file__), '..', 'include', 'rnb_b200.h')).read()
    doc = re.sub(r'\s*\n \*\s*', ' ', header[header.index('rnb_r2p1d_op_info() describes'):])
    for k, name in enumerate(R2P1DEngine.KERNEL_KINDS):
        if k == 0:
            assert '0 = fp32 direct convolution' in doc
            continue
        m = re.search(r'\b%d = (\w+)' % k, doc)
        assert m, 'kind %d is not documented in the header' % k
        kernel = m.group(1)          # e.g. conv_pair_kernel
        assert kernel.startswith(name.replace('conv_tpatch', 'conv_pair')), (k, name, kernel)
