#!/usr/bin/env python
"""Run the reference's own benchmark driver (benchmark.py, unmodified) on this repository's steps.

    python scripts/run_rnb.py -c config/r2p1d-whole.json -v 500 -mi 0 [--wire uint8] [--max-clips 15]
                              [--logdir DIR] [--reference /path/to/rnb]

Everything after the options this wrapper knows is handed to benchmark.py as its own command line
(reference benchmark.py:150-167: -c config, -v videos, -mi mean interval in ms, 0 = bulk client,
-qs queue size).  The reference's client, runner loop, queues, shared tensors, TimeCards and log
files are used as they are; the pipeline file names `rnb_b200.steps.*` classes, which is the
whole integration.  Three things are arranged around the unmodified driver, none by editing it:

  * `py3nvml` (benchmark.py:44) is not installed here; a stand-in module backed by `pynvml` is
    registered under that name.  Its memory query reports 0 bytes used for a GPU that runs no
    compute process, because benchmark.py:105 treats `used == 0` as "free" and current drivers
    never report zero (RNB_B200_ALLOW_BUSY_GPUS=1 reports every GPU as free, for runs started from
    a process that itself holds a CUDA context, such as the test suite).
  * CUDA_VISIBLE_DEVICES (benchmark.py:94) defaults to all GPUs.
  * --runtime b200: `runner` resolves to rnb_b200.runtime (the same worker protocol with the
    event-fenced hand-off) instead of the reference's runner.py.
  * --wire uint8: shared tensors are built by rnb_b200.handoff.TensorEvent with the dtype the
    producing step declares (`output_dtype()`), instead of control.py:41's fixed fp32, so the loader
    ships raw uint8 frames (a quarter of the bytes); control.TensorEvent / control.load_class are
    rebound in this process only, before benchmark.py builds its SharedQueuesAndTensors.

Logs land in <logdir>/logs/<job id>/ exactly as benchmark.py writes them; scripts/latency_percentiles.py
turns them into videos/s, clips/s and p50 / p99 per-video latency.
"""
import argparse
import os
import runpy
import sys
import types

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def find_reference(explicit=None):
    for cand in (explicit, os.environ.get('RNB_REFERENCE'), os.path.join(ROOT, 'baseline', '_ref', 'rnb'),
                 '/root/reference'):
        if cand and os.path.isfile(os.path.join(cand, 'benchmark.py')):
            return os.path.abspath(cand)
    return None


def install_py3nvml_shim():
    import pynvml
    shim = types.ModuleType('py3nvml.py3nvml')

    class _Mem:
        def __init__(self, used, total, free):
            self.used, self.total, self.free = used, total, free

    allow_busy = os.environ.get('RNB_B200_ALLOW_BUSY_GPUS') == '1'

    def mem_info(handle):
        m = pynvml.nvmlDeviceGetMemoryInfo(handle)
        try:
            busy = not allow_busy and len(pynvml.nvmlDeviceGetComputeRunningProcesses(handle)) > 0
        except Exception:  # noqa: BLE001
            busy = False
        return _Mem(m.used if busy else 0, m.total, m.free)

    shim.nvmlInit = pynvml.nvmlInit
    shim.nvmlShutdown = pynvml.nvmlShutdown
    shim.nvmlDeviceGetCount = pynvml.nvmlDeviceGetCount
    shim.nvmlDeviceGetHandleByIndex = pynvml.nvmlDeviceGetHandleByIndex
    shim.nvmlDeviceGetMemoryInfo = mem_info
    pkg = types.ModuleType('py3nvml')
    pkg.py3nvml = shim
    sys.modules['py3nvml'] = pkg
    sys.modules['py3nvml.py3nvml'] = shim


def install_dtype_aware_buffers():
    """control.TensorEvent -> rnb_b200.handoff.TensorEvent with the producing step's dtype."""
    import control
    import torch
    from rnb_b200 import handoff
    state = {'dtype': torch.float32}
    ref_load_class = control.load_class

    def load_class(path):
        cls = ref_load_class(path)
        fn = getattr(cls, 'output_dtype', None)
        state['dtype'] = fn() if fn is not None else torch.float32
        return cls

    def tensor_event(shapes, device, dtype=None):
        return handoff.TensorEvent(shapes, device, dtype=dtype if dtype is not None else state['dtype'])

    control.load_class = load_class
    control.TensorEvent = tensor_event


def main():
    ap = argparse.ArgumentParser(add_help=False)
    ap.add_argument('--reference', default=None)
    ap.add_argument('--wire', choices=['fp32', 'uint8'], default='fp32')
    ap.add_argument('--runtime', choices=['reference', 'b200'], default='reference',
                    help="worker loop: the reference's runner.py, or rnb_b200.runtime (same protocol, "
                         'event-fenced hand-off instead of copy_ + host synchronisation)')
    ap.add_argument('--max-clips', type=int, default=15)
    ap.add_argument('--logdir', default=os.path.join(ROOT, 'gpurun_out', 'rnb'))
    ap.add_argument('--plan-cache', default=None)
    args, rest = ap.parse_known_args()
    ref = find_reference(args.reference)
    if ref is None:
        sys.exit('run_rnb.py: no RnB checkout found (set RNB_REFERENCE or run '
                 'scripts/stage_reference.py where /root/reference exists)')
    # the step classes read these at import, in the main process and in every spawned worker
    os.environ['RNB_B200_WIRE'] = args.wire
    os.environ['RNB_B200_MAX_CLIPS'] = str(args.max_clips)
    if args.plan_cache:
        os.environ['RNB_B200_PLAN_CACHE'] = os.path.abspath(args.plan_cache)
    os.environ['PYTHONPATH'] = os.pathsep.join([ref, ROOT, os.environ.get('PYTHONPATH', '')])
    if 'CUDA_VISIBLE_DEVICES' not in os.environ:
        import pynvml
        pynvml.nvmlInit()
        os.environ['CUDA_VISIBLE_DEVICES'] = ','.join(
            str(i) for i in range(pynvml.nvmlDeviceGetCount()))
    sys.path[:0] = [ref, ROOT]
    install_py3nvml_shim()
    # config paths on the command line are relative to where the user is; logs go to --logdir
    rest = [os.path.abspath(a) if a.endswith('.json') and os.path.exists(a) else a for a in rest]
    os.makedirs(args.logdir, exist_ok=True)
    os.chdir(args.logdir)
    if args.wire != 'fp32' or args.runtime == 'b200':
        install_dtype_aware_buffers()
    if args.runtime == 'b200':
        # benchmark.py does `from runner import runner` (benchmark.py:148): hand it our loop
        import rnb_b200.runtime
        sys.modules['runner'] = rnb_b200.runtime
    sys.argv = [os.path.join(ref, 'benchmark.py')] + rest
    runpy.run_path(sys.argv[0], run_name='__main__')


if __name__ == '__main__':
    main()
