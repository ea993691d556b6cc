#!/usr/bin/env python
"""Throughput of the R(2+1)D hot path on B200: clips/s, roofline of the convolution kernel, an
end-to-end number with host buffers, and the CPU oracle timed beside them.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--batch 16] [--model r18|r34]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 \
        --master-port P bench.py --gpus N --steps K --warmup W
    python bench.py --impl reference ...       # the CPU implementation of the same path

A step is one forward pass over one batch of synthetic uint8 clips.  Default workload =
BASELINE.json configs[1]: R(2+1)D-18, bf16, batch 16, one B200 (one replica per GPU for N > 1,
weak scaling, no collective on the data path).  Rank 0 prints exactly one JSON line.
"""
import argparse
import json
import os
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

LAYER_SIZES = {'r18': [2, 2, 2, 2, 2], 'r34': [2, 3, 4, 6, 3]}  # runner form: stem slot first
GFLOP_PER_CLIP = {'r18': 42.172, 'r34': 76.853}
METRIC = 'R(2+1)D clips/sec (8x112x112 clips)'
CLIP_BYTES = 8 * 3 * 112 * 112
NCU_LAUNCHES = os.path.join(ROOT, 'profiles', 'r02_ncu_launches.json')  # see ncu_traffic()
FALLBACK_PEAK_TFLOPS = 1590.0  # /opt/skills/guides/B200_PROFILING.md, used only without MEASURED_PEAKS.json


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument('--gpus', type=int, default=1)
    ap.add_argument('--steps', type=int, default=200)
    ap.add_argument('--warmup', type=int, default=20)
    ap.add_argument('--batch', type=int, default=16)
    ap.add_argument('--model', choices=sorted(LAYER_SIZES), default='r18')
    ap.add_argument('--dtype', choices=['bf16', 'fp32'], default='bf16')
    ap.add_argument('--impl', choices=['b200', 'reference'], default='b200')
    ap.add_argument('--no-cpu-baseline', action='store_true')
    ap.add_argument('--ref-clips', type=int, default=0,
                    help='clips per step of the CPU arms (0: the whole batch when the run fits a few '
                         'minutes, else a bounded sample of it)')
    ap.add_argument('--profile-table', default=None,
                    help='write the per-kernel timing table (markdown) to this path')
    return ap.parse_args()


def base_config(args, world):
    """The `config` object of the JSON line; both arms print the same one."""
    B = args.batch
    pool_n = input_pool_batches(B)
    return {'workload': workload_name(args), 'batch': B, 'clip': '3x8x112x112 uint8',
            'weights': 'random-init (seed 0), BatchNorm folded (inference statistics)',
            'parallelism': 'replica x%d, no collective' % world,
            'l2': 'inputs larger than L2: %d distinct %.1f MB uint8 batches (%.0f MB) '
                  'round-robin; per-step activation traffic also exceeds the 126 MB L2'
                  % (pool_n, B * CLIP_BYTES / 1e6, pool_n * B * CLIP_BYTES / 1e6),
            'e2e_path': 'rnb_b200.engine.HostPipeline: pinned uint8 frames -> H2D (copy '
                        'stream, overlaps the previous forward) -> R2P1DEngine.forward (fused '
                        'preprocess) -> D2H logits to pinned memory, read on the host every step'}


def input_pool_batches(batch):
    return max(2, -(-160 * 1024 * 1024 // (batch * CLIP_BYTES)))


def ncu_traffic(model, dtype, batch, world):
    """DRAM bytes (read + write) per convolution launch from the committed ncu capture of this
    configuration (profiles/r02_ncu_launches.json, written by tools/ncu_summary.py), or None."""
    try:
        cap = json.load(open(NCU_LAUNCHES))
    except Exception:  # noqa: BLE001
        return None, None
    key = '%s-%s-b%d' % (model, dtype, batch)
    e = cap.get(key)
    if not e or world != 1:
        return None, None
    return e.get('dram_bytes_per_conv_launch'), e.get('source')


def workload_name(args):
    return 'R(2+1)D-%s %s batch=%d single B200 (r2p1d-whole.json, one replica%s)' % (
        args.model[1:], args.dtype, args.batch, ' per GPU' if args.gpus > 1 else '')


# ------------------------------------------------------------------------------ clocks
class ClockSampler(threading.Thread):
    """Samples SM clock and throttle reasons of one GPU through NVML while the bench runs."""
    REASONS = {0x4: 'sw_power_cap', 0x8: 'hw_slowdown', 0x20: 'sw_thermal_slowdown',
               0x40: 'hw_thermal_slowdown', 0x80: 'hw_power_brake_slowdown'}

    def __init__(self, index, period=0.02):
        super().__init__(daemon=True)
        self.index, self.period = index, period
        self.samples = []  # (time, sm_mhz, reasons_bits)
        self.max_mhz = None
        self.ok = False
        self._stop_flag = threading.Event()
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.handle = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = int(pynvml.nvmlDeviceGetMaxClockInfo(self.handle, pynvml.NVML_CLOCK_SM))
            self.ok = True
        except Exception as e:  # noqa: BLE001
            self.err = repr(e)

    def run(self):
        if not self.ok:
            return
        nv = self.nv
        while not self._stop_flag.is_set():
            try:
                mhz = int(nv.nvmlDeviceGetClockInfo(self.handle, nv.NVML_CLOCK_SM))
                try:
                    bits = int(nv.nvmlDeviceGetCurrentClocksEventReasons(self.handle))
                except Exception:  # noqa: BLE001
                    bits = int(nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.handle))
                self.samples.append((time.time(), mhz, bits))
            except Exception:  # noqa: BLE001
                pass
            time.sleep(self.period)

    def stop(self):
        self._stop_flag.set()

    def summary(self, t0, t1):
        if not self.ok:
            return {'sm_mhz': None, 'sm_max_mhz': None, 'reasons': [], 'note': 'NVML unavailable'}
        inside = [s for s in self.samples if t0 <= s[0] <= t1] or self.samples
        mhz = sorted(s[1] for s in inside)
        bits = 0
        for s in inside:
            bits |= s[2]
        return {'sm_mhz': mhz[len(mhz) // 2] if mhz else None, 'sm_max_mhz': self.max_mhz,
                'reasons': [n for b, n in sorted(self.REASONS.items()) if bits & b],
                'samples': len(inside)}


# ------------------------------------------------------------------------------ CPU arm
def time_oracle(model_name, clips_per_step, steps, warmup, threads=None):
    """Times oracle/r2p1d_ref.py (the CPU port of the path) on this host.  Returns clips/s."""
    import torch
    from oracle import r2p1d_ref as oracle
    threads = threads or os.cpu_count() or 1
    torch.set_num_threads(threads)
    ls = LAYER_SIZES[model_name]
    model = oracle.build(layer_sizes=ls, seed=0)
    clips = oracle.synthetic_clips(clips_per_step, seed=1234)
    for _ in range(warmup):
        oracle.forward_u8(model, clips)
    t0 = time.perf_counter()
    for _ in range(steps):
        oracle.forward_u8(model, clips)
    dt = time.perf_counter() - t0
    return clips_per_step * steps / dt, dt / steps, torch.get_num_threads()


def pick_cpu_sample(args, steps, budget_s):
    """Clips per step for a CPU arm: the whole batch if `steps` of it fit `budget_s`, else the
    largest power-of-two sample that does.  Returns (clips, seconds per step at that size)."""
    if args.ref_clips > 0:
        clips = min(args.ref_clips, args.batch)
        _, per_step, _ = time_oracle(args.model, clips, 1, 1)
        return clips, per_step
    clips = args.batch
    while True:
        _, per_step, _ = time_oracle(args.model, clips, 1, 1)
        if per_step * steps <= budget_s or clips == 1:
            return clips, per_step
        clips = max(1, clips // 2)


def run_reference(args, rank):
    """--impl reference: the CPU implementation of the path on the host cores, rank 0 only.

    Times exactly --steps steps after --warmup warm-up steps; a step is the forward pass over the
    same batch the GPU arm uses, or -- when that would not end within a few minutes -- over a bounded
    sample of it (stated in `cpu_baseline.sample`; `ms_per_step` is what was timed)."""
    if rank != 0:
        return
    steps, warmup = max(1, args.steps), max(0, args.warmup)
    clips, _ = pick_cpu_sample(args, steps + warmup, 150.0)
    value, per_step, threads = time_oracle(args.model, clips, steps, warmup)
    sample = ('%d timed steps of %d clips each%s, oracle/r2p1d_ref.py (torch fp32 CPU port; the '
              'reference\'s own submodule is absent)'
              % (steps, clips, '' if clips == args.batch
                 else ' (a sample of the batch-%d step)' % args.batch))
    line = {
        'impl': 'reference', 'metric': METRIC, 'value': value, 'unit': 'clips/s',
        'n_gpus': args.gpus, 'steps': steps, 'warmup': warmup,
        'ms_per_step': per_step * 1e3, 'clips_per_step': clips,
        'higher_is_better': True, 'scaling': 'weak', 'vs_baseline': None, 'dtype': 'f32',
        'data': 'synthetic', 'config': base_config(args, args.gpus),
        'cpu_baseline': {'value': value, 'unit': 'clips/s', 'cores': threads, 'kind': 'port',
                         'sample': sample},
        'e2e': {'value': value, 'unit': 'clips/s', 'h2d_bytes_per_step': 0,
                'd2h_bytes_per_step': 0},
        'gpu_launches': 0,
    }
    _emit(line)


# ------------------------------------------------------------------------------ GPU arm
def peaks():
    path = os.path.join(ROOT, 'MEASURED_PEAKS.json')
    try:
        p = json.load(open(path))
        return float(p['bf16_tflops']), float(p.get('bf16_tflops_sustained', p['bf16_tflops'])), \
            'MEASURED_PEAKS.json bf16_tflops (burst, of measured)'
    except Exception:  # noqa: BLE001
        return FALLBACK_PEAK_TFLOPS, 1400.0, 'fallback 1.59 PFLOP/s (of fallback)'


def run_plugin_leg(args, dev, sd, ls, K, W, rdist, barrier):
    """clips/s through rnb_b200.steps.R2P1DLoader -> handoff -> R2P1DRunner driven the way
    runner.py:80-173 drives them (uint8 wire, event-fenced copies, no queues or processes)."""
    import torch
    os.environ['RNB_B200_MAX_CLIPS'] = str(max(args.batch, 15))
    os.environ['RNB_B200_WIRE'] = 'uint8'
    from rnb_b200 import handoff, steps
    from rnb_b200.timecard import TimeCard
    B = args.batch
    loader = steps.R2P1DLoader(dev, num_clips_population=[B], num_clips_weights=[1], seed=0)
    runner = steps.R2P1DRunner(dev, layer_sizes=ls, state_dict=sd, dtype=args.dtype,
                               max_batch=max(B, 16), prepare_batches=[B], warmup=1)
    ring = [handoff.TensorEvent(steps.R2P1DLoader.output_shape(), dev,
                                dtype=steps.R2P1DLoader.output_dtype()) for _ in range(4)]
    placeholder = torch.zeros_like(ring[0].tensors[0])
    host = [torch.empty((B, 400), dtype=torch.float32).pin_memory() for _ in range(2)]
    done = [torch.cuda.Event() for _ in range(2)]

    def step(i):
        tc = TimeCard(i)
        (frames,), _, tc = loader(None, 'synthetic/video_%07d.mp4' % i, tc)
        te = ring[i % len(ring)]
        n = handoff.publish(te, frames)
        x = handoff.fetch(placeholder, te, n)
        (logits,), _, tc = runner((x,), None, tc)
        done[i % 2].synchronize()          # the host consumed this slot's previous result
        host[i % 2].copy_(logits, non_blocking=True)
        done[i % 2].record()

    for i in range(W):
        step(i)
    barrier()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ev0.record()
    for i in range(K):
        step(i)
    ev1.record()
    barrier()
    value, ms = rdist.whole_job_throughput(K * B, ev0.elapsed_time(ev1), dev)
    return {'value': value, 'unit': 'clips/s', 'ms_per_step': ms / K,
            'path': 'steps.R2P1DLoader.__call__ -> handoff.publish/fetch (uint8 TensorEvent ring, '
                    'CUDA-event fenced) -> steps.R2P1DRunner.__call__ -> D2H logits; device-resident '
                    'synthetic frames as in RnB, where the decoder writes GPU memory'}


def run_b200(args, rank, world, local_rank):
    import torch
    import torch.distributed as dist
    from rnb_b200 import dist as rdist
    from rnb_b200.engine import R2P1DEngine, random_state_dict

    if not torch.cuda.is_available():
        raise RuntimeError('bench.py needs a CUDA device; there is no CPU fallback for the product path')
    torch.cuda.set_device(local_rank)
    dev = torch.device('cuda', local_rank)
    if world > 1:
        dist.init_process_group('nccl', device_id=dev)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    B, K, W = args.batch, max(1, args.steps), max(3, args.warmup)
    ls = LAYER_SIZES[args.model]
    sd = random_state_dict(1, 5, 400, ls, seed=0)
    engine = R2P1DEngine(dev, sd, 1, 5, 400, ls, dtype=args.dtype, max_batch=max(B, 16))
    # kernel selection ahead of the loop (the per-video latency leg below runs 1 and 15 clips)
    engine.prepare(B, 1, 15)

    # inputs larger than L2: a pool of distinct uint8 batches (> 126 MB) visited round-robin
    pool_n = input_pool_batches(B)
    g = torch.Generator(device=dev).manual_seed(1234 + rank)
    pool = torch.randint(0, 256, (pool_n, B, 8, 3, 112, 112), dtype=torch.uint8, device=dev,
                         generator=g)
    sampler = ClockSampler(local_rank)
    sampler.start()

    # ---- device-resident throughput ("value")
    for i in range(W):
        engine(pool[i % pool_n])
    barrier()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    # lets `ncu --profile-from-start off` capture exactly the timed steps (never used for a number)
    mark_profile = os.environ.get('RNB_BENCH_CUDA_PROFILER') == '1'
    if mark_profile:
        torch.cuda.profiler.start()
    t_wall0 = time.time()
    ev0.record()
    for i in range(K):
        engine(pool[i % pool_n])
    ev1.record()
    barrier()
    t_wall1 = time.time()
    if mark_profile:
        torch.cuda.profiler.stop()
    value, ms = rdist.whole_job_throughput(K * B, ev0.elapsed_time(ev1), dev)
    clocks = sampler.summary(t_wall0, t_wall1)

    # ---- end to end through the public API with HOST buffers: pinned uint8 -> H2D -> forward
    #      (fused preprocess) -> D2H logits, every step
    from rnb_b200.engine import HostPipeline
    host_in = [torch.randint(0, 256, (B, 8, 3, 112, 112), dtype=torch.uint8).pin_memory()
               for _ in range(4)]
    pipe = HostPipeline(engine, B)

    prev = None
    for i in range(W):
        slot = pipe.submit(host_in[i % len(host_in)])
        if prev is not None:
            pipe.result(prev)
        prev = slot
    pipe.result(prev)
    barrier()
    ev0.record()
    prev, checksum = None, 0.0
    for i in range(K):
        slot = pipe.submit(host_in[i % len(host_in)])
        if prev is not None:
            # every step's logits are read on the host (one batch stays in flight meanwhile)
            checksum += float(pipe.result(prev)[0, 0])
        prev = slot
    checksum += float(pipe.result(prev)[0, 0])
    ev1.record()
    barrier()
    value_e2e, ms_e2e = rdist.whole_job_throughput(K * B, ev0.elapsed_time(ev1), dev)

    # ---- the same through the plugin classes RnB loads from config/r2p1d-whole.json: loader step
    #      -> shared buffer (publish / fetch) -> runner step -> logits to pinned host memory
    plugin = run_plugin_leg(args, dev, sd, ls, K, W, rdist, barrier)

    # ---- per-video latency (second half of BASELINE's metric): one video = 1 or 15 clips drawn
    #      10:1 like the reference's sampler (models/r2p1d/sampler.py:58-59), host frames in, class
    #      id out, one video at a time (no batching across videos), wall clock around each call
    lat = None
    if rank == 0:
        import random
        rng = random.Random(0)
        n_videos = 300
        host_videos = {n: torch.randint(0, 256, (n, 8, 3, 112, 112), dtype=torch.uint8).pin_memory()
                       for n in (1, 15)}
        dev_video = torch.empty((15, 8, 3, 112, 112), dtype=torch.uint8, device=dev)
        cls_host = torch.empty((), dtype=torch.int32).pin_memory()
        times = {1: [], 15: []}
        for i in range(n_videos + 20):
            n = rng.choices([1, 15], [10, 1])[0]
            t0 = time.perf_counter()
            dev_video[:n].copy_(host_videos[n], non_blocking=True)
            _, _, cls = engine.classify(dev_video[:n])
            cls_host.copy_(cls, non_blocking=True)
            torch.cuda.current_stream().synchronize()
            if i >= 20:
                times[n].append((time.perf_counter() - t0) * 1e3)
        allt = sorted(times[1] + times[15])

        def pct(v, q):
            v = sorted(v)
            return v[min(len(v) - 1, int(q * len(v)))] if v else None
        lat = {'unit': 'ms', 'videos': len(allt), 'p50': pct(allt, 0.50), 'p99': pct(allt, 0.99),
               'p50_1clip': pct(times[1], 0.50), 'p50_15clip': pct(times[15], 0.50),
               'path': 'pinned uint8 frames -> H2D -> forward + on-GPU sum/argmax -> class id D2H, '
                       'one video per call, clips per video 1 or 15 (10:1)'}

    # ---- roofline of the dominant kernel (conv_tc_kernel: all convolution launches), measured
    #      live with CUDA events between launches on the launching stream
    prof_steps = max(3, min(K, 30))
    conv_ms, other_ms = 0.0, 0.0
    per_op = None
    for i in range(prof_steps):
        m = engine.profile(pool[i % pool_n])
        conv_ms += sum(m[1:-1])
        other_ms += m[0] + m[-1]
        per_op = m if per_op is None else [a + b for a, b in zip(per_op, m)]
    conv_ms /= prof_steps
    other_ms /= prof_steps
    per_op = [v / prof_steps for v in per_op]
    sampler.stop()
    info = engine.op_info()
    peak_burst, peak_sust, peak_src = peaks()
    conv_flops = sum(f for _, f, _, _, _ in info) * B
    by_kernel = {}
    for i, (_, fl, _, _, kern) in enumerate(info):
        e = by_kernel.setdefault(kern + '_kernel', {'launches': 0, 'gflop': 0.0, 'us': 0.0})
        e['launches'] += 1
        e['gflop'] += fl * B / 1e9
        e['us'] += per_op[i + 1] * 1e3
    for e in by_kernel.values():
        e['tflops'] = e['gflop'] / e['us'] * 1e3 if e['us'] > 0 else 0.0
        e['frac'] = e['tflops'] / peak_burst
    # serialised diagnostic: the event between launches switches programmatic dependent launch off,
    # so these per-kernel times sum to MORE than the real step
    achieved_serial = conv_flops / (conv_ms * 1e-3) / 1e12
    n_conv = len(info)
    # the figure that can be checked from the line itself: algorithmic FLOPs of the step over the
    # timed step, with the kernels launched back to back exactly as in the timed region
    step_flops = GFLOP_PER_CLIP[args.model] * 1e9 * B
    achieved = step_flops / (ms / K * 1e-3) / 1e12
    traffic, traffic_src = ncu_traffic(args.model, args.dtype, B, world)

    if args.profile_table and rank == 0:
        with open(args.profile_table, 'w') as f:
            f.write('| # | convolution | us | GFLOP | TFLOP/s | min GB | GB/s |\n|---|---|---|---|---|---|---|\n')
            for i, (name, fl, ab, wb, kern) in enumerate(info):
                us = per_op[i + 1] * 1e3
                f.write('| %d | %s (%s) | %.1f | %.2f | %.0f | %.3f | %.0f |\n' % (
                    i, name, kern, us, fl * B / 1e9, fl * B / us / 1e6, (ab * B + wb) / 1e9,
                    (ab * B + wb) / us / 1e3))
            f.write('| | input staging | %.1f | | | | |\n| | pool + classifier | %.1f | | | | |\n'
                    % (per_op[0] * 1e3, per_op[-1] * 1e3))

    # ---- CPU baseline beside it (rank 0, N = 1 only): a bounded sample of the same workload
    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        clips, per_step = pick_cpu_sample(args, 8, 20.0)
        n = max(1, min(8, int(15.0 / max(per_step, 1e-3))))
        v, per_step, threads = time_oracle(args.model, clips, n, 0)
        cpu = {'value': v, 'unit': 'clips/s', 'cores': threads, 'kind': 'port',
               'sample': '%d steps of %d clips (of the batch-%d step), oracle/r2p1d_ref.py torch fp32'
                         % (n, clips, B)}

    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    if rank != 0:
        return
    line = {
        'metric': METRIC, 'value': value, 'unit': 'clips/s',
        'n_gpus': world, 'steps': K, 'warmup': W, 'ms_per_step': ms / K,
        'higher_is_better': True, 'scaling': 'weak', 'vs_baseline': None, 'dtype': args.dtype,
        'data': 'synthetic',
        'config': base_config(args, world),
        'e2e': {'value': value_e2e, 'unit': 'clips/s',
                'h2d_bytes_per_step': pipe.h2d_bytes, 'd2h_bytes_per_step': pipe.d2h_bytes},
        'gpu_launches': K * engine.num_launches,
        'clocks': clocks,
        'roofline': {'bound': 'tensor', 'achieved': achieved, 'peak': peak_burst,
                     'unit': 'TFLOP/s', 'frac': achieved / peak_burst,
                     'how': 'whole step: %.3f GFLOP/clip x %d clips / ms_per_step (the %d convolution '
                            'launches are %.1f%% of the step; kernels back to back with programmatic '
                            'dependent launch, as timed)' % (GFLOP_PER_CLIP[args.model], B, n_conv,
                                                             100 * conv_ms / (conv_ms + other_ms)),
                     'traffic': traffic, 'traffic_unit': 'bytes per convolution launch',
                     'traffic_source': traffic_src,
                     'algorithmic_bytes_per_launch': sum(a * B + w for _, _, a, w, _ in info) / n_conv,
                     'kernel': 'tcgen05 implicit-GEMM convolutions: conv_tc_kernel + conv_pair_kernel + conv_splitk_kernel + '
                               'conv_patch_kernel (%d launches per step)' % n_conv,
                     'peak_source': peak_src, 'frac_of_sustained': achieved / peak_sust,
                     'flops_per_launch_avg': conv_flops / n_conv,
                     'us_per_launch_avg': ms / K * 1e3 * conv_ms / (conv_ms + other_ms) / n_conv,
                     'kernel_share_of_step': conv_ms / (conv_ms + other_ms),
                     'serialised': {'note': 'CUDA event between launches = programmatic dependent '
                                            'launch OFF; sums exceed the real step; diagnostic only',
                                    'achieved': achieved_serial, 'frac': achieved_serial / peak_burst,
                                    'conv_us_sum': conv_ms * 1e3, 'other_us_sum': other_ms * 1e3,
                                    'by_kernel': by_kernel}},
        'e2e_plugin': plugin,
        'cpu_baseline': cpu,
        'latency': lat,
    }
    _emit(line)


def _emit(line):
    """The one JSON line, on the process's original stdout."""
    os.write(_REAL_STDOUT, (json.dumps(line) + '\n').encode())


# Libraries chat on stdout (NCCL prints its version there); the contract is ONE JSON line, so
# everything else is sent to stderr and the line is written to a saved copy of the descriptor.
_REAL_STDOUT = os.dup(1)
os.dup2(2, 1)


def main():
    args = parse_args()
    rank = int(os.environ.get('RANK', '0'))
    world = int(os.environ.get('WORLD_SIZE', '1'))
    local_rank = int(os.environ.get('LOCAL_RANK', '0'))
    if world == 1 and args.gpus > 1:
        # launched without torchrun: re-exec under it, as the driver would
        import subprocess
        port = 29500 + (os.getpid() % 2000)
        cmd = [sys.executable, '-m', 'torch.distributed.run', '--nnodes=1',
               '--nproc-per-node', str(args.gpus), '--master-addr', '127.0.0.1',
               '--master-port', str(port), os.path.abspath(__file__)] + sys.argv[1:]
        # (this process's fd 1 already points at stderr; hand the workers the real stdout)
        sys.exit(subprocess.call(cmd, stdout=_REAL_STDOUT))
    if args.impl == 'reference':
        run_reference(args, rank)
    else:
        run_b200(args, rank, world, local_rank)


if __name__ == '__main__':
    main()
