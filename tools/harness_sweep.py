#!/usr/bin/env python
"""Harness-level numbers: the pipeline files of config/ run under the reference's unmodified
benchmark.py (through scripts/run_rnb.py) on however many GPUs this box has, with the bulk client
for throughput and the Poisson client for latency.  One JSON object per run on stdout.

    python tools/harness_sweep.py [--videos 2000] [--gpus 1,2,4,8] [--only whole,aggressive,...]

Topologies (BASELINE.json configs 2-5):
  whole        config/r2p1d-whole.json, loader + runner on GPU 0
  aggressive   config/r2p1d-aggressive.json cut to the first N GPUs: one loader->runner pair and
               one queue per GPU
  nopipeline   config/r2p1d-nopipeline.json cut to N GPUs: R2P1DSingleStep replicas, one queue
  fourgroups   config/r2p1d-aggressive-four-groups.json (R(2+1)D-34) cut to N GPUs, 2 GPUs per group
  segment      config/r2p1d-segment.json (R(2+1)D-34): loader on GPU 0, num_segments 3, runners on
               GPUs 1..N-1, aggregator on the CPU
"""
import argparse
import glob
import json
import os
import subprocess
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, 'scripts'))
MEAN_CLIPS = (10 * 1 + 1 * 15) / 11.0   # sampler.py:58-59
BATCH = 3                               # config/rnb.json:15; --batch overrides


def load(name):
    return json.load(open(os.path.join(ROOT, 'config', name)))


def make_config(kind, n):
    if kind == 'whole':
        return load('r2p1d-whole.json')
    if kind == 'aggressive':
        c = load('r2p1d-aggressive.json')
        for st in c['pipeline']:
            st['queue_groups'] = st['queue_groups'][:n]
        return c
    if kind == 'nopipeline':
        c = load('r2p1d-nopipeline.json')
        c['pipeline'][0]['queue_groups'][0]['gpus'] = list(range(n))
        return c
    if kind == 'fourgroups':
        c = load('r2p1d-aggressive-four-groups.json')
        groups = max(1, n // 2)
        for st in c['pipeline']:
            st['queue_groups'] = st['queue_groups'][:groups]
            if n == 1:
                for g in st['queue_groups']:
                    g['gpus'] = [0]
        return c
    if kind == 'rnb':
        # config/rnb.json: loaders -> LargeSmallSelector -> Batcher (1-clip videos, `batch` of them
        # per forward) / pass-through (15-clip videos) -> runners
        c = load('rnb.json')
        gpus = list(range(n))
        c['pipeline'][0]['queue_groups'][0]['gpus'] = gpus
        c['pipeline'][1]['queue_groups'][0]['gpus'] = [0]
        c['pipeline'][1]['queue_groups'][0]['batch'] = BATCH
        c['pipeline'][1]['queue_groups'][1]['gpus'] = [1 if n > 1 else 0]
        c['pipeline'][2]['queue_groups'][0]['gpus'] = gpus
        return c
    if kind == 'segment':
        c = load('r2p1d-segment.json')
        c['pipeline'][1]['queue_groups'][0]['gpus'] = list(range(1, n)) if n > 1 else [0]
        return c
    raise ValueError(kind)


def run(kind, n, videos, mi, wire, max_clips=15, timeout=1200, runtime='reference'):
    import latency_percentiles
    cfg = make_config(kind, n)
    tmp = tempfile.mkdtemp(prefix='rnb_%s_%d_' % (kind, n))
    path = os.path.join(tmp, '%s-%dgpu.json' % (kind, n))
    json.dump(cfg, open(path, 'w'))
    env = dict(os.environ, CUDA_VISIBLE_DEVICES=','.join(str(i) for i in range(max(n, 1))),
               RNB_B200_ALLOW_BUSY_GPUS='1')
    cmd = [sys.executable, os.path.join(ROOT, 'scripts', 'run_rnb.py'), '--logdir', tmp, '--wire', wire,
           '--runtime', runtime, '--max-clips', str(max_clips), '-c', path, '-v', str(videos), '-mi', str(mi)]
    t0 = time.time()
    proc = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True,
                          timeout=timeout, env=env)
    rec = {'topology': kind, 'gpus': n, 'runtime': runtime, 'batcher_batch': BATCH if kind == 'rnb' else None, 'videos': videos, 'mean_interval_ms': mi, 'wire': wire,
           'client': 'bulk' if mi == 0 else 'poisson', 'wall_s': time.time() - t0,
           'rc': proc.returncode}
    jobs = glob.glob(os.path.join(tmp, 'logs', '*'))
    if proc.returncode == 0 and jobs:
        out = latency_percentiles.parse_job(jobs[0], skip=10)
        rec.update({'videos_per_s': out.get('videos_per_s'), 'seconds': out.get('seconds'),
                    'termination_flag': out.get('termination_flag'),
                    'latency_ms': out['latency_ms'], 'videos_logged': out['videos_logged']})
        if rec['videos_per_s']:
            rec['clips_per_s'] = rec['videos_per_s'] * MEAN_CLIPS
    else:
        rec['tail'] = proc.stdout[-1500:]
    print(json.dumps(rec), flush=True)
    return rec


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--videos', type=int, default=2000)
    ap.add_argument('--gpus', default='')
    ap.add_argument('--only', default='whole,aggressive,nopipeline,fourgroups,segment')
    ap.add_argument('--runtime', default='reference', choices=['reference', 'b200'])
    ap.add_argument('--batch', type=int, default=3, help="videos per forward of the 'rnb' topology's Batcher")
    ap.add_argument('--poisson', default='', help='comma-separated mean intervals (ms) for latency runs')
    args = ap.parse_args()
    global BATCH
    BATCH = args.batch
    import torch
    have = torch.cuda.device_count()
    counts = [int(v) for v in args.gpus.split(',')] if args.gpus else [c for c in (1, 2, 4, 8) if c <= have]
    kinds = args.only.split(',')
    for kind in kinds:
        for n in counts:
            if kind == 'whole' and n != 1:
                continue
            if kind == 'segment' and n < 2:
                continue
            if kind == 'fourgroups' and n not in (1, 2, 4, 8):
                continue
            for wire in (('fp32', 'uint8') if kind in ('whole', 'aggressive', 'segment') else ('fp32',)):
                rec = run(kind, n, args.videos, 0, wire, runtime=args.runtime)
                if args.poisson and rec.get('videos_per_s') and wire == 'uint8' or \
                        (args.poisson and kind in ('nopipeline', 'fourgroups') and rec.get('videos_per_s')):
                    for mi in [int(v) for v in args.poisson.split(',')]:
                        run(kind, n, min(args.videos, 1500), mi, wire, runtime=args.runtime)


if __name__ == '__main__':
    main()
