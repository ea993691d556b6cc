#!/usr/bin/env python
"""p50 / p99 per-video latency and throughput from the log directory of one RnB run.

The reference writes, per final-step process, `logs/<job>/g<gpu>-group<g>-<i>.txt`: a header line of
TimeCard keys (plus gpu columns) and one row of wall-clock stamps per video
(rnb_logging.py:188-214), and `log-meta.txt` with the start / end time of the run
(benchmark.py:297-300).  It has no percentile code.  Per-video latency here is the LAST stamp of a
row minus its `enqueue_filename` stamp.

    python scripts/latency_percentiles.py <logs/job-dir> [--skip 10] [--json]
"""
import argparse
import glob
import json
import os
import re


def parse_job(job_dir, skip=10):
    lat, clips_rows = [], 0
    keys_seen = None
    for path in sorted(glob.glob(os.path.join(job_dir, 'g*-group*-*.txt'))):
        with open(path) as f:
            lines = [ln.split() for ln in f.read().splitlines() if ln.strip()]
        if not lines:
            continue
        header = lines[0]
        keys = [k for k in header if not k.startswith('gpu')]
        keys_seen = keys
        i0 = keys.index('enqueue_filename')
        rows = [[float(v) for v in row[:len(keys)]] for row in lines[1:]]
        # (the reference's own summary skips the first NUM_SUMMARY_SKIPS rows of a process, runner.py:4)
        for row in rows[skip:]:
            lat.append((max(row) - row[i0]) * 1e3)
        clips_rows += len(rows)
    meta = {}
    meta_path = os.path.join(job_dir, 'log-meta.txt')
    if os.path.exists(meta_path):
        text = open(meta_path).read().splitlines()
        m = re.search(r'videos=(\d+)', text[0]) if text else None
        if m:
            meta['videos_requested'] = int(m.group(1))
        if len(text) > 1:
            t0, t1 = [float(v) for v in text[1].split()[:2]]
            meta['seconds'] = t1 - t0
        if len(text) > 2:
            meta['termination_flag'] = int(text[2].split(':')[1])
    lat.sort()

    def pct(q):
        return lat[min(len(lat) - 1, int(q * len(lat)))] if lat else None
    out = {'job': os.path.basename(os.path.normpath(job_dir)), 'videos_logged': clips_rows,
           'latency_ms': {'p50': pct(0.50), 'p90': pct(0.90), 'p99': pct(0.99),
                          'mean': sum(lat) / len(lat) if lat else None, 'n': len(lat)},
           'keys': keys_seen}
    out.update(meta)
    if meta.get('seconds') and meta.get('videos_requested'):
        out['videos_per_s'] = meta['videos_requested'] / meta['seconds']
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('job_dir')
    ap.add_argument('--skip', type=int, default=10)
    ap.add_argument('--clips-per-video', type=float, default=None,
                    help='mean clips per video of the run, to print clips/s as well')
    args = ap.parse_args()
    out = parse_job(args.job_dir, args.skip)
    if args.clips_per_video and out.get('videos_per_s'):
        out['clips_per_s'] = out['videos_per_s'] * args.clips_per_video
    print(json.dumps(out))


if __name__ == '__main__':
    main()
