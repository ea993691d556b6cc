#!/usr/bin/env python
"""Per-kernel timeline of the real step (programmatic dependent launch intact) through the CUPTI
bridge: for every kernel of a forward pass its duration and its EXPOSED time, i.e. how much later
the chain ends because of it (end_i - max(end_{i-1}, start_i)); exposed times sum to the step.

    python tools/timeline.py [--batch 16] [--model r18] [--steps 30] > profiles/rNN_timeline.md
"""
import argparse
import os
import re
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
LAYER_SIZES = {'r18': [2, 2, 2, 2, 2], 'r34': [2, 3, 4, 6, 3]}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--batch', type=int, default=16)
    ap.add_argument('--model', default='r18')
    ap.add_argument('--steps', type=int, default=30)
    ap.add_argument('--raw', action='store_true',
                    help='launch list without layer names (chunked schedules: RNB_B200_CHUNK_MB)')
    args = ap.parse_args()
    import torch
    from rnb_b200 import cupti
    from rnb_b200.engine import R2P1DEngine, random_state_dict
    dev = torch.device('cuda:0')
    B, ls = args.batch, LAYER_SIZES[args.model]
    eng = R2P1DEngine(dev, random_state_dict(1, 5, 400, ls, seed=0), 1, 5, 400, ls, max_batch=B)
    eng.prepare(B)
    g = torch.Generator(device=dev).manual_seed(1)
    pool = torch.randint(0, 256, (12, B, 8, 3, 112, 112), dtype=torch.uint8, device=dev, generator=g)
    for i in range(10):
        eng(pool[i % 12])
    torch.cuda.synchronize()
    cupti.initialize()
    cupti.report()
    for i in range(args.steps):
        eng(pool[i % 12])
    torch.cuda.synchronize()
    cupti.flush()
    recs = [r for r in cupti.report(details=True) if 'rnb::' in r[0]]
    cupti.finalize()
    n = eng.num_launches
    assert len(recs) == n * args.steps, (len(recs), n, args.steps)
    names, flops = ['input staging'], [0.0]
    info = eng.op_info()
    k = 0
    if args.raw:
        names, flops, info = ['-'] * n, [0.0] * n, []
    while k < len(info):
        nm, f, _, _, kind = info[k]
        if kind in ('conv_patch_dual', 'conv_stem_fused') and k + 1 < len(info) \
                and info[k + 1][4] == kind and 'spatial' in nm:
            names.append(nm.replace('.spatial_conv', '') + ' (spatial + temporal, one launch)')
            flops.append((f + info[k + 1][1]) * B)
            k += 2
        else:
            names.append(nm)
            flops.append(f * B)
            k += 1
    if not args.raw:
        names.append('pool + classifier')
        flops.append(0.0)
    dur = [0.0] * n
    exp = [0.0] * n
    gap = [0.0] * n
    short = []
    for s in range(1, args.steps):          # skip the first step (no predecessor in the window)
        step = recs[s * n:(s + 1) * n]
        prev_end = recs[s * n - 1][2]
        for k, r in enumerate(step):
            dur[k] += r[2] - r[1]
            exp[k] += max(0, r[2] - max(prev_end, r[1])) if r[2] > prev_end else 0
            gap[k] += max(0, r[1] - prev_end)
            prev_end = max(prev_end, r[2])
        short.append(step[-1][2] - recs[s * n - 1][2])
    m = args.steps - 1
    kern = [re.sub(r'^void |rnb::|\(anonymous namespace\)::|\(.*$', '', recs[k][0]) for k in range(n)]
    print('# Kernel timeline of one step, R(2+1)D-%s bf16 batch %d (CUPTI, %d steps, PDL chain intact)\n'
          % (args.model[1:], B, m))
    print('step (end of classifier to end of next classifier): %.1f us mean\n' % (sum(short) / m / 1e3))
    print('| # | layer | kernel | grid | duration us | exposed us | idle before us | GFLOP | TFLOP/s over exposed |')
    print('|---|---|---|---|---|---|---|---|---|')
    for k in range(n):
        e = (exp[k] + gap[k]) / m / 1e3
        print('| %d | %s | `%s` | %d | %.1f | %.1f | %.1f | %.2f | %s |' % (
            k, names[k], kern[k][:40], recs[k][5][0], dur[k] / m / 1e3, exp[k] / m / 1e3,
            gap[k] / m / 1e3, flops[k] / 1e9, ('%.0f' % (flops[k] / e / 1e6)) if flops[k] and e > 0 else '-'))
    print('\nsum of exposed + idle: %.1f us' % ((sum(exp) + sum(gap)) / m / 1e3))


if __name__ == '__main__':
    main()
