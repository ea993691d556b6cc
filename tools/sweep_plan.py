#!/usr/bin/env python
"""Greedy per-layer kernel sweep timed on the WHOLE step.

rnb_r2p1d_prepare times every candidate of a layer in isolation; with programmatic dependent launch
the kernels of a step overlap at their boundaries, so the isolated winner is not always the one
that makes the step fastest.  This tool starts from the prepared plan and, layer by layer, pins each
alternative (rnb_r2p1d_force_kernel), times the full forward and keeps a change only when the step
gets faster by more than the noise margin.  Prints what it kept; --save appends the resulting plan
to a plan-cache file that RNB_B200_PLAN_CACHE can point at.

    python tools/sweep_plan.py [--batch 16] [--model r18] [--iters 60] [--margin 0.004]
"""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

LAYER_SIZES = {'r18': [2, 2, 2, 2, 2], 'r34': [2, 3, 4, 6, 3]}

CANDIDATES = (
    [('conv_tc', dict(cm=cm, cn=cn)) for cm, cn in ((1, 1), (2, 1), (1, 2), (2, 2), (4, 1), (1, 4))] +
    [('conv_tc', dict(cm=1, cn=1, block_n=bn)) for bn in (64, 96, 128, 192)] +
    [('conv_patch', dict(pair=0)), ('conv_patch', dict(pair=1))] +
    [('conv_pair', dict(block_n=bn)) for bn in (256, 192, 128, 64)] +
    [('conv_tpatch', dict(block_n=bn)) for bn in (256, 128)] +
    [('conv_splitk', dict(ks=ks, block_n=bn)) for ks, bn in ((2, 256), (3, 256), (2, 128), (3, 128),
                                                              (4, 128), (4, 256))]
)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--batch', type=int, default=16)
    ap.add_argument('--model', default='r18')
    ap.add_argument('--iters', type=int, default=60)
    ap.add_argument('--margin', type=float, default=0.004)
    ap.add_argument('--ops', default='', help='comma-separated op indices (default: all)')
    args = ap.parse_args()
    import torch
    from rnb_b200 import _lib
    from rnb_b200.engine import R2P1DEngine, random_state_dict
    dev = torch.device('cuda:0')
    B = args.batch
    ls = LAYER_SIZES[args.model]
    eng = R2P1DEngine(dev, random_state_dict(1, 5, 400, ls, seed=0), 1, 5, 400, ls, max_batch=B)
    eng.prepare(B)
    g = torch.Generator(device=dev).manual_seed(1)
    pool = torch.randint(0, 256, (12, B, 8, 3, 112, 112), dtype=torch.uint8, device=dev, generator=g)

    def step_us():
        best = None
        for rep in range(3):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            for i in range(5):
                eng(pool[i % 12])
            e0.record()
            for i in range(args.iters):
                eng(pool[i % 12])
            e1.record()
            torch.cuda.synchronize()
            us = e0.elapsed_time(e1) / args.iters * 1e3
            best = us if best is None else min(best, us)
        return best

    names = [n for n, *_ in eng.op_info()]
    ops = [int(v) for v in args.ops.split(',')] if args.ops else range(len(names))
    base = step_us()
    print('prepared plan: %.1f us' % base, file=sys.stderr)
    kept = []
    for op in ops:
        cur_kind, cur_kw, _ = eng.plan_entry(B, op)
        best_us, best = base, (cur_kind, cur_kw)
        for kind, kw in CANDIDATES:
            try:
                eng.force_kernel(B, op, kind, **kw)
            except _lib.RnbError:
                continue
            got_kind, got_kw, _ = eng.plan_entry(B, op)
            if (got_kind, got_kw) == (cur_kind, cur_kw):
                continue
            us = step_us()
            if us < best_us * (1 - args.margin):
                best_us, best = us, (got_kind, got_kw)
        eng.force_kernel(B, op, best[0], **best[1])
        if best != (cur_kind, cur_kw):
            # confirm against a fresh baseline measurement before keeping it
            us = step_us()
            eng.force_kernel(B, op, cur_kind, **cur_kw)
            ref = step_us()
            if us < ref * (1 - args.margin):
                eng.force_kernel(B, op, best[0], **best[1])
                kept.append({'op': op, 'name': names[op], 'from': [cur_kind, cur_kw],
                             'to': list(best), 'us': us, 'ref_us': ref})
                base = us
                print('op %2d %-55s %s %s -> %s %s : %.1f -> %.1f us' % (
                    op, names[op], cur_kind, cur_kw, best[0], best[1], ref, us), file=sys.stderr)
            else:
                base = ref
    final = step_us()
    plan = [eng.plan_entry(B, op)[:2] for op in range(len(names))]
    print(json.dumps({'batch': B, 'model': args.model, 'final_us': final, 'kept': kept,
                      'plan': [[names[i]] + list(p) for i, p in enumerate(plan)]}))


if __name__ == '__main__':
    main()
