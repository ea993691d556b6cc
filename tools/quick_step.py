#!/usr/bin/env python
"""Step time of the whole network at one batch size, for A/B runs of environment knobs:
    RNB_B200_L2_POLICY=residual python tools/quick_step.py [--batch 16] [--model r18] [--iters 100]
"""
import argparse, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
LAYER_SIZES = {'r18': [2, 2, 2, 2, 2], 'r34': [2, 3, 4, 6, 3]}
ap = argparse.ArgumentParser()
ap.add_argument('--batch', type=int, default=16)
ap.add_argument('--model', default='r18')
ap.add_argument('--iters', type=int, default=100)
args = ap.parse_args()
import torch
from rnb_b200.engine import R2P1DEngine, random_state_dict
dev = torch.device('cuda:0')
B, ls = args.batch, LAYER_SIZES[args.model]
eng = R2P1DEngine(dev, random_state_dict(1, 5, 400, ls, seed=0), 1, 5, 400, ls, max_batch=B).prepare(B)
g = torch.Generator(device=dev).manual_seed(1)
pool = torch.randint(0, 256, (12, B, 8, 3, 112, 112), dtype=torch.uint8, device=dev, generator=g)
for i in range(10):
    eng(pool[i % 12])
torch.cuda.synchronize()
best = None
for rep in range(5):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(args.iters):
        eng(pool[i % 12])
    e1.record()
    torch.cuda.synchronize()
    us = e0.elapsed_time(e1) / args.iters * 1e3
    best = us if best is None else min(best, us)
ents = [eng.plan_entry(B, i) for i in range(eng.num_convs)]
print('kernels:', ' '.join('%d:%s%s' % (i, e[0].replace('conv_', ''),
                                        ('(ks%d,n%d)' % (e[1]['ks'], e[1]['block_n'])) if e[0] == 'conv_splitk' else '')
                           for i, e in enumerate(ents)))
print('%s batch %d: %.1f us/step  %.0f clips/s  launches %d  env %s' % (
    args.model, B, best, B / best * 1e6, eng.num_launches,
    {k: v for k, v in os.environ.items() if k.startswith('RNB_B200')}))
