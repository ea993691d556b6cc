#!/usr/bin/env python
"""Where the step time goes, measured with the programmatic-dependent-launch chain intact.

bench.py's per-kernel table records an event between launches, which serialises the chain.  Here
each layer range 1..k (k = 1..5) is its own engine and is timed as a whole, back to back, so the
difference between consecutive ranges is the cost a stage adds to the real step.  Partial ranges
hand back the private bf16 channels-last format (a plain device copy whose cost is measured
separately and subtracted).

    python tools/stage_times.py [--batch 16] [--model r18] [--iters 100]
"""
import argparse
import ctypes
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

LAYER_SIZES = {'r18': [2, 2, 2, 2, 2], 'r34': [2, 3, 4, 6, 3]}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--batch', type=int, default=16)
    ap.add_argument('--model', default='r18')
    ap.add_argument('--iters', type=int, default=100)
    args = ap.parse_args()
    import torch
    from rnb_b200 import _lib
    from rnb_b200.engine import R2P1DEngine, random_state_dict, LAYER_OUTPUT_SHAPE
    dev = torch.device('cuda:0')
    B = args.batch
    ls = LAYER_SIZES[args.model]
    g = torch.Generator(device=dev).manual_seed(1)
    pool = torch.randint(0, 256, (12, B, 8, 3, 112, 112), dtype=torch.uint8, device=dev, generator=g)
    rows = []
    prev = 0.0

    def timed(fn):
        best = None
        for rep in range(3):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for i in range(args.iters):
                fn(i)
            e1.record()
            torch.cuda.synchronize()
            ms = e0.elapsed_time(e1) / args.iters
            best = ms if best is None else min(best, ms)
        return best

    for end in range(1, 6):
        sd = random_state_dict(1, end, 400, ls[:end], seed=0)
        eng = R2P1DEngine(dev, sd, 1, end, 400, ls[:end], dtype='bf16', max_batch=B)
        eng.prepare(B)
        if end < 5:
            c, t, h, w = LAYER_OUTPUT_SHAPE[end]
            out = torch.empty((B, t, h, w, (c + 15) // 16 * 16), dtype=torch.bfloat16, device=dev)
            fmt = _lib.FMT_BF16_NDHWC
        else:
            out = torch.empty((B, 400), dtype=torch.float32, device=dev)
            fmt = _lib.FMT_F32_LOGITS
        stream = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)

        def fwd(x):
            _lib.check(eng.lib.rnb_r2p1d_forward(eng._handle, ctypes.c_void_p(x.data_ptr()),
                                                 _lib.FMT_U8_NTCHW, B, ctypes.c_void_p(out.data_ptr()),
                                                 fmt, stream))
        for i in range(10):
            fwd(pool[i % 12])
        torch.cuda.synchronize()
        best = timed(lambda i: fwd(pool[i % 12]))
        if end < 5:
            src = torch.empty_like(out)
            best -= timed(lambda i: out.copy_(src))
        rows.append({'range': '1..%d' % end, 'us': best * 1e3, 'delta_us': (best - prev) * 1e3,
                     'gflop': eng.flops(B) / 1e9, 'launches': eng.num_launches,
                     'kernels': [k for *_, k in eng.op_info()]})
        prev = best
        eng.close()
    print(json.dumps({'batch': B, 'model': args.model, 'stages': rows}))
    for r in rows:
        print('%-6s %8.1f us  (+%7.1f)  %7.1f GFLOP  %d launches' % (
            r['range'], r['us'], r['delta_us'], r['gflop'], r['launches']), file=sys.stderr)


if __name__ == '__main__':
    main()
