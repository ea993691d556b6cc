#!/usr/bin/env python
"""Achieved bandwidth of the hand-off copy (rnb_peer_copy_fenced) between two GPUs, for the
message sizes of BASELINE config 4 (one 5-clip segment as uint8 / bf16 / fp32 = 1.5 / 3 / 6 MB)
and one 16-clip conv2 activation in the bf16 wire format (51 MB).  Prints a markdown table.

    python tools/nvlink_bw.py            # needs >= 2 GPUs; with 1 GPU it times the same-device copy
"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    import torch
    from rnb_b200 import ops
    n = torch.cuda.device_count()
    src_dev = torch.device('cuda:0')
    dst_dev = torch.device('cuda:%d' % (1 if n > 1 else 0))
    sizes = [('5-clip segment, uint8', 5 * 8 * 3 * 112 * 112),
             ('5-clip segment, bf16', 5 * 8 * 3 * 112 * 112 * 2),
             ('5-clip segment, fp32', 5 * 8 * 3 * 112 * 112 * 4),
             ('15-clip video, fp32', 15 * 8 * 3 * 112 * 112 * 4),
             ('16-clip conv2 activation, bf16', 16 * 8 * 56 * 56 * 64 * 2),
             ('256 MB', 256 << 20)]
    print('| message | bytes | us per copy | GB/s | path |')
    print('|---|---|---|---|---|')
    path = 'cuda:0 -> cuda:1 (peer access enabled)' if n > 1 else 'same device'
    for name, nbytes in sizes:
        src = torch.empty(nbytes, dtype=torch.uint8, device=src_dev)
        dst = torch.empty(nbytes, dtype=torch.uint8, device=dst_dev)
        with torch.cuda.device(dst_dev):
            for _ in range(5):
                ops.peer_copy(dst, src)
            torch.cuda.synchronize(dst_dev)
            iters = 50
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(iters):
                ops.peer_copy(dst, src)
            e1.record()
            torch.cuda.synchronize(dst_dev)
            us = e0.elapsed_time(e1) / iters * 1e3
        print('| %s | %d | %.1f | %.1f | %s |' % (name, nbytes, us, nbytes / us / 1e3, path))


if __name__ == '__main__':
    main()
