#!/usr/bin/env python
"""Two steps of the network between cudaProfilerStart/Stop, for an ncu pass that keeps the caches warm:
    ncu --profile-from-start off --cache-control none --clock-control none \
        --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum --csv \
        --log-file gpurun_out/dram.csv python tools/dram_probe.py
(per-kernel DRAM bytes with the L2 contents the previous launch left behind)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from rnb_b200.engine import R2P1DEngine, random_state_dict
dev = torch.device('cuda:0')
B, ls = 16, [2, 2, 2, 2, 2]
eng = R2P1DEngine(dev, random_state_dict(1, 5, 400, ls, seed=0), 1, 5, 400, ls, max_batch=B).prepare(B)
g = torch.Generator(device=dev).manual_seed(1)
pool = torch.randint(0, 256, (4, B, 8, 3, 112, 112), dtype=torch.uint8, device=dev, generator=g)
for i in range(4):
    eng(pool[i])
torch.cuda.synchronize()
torch.cuda.profiler.start()
for i in range(2):
    eng(pool[i])
torch.cuda.synchronize()
torch.cuda.profiler.stop()
