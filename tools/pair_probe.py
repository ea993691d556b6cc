import sys, torch
sys.path.insert(0, '.')
from rnb_b200.engine import R2P1DEngine, random_state_dict
dev = torch.device('cuda:0'); B = 16
eng = R2P1DEngine(dev, random_state_dict(seed=0), max_batch=B).prepare(B)
g = torch.Generator(device=dev).manual_seed(1)
pool = torch.randint(0, 256, (12, B, 8, 3, 112, 112), dtype=torch.uint8, device=dev, generator=g)
def step_us(iters=100):
    best = None
    for rep in range(4):
        for i in range(5): eng(pool[i % 12])
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for i in range(iters): eng(pool[i % 12])
        e1.record(); torch.cuda.synchronize()
        us = e0.elapsed_time(e1) / iters * 1e3
        best = us if best is None else min(best, us)
    return best
info = eng.op_info()
print('base', step_us())
sp = [i for i, (n, *_r) in enumerate(info) if 'res2plus1d.conv2.' in n and 'spatial' in n]
tp = [i for i, (n, *_r) in enumerate(info) if 'res2plus1d.conv2.' in n and 'temporal' in n]
for i in sp: eng.force_kernel(B, i, 'conv_patch', pair=1)
print('conv2 spatial pair', step_us())
for i in tp: eng.force_kernel(B, i, 'conv_patch', pair=1)
print('conv2 spatial+temporal pair', step_us())
for i in sp: eng.force_kernel(B, i, 'conv_patch', pair=0)
print('conv2 temporal pair only', step_us())
for i in tp: eng.force_kernel(B, i, 'conv_patch', pair=0)
print('base again', step_us())
