"""Measure the write-only HBM ceiling this workload can be compared with (SURVEY.md §8d asks for it beside the copy
peak in MEASURED_PEAKS.json): fill 201 MB slots of a 604 MB ring (the engine's footprint at 4096 envs) and one big
buffer, with torch's fill kernel and with cudaMemsetAsync; CUDA events, best and median of 20."""
import json
import sys

import torch

def timed(fn, reps=20, inner=9):
    ts = []
    for _ in range(reps):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for k in range(inner):
            fn(k)
        b.record()
        torch.cuda.synchronize()
        ts.append(a.elapsed_time(b) / inner)
    ts.sort()
    return ts[0], ts[len(ts) // 2]

out = {}
slot = 4096 * 49152
ring = torch.empty((3, slot), dtype=torch.uint8, device='cuda')
big = torch.empty(4 << 30, dtype=torch.uint8, device='cuda')
src = torch.empty((3, slot), dtype=torch.uint8, device='cuda')
for name, fn, nbytes in (
        ('fill_ring_201MB', lambda k: ring[k % 3].fill_(7), slot),
        ('memset_ring_201MB', lambda k: ring[k % 3].zero_(), slot),
        ('fill_4GiB', lambda k: big.fill_(7), big.numel()),
        ('copy_ring_201MB(read+write)', lambda k: ring[k % 3].copy_(src[(k + 1) % 3]), 2 * slot)):
    for _ in range(3):
        fn(0)
    best, med = timed(fn, inner=9 if nbytes < (1 << 30) else 2)
    out[name] = {'best_GBs': nbytes / best / 1e6, 'median_GBs': nbytes / med / 1e6, 'ms_best': best}
print(json.dumps(out))
