"""Kernel timeline of one Dataset.sort step on gensort records (torch.profiler / CUPTI): prints every kernel with its
start time and the idle gap before it, to find where the device waits for the host.
    python scripts/timeline.py [rows] [prefix 0|1]"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from torch.profiler import ProfilerActivity, profile

import ray_project_ray_b200.data as rd

n = int(sys.argv[1]) if len(sys.argv) > 1 else 125_000_000
prefix = ["key"] if (len(sys.argv) < 3 or sys.argv[2] == "1") else None
dev = torch.device("cuda", 0)
g = torch.Generator(device=dev).manual_seed(1)
key = torch.randint(0, 256, (n, 10), dtype=torch.uint8, device=dev, generator=g)
pay = torch.randint(0, 256, (n, 90), dtype=torch.uint8, device=dev, generator=g)
ds = rd.from_torch_columns({"key": key, "payload": pay}, override_num_blocks=8, records=True, key_prefix=prefix)
del key, pay
for _ in range(3):
    out = ds.sort("key")
torch.cuda.synchronize()
with profile(activities=[ProfilerActivity.CUDA, ProfilerActivity.CPU]) as prof:
    for _ in range(2):
        out = ds.sort("key")
    torch.cuda.synchronize()
ev = [e for e in prof.events() if e.device_type == torch.autograd.DeviceType.CUDA]
ev.sort(key=lambda e: e.time_range.start)
t0 = ev[0].time_range.start
prev_end = t0
for e in ev:
    gap = e.time_range.start - prev_end
    dur = e.time_range.end - e.time_range.start
    if dur > 20 or gap > 20:
        print(f"{(e.time_range.start - t0) / 1e3:9.3f} ms  gap {gap / 1e3:7.3f}  dur {dur / 1e3:7.3f}  {e.name[:70]}")
    prev_end = max(prev_end, e.time_range.end)
print("total", (prev_end - t0) / 1e3, "ms for 2 steps")
