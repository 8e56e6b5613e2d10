"""Does the sort (instruction / latency bound) overlap with a row gather (DRAM bound) on a second stream?"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from ray_project_ray_b200 import ops
n = 100_000_000
dev = torch.device("cuda:0")
g = torch.Generator(device=dev).manual_seed(0)
k = torch.randint(0, 2**63 - 1, (n,), dtype=torch.int64, device=dev, generator=g)
pay = torch.empty((n, 64), dtype=torch.uint8, device=dev)
perm = torch.randperm(n, device=dev).to(torch.int32)
side = torch.cuda.Stream(device=dev)
def both():
    ev = torch.cuda.Event(); ev.record()
    with torch.cuda.stream(side):
        side.wait_event(ev)
        ops.gather(pay, perm)
    ops.argsort(k)
    torch.cuda.current_stream().wait_stream(side)
def timeit(fn, iters=5):
    for _ in range(2): fn()
    torch.cuda.synchronize(); ts = []
    for _ in range(iters):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); fn(); b.record(); torch.cuda.synchronize(); ts.append(a.elapsed_time(b))
    return min(ts)
ts = timeit(lambda: ops.argsort(k)); tg = timeit(lambda: ops.gather(pay, perm)); tb = timeit(both)
print(f"sort alone {ts:.3f} ms, gather alone {tg:.3f} ms, sum {ts + tg:.3f} ms, both on two streams {tb:.3f} ms")
