"""Staged scatter by target on ONE GPU (the 'peers' are local buffers): cost per row width."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from ray_project_ray_b200 import ops
n = int(sys.argv[1]) if len(sys.argv) > 1 else 100_000_000
world = 8
dev = torch.device("cuda:0")
g = torch.Generator(device=dev).manual_seed(0)
keys = torch.randint(0, 2**63 - 1, (n,), dtype=torch.int64, device=dev, generator=g)
split = torch.from_numpy((np.arange(1, world, dtype=np.uint64) * np.uint64(2**63 // world) + np.uint64(2**63)).view(np.int64)).to(dev)
target, counts = ops.partition_targets(keys, False, split)
cnt = counts.cpu().tolist()[:world]
def timeit(fn, iters=4):
    for _ in range(2): fn()
    torch.cuda.synchronize(); ts = []
    for _ in range(iters):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); fn(); b.record(); torch.cuda.synchronize(); ts.append(a.elapsed_time(b))
    return min(ts)
for w in (64, 96, 90, 100, 82, 128):
    pay = torch.empty((n, w), dtype=torch.uint8, device=dev)
    bufs = [torch.empty((c + 1, w), dtype=torch.uint8, device=dev) for c in cnt]
    t = timeit(lambda: ops.scatter_rows_by_target(pay, target, [b.data_ptr() for b in bufs], [0] * world))
    print(f"scatter rows of {w:3d} B to {world} local buffers: {t:7.3f} ms  {2 * n * w / t / 1e6:7.0f} GB/s", flush=True)
    del pay, bufs
