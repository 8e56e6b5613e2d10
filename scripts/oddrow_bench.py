"""Row movers on gensort-shaped columns (10-byte key, 90-byte payload) against 16-byte-multiple rows."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from ray_project_ray_b200 import ops
n = int(sys.argv[1]) if len(sys.argv) > 1 else 100_000_000
dev = torch.device("cuda:0")
def timeit(fn, iters=4):
    for _ in range(2): fn()
    torch.cuda.synchronize(); ts = []
    for _ in range(iters):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); fn(); b.record(); torch.cuda.synchronize(); ts.append(a.elapsed_time(b))
    return min(ts)
perm = torch.randperm(n, device=dev).to(torch.int32)
for w in (90, 96, 100, 112, 128, 10, 16, 82, 80):
    x = torch.empty((n, w), dtype=torch.uint8, device=dev)
    t = timeit(lambda: ops.gather(x, perm))
    print(f"gather rows of {w:3d} B: {t:7.3f} ms  {n * 2 * w / t / 1e6:7.0f} GB/s useful", flush=True)
    del x
