"""groupby().aggregate(Sum, Mean, Count) over 200 M rows with 1 M distinct int64 keys (BASELINE configs[2]), uniform or
Zipf(1.1) keys, on one GPU: the dense-range path against the sorted path.
    python scripts/groupby_bench.py [rows] [uniform|zipf]"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

import ray_project_ray_b200.data as rd
from ray_project_ray_b200 import aggregate as A
from ray_project_ray_b200.aggregate import Count, Mean, Sum

n = int(sys.argv[1]) if len(sys.argv) > 1 else 200_000_000
kind = sys.argv[2] if len(sys.argv) > 2 else "uniform"
groups = 1_000_000
dev = torch.device("cuda", 0)
gen = torch.Generator(device=dev).manual_seed(4)
if kind == "uniform":
    k = torch.randint(0, groups, (n,), dtype=torch.int64, device=dev, generator=gen)
else:
    w = torch.arange(1, groups + 1, dtype=torch.float64, device=dev).pow(-1.1)
    cdf = torch.cumsum(w / w.sum(), 0)
    ranks = torch.searchsorted(cdf, torch.rand(n, dtype=torch.float64, device=dev, generator=gen)).clamp_(max=groups - 1)
    k = torch.randperm(groups, device=dev, generator=gen)[ranks]
    del w, cdf, ranks
v = torch.randint(-2**31, 2**31, (n,), dtype=torch.int64, device=dev, generator=gen)
ds = rd.from_torch_columns({"k": k, "v": v}, override_num_blocks=8)


def timed(fn, reps=3):
    fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(reps):
        r = fn()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / reps, r


ms, r = timed(lambda: ds.groupby("k").aggregate(Sum("v"), Mean("v"), Count()))
print(f"{kind}: dense-range path {ms:.3f} ms, {r.count()} groups")
saved, A.DENSE_MAX_RANGE = A.DENSE_MAX_RANGE, 0
ms2, r2 = timed(lambda: ds.groupby("k").aggregate(Sum("v"), Mean("v"), Count()))
A.DENSE_MAX_RANGE = saved
print(f"{kind}: sorted path      {ms2:.3f} ms")
a, b = r.to_numpy(), r2.to_numpy()
assert all((a[c] == b[c]).all() for c in a), "paths differ"
print("identical results")
