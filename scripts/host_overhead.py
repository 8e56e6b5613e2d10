import sys, os, time, cProfile, pstats
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import ray_project_ray_b200.data as rd
n = 100_000_000
dev = torch.device("cuda:0")
g = torch.Generator(device=dev).manual_seed(0)
k = torch.randint(0, 2**63 - 1, (n,), dtype=torch.int64, device=dev, generator=g)
p = torch.empty((n, 64), dtype=torch.uint8, device=dev)
ds = rd.from_torch_columns({"id": k, "payload": p}, override_num_blocks=8)
for _ in range(3):
    ds.sort("id")
torch.cuda.synchronize()
t0 = time.perf_counter()
for _ in range(10):
    out = ds.sort("id")
torch.cuda.synchronize()
print("wall per step ms", (time.perf_counter() - t0) / 10 * 1e3)
pr = cProfile.Profile()
pr.enable()
for _ in range(10):
    out = ds.sort("id")
torch.cuda.synchronize()
pr.disable()
pstats.Stats(pr).sort_stats("cumulative").print_stats(28)
