import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from ray_project_ray_b200 import ops
n = int(sys.argv[1]) if len(sys.argv) > 1 else 50_000_000
dev = torch.device("cuda:0")
g = torch.Generator(device=dev).manual_seed(0)
k = torch.randint(0, 2**63 - 1, (n,), dtype=torch.int64, device=dev, generator=g)
for _ in range(2):
    p, c, _ = ops.argsort(k)
torch.cuda.synchronize()
print("ok", int(p[0].item()))
