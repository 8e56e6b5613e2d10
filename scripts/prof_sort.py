"""Per-kernel timing of one argsort (b200x_profile_*: CUDA events around each profiled launch)."""
import ctypes
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from ray_project_ray_b200 import _lib, ops

NAMES = {0: "pass", 1: "gather", 2: "finish", 3: "digit hist", 4: "permute", 5: "segreduce", 6: "lsd hist"}
n = int(sys.argv[1]) if len(sys.argv) > 1 else 100_000_000
kind = sys.argv[2] if len(sys.argv) > 2 else "i64"
dev = torch.device("cuda:0")
g = torch.Generator(device=dev).manual_seed(0)
if kind == "i64":
    k = torch.randint(0, 2**63 - 1, (n,), dtype=torch.int64, device=dev, generator=g)
elif kind == "f64":
    k = torch.randn(n, dtype=torch.float64, device=dev, generator=g)
else:
    raise SystemExit("kind: i64 | f64")
for _ in range(2):
    ops.argsort(k)
torch.cuda.synchronize()
dbg = {}
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
ops.argsort(k)
e1.record()
torch.cuda.synchronize()
print(f"argsort {kind} n={n}: {e0.elapsed_time(e1):.3f} ms")
ops.argsort(k, _debug=dbg)
print(dbg)
_lib.call("b200x_profile_enable", 1)
ops.argsort(k)
torch.cuda.synchronize()
for kid, name in NAMES.items():
    ms, cnt = ctypes.c_double(), ctypes.c_int()
    _lib.call("b200x_profile_read", kid, ctypes.byref(ms), ctypes.byref(cnt))
    if cnt.value:
        print(f"  {name:12s} {ms.value:8.3f} ms in {cnt.value} launches")
_lib.call("b200x_profile_enable", 0)
