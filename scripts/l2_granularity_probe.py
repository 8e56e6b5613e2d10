import sys, os, ctypes
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from ray_project_ray_b200 import ops
dev = torch.device("cuda:0")
torch.zeros(1, device=dev)
rt = ctypes.CDLL("libcudart.so.12")
val = ctypes.c_size_t(0)
print("get rc", rt.cudaDeviceGetLimit(ctypes.byref(val), 5), "L2 fetch granularity", val.value)
n = 100_000_000
g = torch.Generator(device=dev).manual_seed(0)
perm = torch.randperm(n, device=dev, generator=g).to(torch.int32)
pay = torch.empty((n, 64), dtype=torch.uint8, device=dev)
def t():
    for _ in range(2): ops.gather(pay, perm)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(); ops.gather(pay, perm); e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1)
print("default", t())
for gran in (32, 64, 128):
    print("set", gran, "rc", rt.cudaDeviceSetLimit(5, ctypes.c_size_t(gran)))
    rt.cudaDeviceGetLimit(ctypes.byref(val), 5)
    print("  now", val.value, "gather ms", t())
