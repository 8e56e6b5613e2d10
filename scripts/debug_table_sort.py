"""Developer aid: run sizes and bucket table of the table-binned hybrid sort (reads the scratch layout of csrc/radix_sort.cu)."""
import ctypes, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
from ray_project_ray_b200 import _lib, ops

n = int(sys.argv[1]) if len(sys.argv) > 1 else 100_000_000
dev = torch.device("cuda:0")
g = torch.Generator(device=dev).manual_seed(0)
k = torch.randn(n, dtype=torch.float64, device=dev, generator=g)
perm = torch.empty(n, dtype=torch.int32, device=dev)
need = _lib.query_temp("b200x_radix_argsort", k.data_ptr(), _lib.KEY_F64, 0, None, perm.data_ptr(), None, None, n, 0, 0,
                       stream=torch.cuda.current_stream().cuda_stream)
temp = torch.zeros(need, dtype=torch.uint8, device=dev)
nb = ctypes.c_size_t(need)
_lib.call("b200x_radix_argsort", k.data_ptr(), _lib.KEY_F64, 0, None, perm.data_ptr(), None, None, n, 0, 0, temp.data_ptr(),
          ctypes.byref(nb), torch.cuda.current_stream().cuda_stream)
torch.cuda.synchronize()
OFF_JOINT = 768 + 8192 + 2048
OFF_RUNS = OFF_JOINT + (512 + 3 * 65536) * 4
OFF_TAB = OFF_JOINT + (512 + 2 * 65536) * 4
runs = temp[OFF_RUNS:OFF_RUNS + 65537 * 4].cpu().numpy().view(np.uint32).astype(np.int64)
tab = temp[OFF_TAB:OFF_TAB + 65536 * 4].cpu().numpy().view(np.uint32)
sampU = temp[OFF_JOINT + 2048:OFF_JOINT + 2048 + 65536 * 4].cpu().numpy().view(np.uint32)
sampT = temp[OFF_JOINT + 2048 + 65536 * 4:OFF_JOINT + 2048 + 2 * 65536 * 4].cpu().numpy().view(np.uint32)
sizes = np.diff(runs)
print("runs: nonempty", (sizes > 0).sum(), "max", sizes.max(), "mean", sizes[sizes > 0].mean())
print("rows in runs >2048:", sizes[sizes > 2048].sum() / n, ">4096:", sizes[sizes > 4096].sum() / n)
bins, sb = tab >> 5, tab & 31
print("table: total bins", int(bins[-1]) + (1 << int(sb[-1]) if sb[-1] != 31 else 1), "dense buckets", (sb != 31).sum(),
      "s histogram", np.bincount(sb, minlength=32)[:17])
for r in np.argsort(sizes)[-6:]:
    j1 = np.searchsorted(bins, r, side="right") - 1
    print("run", r, "size", sizes[r], "bucket", j1, "s", sb[j1], "first bin", bins[j1], "sampled U", sampU[j1], "sampled T", sampT[r])
