"""Receiver-side gather of the record exchange in isolation (one GPU): b200x_gather_rows_pitched against
b200x_gather_records_prefixed as one pass and as K passes restricted to slot ranges (the chunked schedule).
    python scripts/gather_bench.py [rows]"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from ray_project_ray_b200 import ops

n = int(sys.argv[1]) if len(sys.argv) > 1 else 125_000_000
dev = torch.device("cuda", 0)
recv = torch.randint(0, 256, (n, 128), dtype=torch.uint8, device=dev)
perm = torch.randperm(n, device=dev, dtype=torch.int32) if n < 2**31 else None
keys = torch.sort(torch.randint(0, 2**62, (n,), dtype=torch.int64, device=dev)).values.view(torch.uint64)
out = torch.empty_like(recv)


def timed(fn, reps=3):
    fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(reps):
        fn()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / reps


print("gather_rows_pitched 128 B            ", round(timed(lambda: ops.gather_records(recv, perm)), 3), "ms")
print("prefixed, shift 0, one pass          ", round(timed(lambda: ops.gather_records_prefixed(recv, perm, out, 128, 128)), 3), "ms")
print("prefixed, shift 8 copy 96, one pass  ", round(timed(lambda: ops.gather_records_prefixed(recv, perm, out, 96, 100, keys, True, 8)), 3), "ms")
from ray_project_ray_b200 import _lib
_lib.call("b200x_set_tuning", 2, 4)
print("  the same with 4 loads in flight     ", round(timed(lambda: ops.gather_records_prefixed(recv, perm, out, 96, 100, keys, True, 8)), 3), "ms")
_lib.call("b200x_set_tuning", 2, 2)
print("  + tie_mode 1                       ", round(timed(lambda: ops.gather_records_prefixed(recv, perm, out, 96, 100, keys, True, 8, tie_mode=1)), 3), "ms")
print("  tie_mode 2 only                    ", round(timed(lambda: ops.gather_records_prefixed(recv, perm, out, 96, 100, keys, True, 8, tie_mode=2)), 3), "ms")
for K in (2, 4, 8):
    W = 8
    per = n // W
    starts = [r * per + c * per // K for r in range(W) for c in range(K)]
    sizes = [sum((r * per + (c + 1) * per // K if c + 1 < K else (r + 1) * per if r + 1 < W else n) - starts[r * K + c]
                 for r in range(W)) for c in range(K)]
    print(f"  split positions into {K} chunk lists   ", round(timed(lambda: ops.split_positions_by_chunk(perm, starts, W, K, sizes)), 3), "ms")
    lists = ops.split_positions_by_chunk(perm, starts, W, K, sizes)

    def passes():
        for c in range(K):
            ops.gather_records_prefixed(recv, perm, out, 96, 100, keys, True, 8, positions=lists[c])
    print(f"  {K} list passes                        ", round(timed(passes), 3), "ms")
