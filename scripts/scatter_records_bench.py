"""The sender side of the record exchange in isolation, on ONE GPU (the "peers" are local buffers, so this measures the
kernel against HBM, not NVLink): b200x_partition_targets + b200x_scatter_records_by_target for gensort records.
    python scripts/scatter_records_bench.py [rows] [world]"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch

from ray_project_ray_b200 import _lib, ops

n = int(sys.argv[1]) if len(sys.argv) > 1 else 125_000_000
world = int(sys.argv[2]) if len(sys.argv) > 2 else 8
dev = torch.device("cuda", 0)
g = torch.Generator(device=dev).manual_seed(3)
rec = torch.randint(0, 256, (n, 128), dtype=torch.uint8, device=dev, generator=g)
hi = torch.randint(-2**63, 2**63 - 1, (n,), dtype=torch.int64, device=dev, generator=g).view(torch.uint64)
split = torch.from_numpy((np.arange(1, world, dtype=np.uint64) * np.uint64(2**64 // world)).view(np.int64)).to(dev)


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


print("partition_targets (one pass)        ", round(timed(lambda: ops.partition_targets(hi, False, split)), 3), "ms")
target, counts = ops.partition_targets(hi, False, split)
cnt = counts.cpu().tolist()[:world]
kb = [torch.empty(c, dtype=torch.int64, device=dev) for c in cnt]
print("key column scatter (8-byte rows)    ", round(timed(lambda: ops.scatter_rows_by_target(hi, target, [b.data_ptr() for b in kb], [0] * world)), 3), "ms")
for pitch, label in ((96, "dense 96-byte slots"), (128, "128-byte slots     ")):
    bufs = [torch.empty((c, pitch), dtype=torch.uint8, device=dev) for c in cnt]
    for bulk in (1, 0):
        if bulk and pitch != 96:
            continue
        _lib.call("b200x_set_tuning", 4, bulk)
        for bps in (0, 2):
            _lib.call("b200x_set_tuning", 1, bps)
            ms = timed(lambda: ops.scatter_records_by_target(rec, target, [b.data_ptr() for b in bufs], [0] * world, pitch,
                                                            copy_bytes=96, src_offset=8))
            print(f"records, 96 bytes into {label}, {'bulk copies' if bulk else '16-byte stores'}, "
                  f"{'one block per tile' if bps == 0 else '2 blocks per SM   '}: {ms:.3f} ms "
                  f"({n * (96 + 96 + 4) / ms / 1e6:.0f} GB/s of 4 + 96 + 96 bytes per record)")
    del bufs
_lib.call("b200x_set_tuning", 1, 0)
_lib.call("b200x_set_tuning", 4, 1)
