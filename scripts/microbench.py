"""Developer micro-benchmarks of the individual kernels (CUDA events, L2-exceeding inputs)."""
import argparse
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from ray_project_ray_b200 import ops


def timeit(fn, iters=5, warmup=2):
    for _ in range(warmup):
        fn()
    torch.cuda.synchronize()
    ts = []
    for _ in range(iters):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        fn()
        e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    return min(ts), sum(ts) / len(ts)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--n", type=int, default=100_000_000)
    ap.add_argument("--only", default="")
    a = ap.parse_args()
    n = a.n
    dev = torch.device("cuda:0")
    g = torch.Generator(device=dev).manual_seed(0)
    res = {}

    def run(name, fn, bytes_):
        if a.only and a.only not in name:
            return
        best, avg = timeit(fn)
        print(f"{name:40s} best {best:8.3f} ms avg {avg:8.3f} ms  {bytes_ / best / 1e6:8.1f} GB/s (algorithmic)", flush=True)

    kfull = torch.randint(0, 2**63 - 1, (n,), dtype=torch.int64, device=dev, generator=g)
    kperm = torch.randperm(n, device=dev, generator=g)
    k20 = torch.randint(0, 1_000_000, (n,), dtype=torch.int64, device=dev, generator=g)
    run("argsort i64 full-range (8 passes)", lambda: ops.argsort(kfull), n * (8 + 8 * 24))
    run("argsort i64 permutation (4 passes)", lambda: ops.argsort(kperm), n * (8 + 4 * 24))
    run("argsort i64 1M distinct (3 passes)", lambda: ops.argsort(k20, want_codes=True), n * (8 + 3 * 24))
    heavy = torch.randint(0, 2**63 - 1, (2000,), dtype=torch.int64, device=dev, generator=g)[torch.randint(0, 2000, (n,), device=dev, generator=g)]
    run("argsort i64 2000 heavy keys, wide range (LSD after sampled warning)", lambda: ops.argsort(heavy), n * (8 + 8 * 24))
    del heavy
    mid = torch.randint(0, 2**63 - 1, (20000,), dtype=torch.int64, device=dev, generator=g)[torch.randint(0, 20000, (n,), device=dev, generator=g)]
    run("argsort i64 20000 keys x 5000 rows, wide range", lambda: ops.argsort(mid), n * (8 + 8 * 24))
    del mid
    f64 = torch.randn(n, dtype=torch.float64, device=dev, generator=g)
    run("argsort f64 normal", lambda: ops.argsort(f64), n * (8 + 8 * 24))
    del f64
    run("argsort i64 perm desc", lambda: ops.argsort(kperm, True), n * (8 + 4 * 24))
    perm, _, _ = ops.argsort(kperm)
    pay64 = torch.empty((n, 64), dtype=torch.uint8, device=dev)
    run("gather rows 64B", lambda: ops.gather(pay64, perm), n * 132)
    run("gather col 8B", lambda: ops.gather(kfull, perm), n * 20)
    run("permute_rows 64B", lambda: ops.permute_rows(pay64, 1), n * 128)
    run("permute_rows 8B", lambda: ops.permute_rows(kfull, 1), n * 16)
    del pay64
    pay72 = torch.empty((n, 72), dtype=torch.uint8, device=dev)
    run("gather rows 72B", lambda: ops.gather(pay72, perm), n * 148)
    run("permute_rows 72B", lambda: ops.permute_rows(pay72, 1), n * 144)
    del pay72
    cols = [torch.empty(n, dtype=torch.int64, device=dev) for _ in range(8)]
    run("gather 8 cols x 8B (one launch)", lambda: ops.gather_columns(cols, perm), n * (4 + 128))
    del cols
    # groupby pieces
    p20, c20, _ = ops.argsort(k20, want_codes=True)
    vals = torch.randint(-2**31, 2**31, (n,), dtype=torch.int64, device=dev, generator=g)
    run("group_boundaries", lambda: ops.group_boundaries(c20), n * 16)
    starts, ng = ops.group_boundaries(c20)
    G = int(ng.item())
    run("segmented_reduce i64 (perm gather)", lambda: ops.segmented_reduce(starts, ng, G, p20, vals, n), n * 12)
    fv = vals.to(torch.float64)
    run("segmented_reduce f64 (perm gather)", lambda: ops.segmented_reduce(starts, ng, G, p20, fv, n), n * 12)
    x = torch.empty(n, dtype=torch.int64, device=dev)
    run("torch copy 8B (reference bw)", lambda: x.copy_(kfull), n * 16)


if __name__ == "__main__":
    main()
