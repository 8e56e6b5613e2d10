"""Exoshuffle / gensort-shaped sort (BASELINE.json configs[3]): 100-byte records =
10-byte key + 90-byte payload, sorted by the key with `Dataset.sort("key")` on 1..8 GPUs.

    python scripts/gensort_bench.py --rows 125000000
    python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 \
        --master-port 29531 scripts/gensort_bench.py --rows 125000000      # 1e9 records = 100 GB

Records are synthetic (seeded per rank, generated on the device): key = 10 random bytes, payload =
8-byte little-endian record number + 82 filler bytes. Checked after the timed steps: global order of
the keys (within every rank and across rank boundaries), row count, and an order-independent checksum
of the record numbers. Prints one JSON line (rows/s, GB/s, fraction of the NVLink / HBM roofline of
SURVEY.md 8d: 2 x row bytes per row of HBM traffic, (G-1)/G x row bytes per row over NVLink).
"""
import argparse
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

import torch
import torch.distributed as dist

KEY_BYTES, PAYLOAD_BYTES = 10, 90


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--rows", type=int, default=125_000_000, help="records per GPU")
    ap.add_argument("--steps", type=int, default=4)
    ap.add_argument("--warmup", type=int, default=2)
    ap.add_argument("--records", type=int, default=1, help="1 = 128-byte record layout, 0 = two dense columns")
    a = ap.parse_args()
    import ray_project_ray_b200.data as rd
    from ray_project_ray_b200 import ops

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)
    n = a.rows
    g = torch.Generator(device=dev).manual_seed(77 + rank)
    key = torch.randint(0, 256, (n, KEY_BYTES), dtype=torch.uint8, device=dev, generator=g)
    payload = torch.randint(0, 256, (n, PAYLOAD_BYTES), dtype=torch.uint8, device=dev, generator=g)
    recno = torch.arange(rank * n, (rank + 1) * n, dtype=torch.int64, device=dev)
    payload[:, :8] = recno.view(torch.uint8).reshape(n, 8)
    want_sum = int(recno.sum().item())
    del recno
    ds = rd.from_torch_columns({"key": key, "payload": payload}, override_num_blocks=4, records=bool(a.records))
    del key, payload

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    out = None
    for _ in range(a.warmup):
        out = ds.sort("key")
    barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(a.steps):
        out = ds.sort("key")
    e1.record()
    barrier()
    ms = e0.elapsed_time(e1) / a.steps
    # ---- checks (outside the timed region)
    blocks = [b for b in out.get_internal_block_refs() if b.num_rows]
    rows = sum(b.num_rows for b in blocks)
    got_sum, first, last = 0, None, None
    prev_hi = prev_lo = None
    for b in blocks:
        k = b.columns["key"]
        hi = ops.pack_key_bytes(k, 0, 8).view(torch.int64) ^ (-0x8000000000000000)   # unsigned order as signed
        lo = ops.pack_key_bytes(k, 8, 2).view(torch.int16).to(torch.int32) & 0xFFFF
        ok = (hi[1:] > hi[:-1]) | ((hi[1:] == hi[:-1]) & (lo[1:] >= lo[:-1]))
        assert bool(ok.all()), "keys out of order inside a block"
        if prev_hi is not None:
            assert (int(hi[0]), int(lo[0])) >= (prev_hi, prev_lo), "keys out of order across blocks"
        prev_hi, prev_lo = int(hi[-1]), int(lo[-1])
        if first is None:
            first = (int(hi[0]), int(lo[0]))
        last = (prev_hi, prev_lo)
        got_sum += int(b.columns["payload"][:, :8].contiguous().view(torch.int64).sum().item())
    if world > 1:
        t = torch.tensor([ms], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t.item())
        tot = torch.tensor([rows, got_sum, want_sum], dtype=torch.int64, device=dev)
        dist.all_reduce(tot)
        rows, got_sum, want_sum = (int(x) for x in tot.tolist())
        edge = torch.tensor([first[0], first[1], last[0], last[1]] if first else [0, 0, 0, 0], dtype=torch.int64, device=dev)
        edges = [torch.empty_like(edge) for _ in range(world)]
        dist.all_gather(edges, edge)
        e = [tuple(int(v) for v in x.tolist()) for x in edges]
        for r in range(1, world):
            assert (e[r][0], e[r][1]) >= (e[r - 1][2], e[r - 1][3]), "keys out of order across ranks"
    assert rows == n * world, f"rows lost: {rows} != {n * world}"
    assert got_sum == want_sum, "record-number checksum differs"
    if rank == 0:
        row_bytes = KEY_BYTES + PAYLOAD_BYTES
        total = n * world
        peaks = {}
        try:
            with open(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "MEASURED_PEAKS.json")) as f:
                peaks = json.load(f)
        except OSError:
            pass
        hbm = float(peaks.get("hbm_gbs", 6548.8))
        t_hbm = 2 * row_bytes * n / (hbm * 1e9)
        t_nvl = row_bytes * n * (world - 1) / world / 900e9
        bound = "nvlink" if t_nvl > t_hbm else "hbm"
        print(json.dumps({
            "workload": f"gensort {row_bytes}-byte records, {KEY_BYTES}-byte key, {n} records per GPU",
            "n_gpus": world, "records": total, "gigabytes": total * row_bytes / 1e9, "ms_per_step": ms,
            "rows_per_s": total / (ms / 1e3), "gb_per_s": total * row_bytes / (ms / 1e3) / 1e9,
            "roofline": {"bound": bound, "roofline_ms": max(t_hbm, t_nvl) * 1e3,
                         "frac": max(t_hbm, t_nvl) * 1e3 / ms,
                         "peaks": f"HBM {hbm} GB/s measured, NVLink 900 GB/s per direction nominal"},
            "checks": "sorted within and across ranks, row count, record-number checksum"}))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
