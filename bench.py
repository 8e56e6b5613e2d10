#!/usr/bin/env python
"""Headline benchmark: Dataset.sort throughput on B200 (BASELINE.json `metric`), on the workload
`north_star` names: synthetic gensort-shaped records.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...

Workload (config.workload = "gensort_125Mx100B"): per GPU 125 M records of 100 bytes = a 10-byte key
(`key`, [n, 10] uint8, random bytes) + a 90-byte payload (`payload`, [n, 90] uint8: 8-byte record number,
8-byte tag derived from key and record number, 74 filler bytes) - BASELINE.json configs[3], the reference's
own sort benchmark uses 100-byte rows "to approximately match Cloudsort"
(release/nightly_tests/dataset/sort.py:96-102). The dataset is held in the device block format's record
layout (one 128-byte record per row) and sorted by `Dataset.sort("key")`: pyarrow's lexicographic byte
order, stable. One row in 1024 repeats its predecessor's key and one in 4096 takes its key from a pool
shared by all ranks, so that the order of equal keys (within and across ranks) is checked on the timed
data. A step = one full sort of the resident dataset. N > 1: every rank holds 125 M records (weak scaling,
100 GB at N = 8) and the sort is global (range-partitioned exchange over NVLink peer memory).

Printed JSON (rank 0, one line): see DESIGN.md §Measurement for every key.
  value      rows/s, inputs resident in HBM, CUDA events around K steps, max over ranks
  e2e        same metric through the public API from pinned HOST buffers (H2D + sort + D2H per step)
  roofline   N = 1: the kernel with the largest share of the step (CUDA events on its launching stream) against
             the measured HBM peak; N > 1: the step against the slower of the HBM / NVLink rooflines of SURVEY.md 8d
  verified   checks of the output of the last timed step, at every N (outside the timed region)
  cpu_baseline  the reference's CPU path (port calling the same pyarrow kernels) on the host cores
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

ROWS_PER_GPU = 125_000_000
KEY_BYTES, PAYLOAD_BYTES = 10, 90
ROW_BYTES = KEY_BYTES + PAYLOAD_BYTES
RECORD_PITCH = 128
CPU_SAMPLE_ROWS = 25_000_000          # rows the CPU arm sorts per step (a bounded sample of one GPU's 125 M)
NVLINK_GBS = 900.0                    # nominal NVLink 5 bandwidth per direction and GPU
TAG_MULT = 0x9E3779B97F4A7C15 - (1 << 64)   # odd 64-bit constant (as int64): tag = key_hi ^ (recno * TAG_MULT)
# Algorithmic bytes per row of one launch of each timed kernel class, and the DRAM traffic
# per row ncu measured for it (profiles/, dram__bytes_read.sum + dram__bytes_write.sum; None = not captured).
KERNELS = {
    0: ("onesweep_pass_kernel", 22.0, 26.0),         # HYBRID: pass A reads 8 writes 12, pass B reads 12 writes 12
    1: ("record gather (gather_rows_pitched_kernel / gather_records_list_kernel, 128-byte records)", 4 + 2 * ROW_BYTES, 259.7),  # index + read record + write record; ncu: 16.5 GB read + 16.0 GB written per 125 M records (2 x the 128-byte record)
    2: ("finish_runs_kernel", 12 + 4 + 8, 23.5),      # read pair, write row index + decoded key
    3: ("digit2_hist_kernel", 8.0, 8.0),              # read key column
    6: ("hist_kernel", 8.0, None),
    7: ("peer scatter (scatter_rows_staged_kernel for the keys, scatter_records_staged_kernel for the records)", None, None),
}


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--rows", type=int, default=ROWS_PER_GPU, help="records per GPU")
    ap.add_argument("--cpu-rows", type=int, default=CPU_SAMPLE_ROWS)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-extras", action="store_true", help="skip the side measurements (72-byte sort, random_shuffle, groupby)")
    return ap.parse_args()


class ClockSampler:
    """Samples nvidia-smi clocks / throttle reasons while the timed region runs."""

    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index: int):
        self.gpu, self.proc, self.lines = gpu_index, None, []

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-lms", "100", "-i", str(self.gpu)], stdout=subprocess.PIPE,
                                         stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except subprocess.TimeoutExpired:
            self.proc.kill()
        sm, smax, reasons, power = [], [], set(), []
        for ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 9:
                continue
            try:
                sm.append(float(f[1]))
                smax.append(float(f[2]))
                power.append(float(f[3]))
            except ValueError:
                continue
            for name, val in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), f[5:9]):
                if val.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(smax) if smax else None,
                "power_w_max": max(power) if power else None, "samples": len(sm), "reasons": sorted(reasons)}


def workload_config(rows_per_gpu: int, world: int, nblocks: int):
    return {"workload": "gensort_125Mx100B", "rows_per_gpu": rows_per_gpu, "row_bytes": ROW_BYTES,
            "key": "10 random bytes ([n, 10] uint8), lexicographic, stable", "key_bytes": KEY_BYTES,
            "payload_bytes": PAYLOAD_BYTES, "record_pitch": RECORD_PITCH, "blocks_per_gpu": nblocks,
            "layout": "one 128-byte record per row (key + payload) + the key's first 8 bytes as a dense big-endian uint64 "
                      "key prefix column (part of the block format: built at ingest, which e2e times; written by every sort)",
            "planted_ties": "1/1024 rows repeat the previous key, 1/4096 rows draw from a 65536-key pool shared by all ranks",
            "l2": f"inputs ({rows_per_gpu * RECORD_PITCH / 1e9:.1f} GB per GPU) exceed the 126 MB L2; no flush needed",
            "parallelism": f"range-partitioned x{world}" if world > 1 else "single GPU"}


def cpu_arm(rows: int, steps: int, warmup: int):
    """The reference's CPU path (oracle port) on the host cores: seconds per sort of `rows` gensort records."""
    from oracle import cpu_baseline as cb

    threads = cb.host_cores()
    blocks = cb.make_gensort_blocks(rows, max(threads, 8))
    times = []
    for i in range(warmup + steps):
        t0 = time.perf_counter()
        outs = cb.sort_blocks(blocks, ["hi", "lo"], False, threads)
        dt = time.perf_counter() - t0
        assert sum(o.num_rows for o in outs) == rows
        if i >= warmup:
            times.append(dt)
    return sum(times) / len(times), threads, len(blocks)


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    import pyarrow

    rows = args.cpu_rows
    sec, threads, nblocks = cpu_arm(rows, args.steps, args.warmup)
    value = rows / sec
    sample = (f"{rows} gensort records x {ROW_BYTES} B per step (key as (hi uint64, lo uint16), sort(['hi','lo']), "
              f"{nblocks} blocks), pull-based map/reduce on a {threads}-thread pool, pyarrow {pyarrow.__version__}; "
              f"rows/s is size-normalised (the GPU arm sorts {args.rows} records per GPU and step)")
    line = {
        "impl": "reference", "metric": "Dataset.sort rows/s", "value": value, "unit": "rows/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": sec * 1e3, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "u8", "data": "synthetic",
        "gb_per_s": value * 2 * ROW_BYTES / 1e9,
        "config": workload_config(args.rows, max(args.gpus, 1), 8),      # the GPU arm's config, key for key
        "cpu_baseline": {"value": value, "unit": "rows/s", "cores": threads, "kind": "port", "sample": sample},
        "e2e": {"value": value, "unit": "rows/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    print(json.dumps(line), flush=True)


def make_records(n: int, rank: int, dev):
    """This rank's synthetic records on the device: (key [n, 10] uint8, payload [n, 90] uint8)."""
    import torch

    from ray_project_ray_b200 import ops

    g = torch.Generator(device=dev).manual_seed(77 + rank)
    key = torch.randint(0, 256, (n, KEY_BYTES), dtype=torch.uint8, device=dev, generator=g)
    payload = torch.randint(0, 256, (n, PAYLOAD_BYTES), dtype=torch.uint8, device=dev, generator=g)
    pool = torch.randint(0, 256, (65536, KEY_BYTES), dtype=torch.uint8, device=dev,
                         generator=torch.Generator(device=dev).manual_seed(4242))      # the same on every rank
    i = torch.arange(5, n, 4096, device=dev)
    key[i] = pool[(((i // 4096 + 7919 * rank) * 2654435761) % (1 << 32)) >> 16]     # ~0.5 rows per pool key and rank
    i = torch.arange(1027, n, 1024, device=dev)
    key[i] = key[i - 1]
    del i, pool
    recno = torch.arange(rank * n, (rank + 1) * n, dtype=torch.int64, device=dev)
    hi = ops.pack_key_bytes(key, 0, 8).view(torch.int64)
    payload[:, 0:8] = recno.view(torch.uint8).reshape(n, 8)
    payload[:, 8:16] = (hi ^ (recno * TAG_MULT)).view(torch.uint8).reshape(n, 8)
    return key, payload


def checksums(blocks, dev):
    """Order-independent sums over a list of blocks: rows, sum/xor of the key's first 8 bytes, sum of its last 2,
    sum/xor of the record numbers (int64 wrap-around)."""
    import torch

    from ray_project_ray_b200 import ops
    from ray_project_ray_b200.block import _contig

    acc = torch.zeros(6, dtype=torch.int64, device=dev)
    for b in blocks:
        if b.num_rows == 0:
            continue
        hi = ops.pack_key_bytes(b.columns["key"], 0, 8).view(torch.int64)
        lo = ops.pack_key_bytes(b.columns["key"], 8, 2).view(torch.int16).to(torch.int64) & 0xFFFF
        rec = _contig(b.columns["payload"][:, 0:8]).view(torch.int64).reshape(-1)
        acc[0] += b.num_rows
        acc[1] += hi.sum()
        acc[2] ^= _xor_reduce(hi)
        acc[3] += lo.sum()
        acc[4] += rec.sum()
        acc[5] ^= _xor_reduce(rec)
    return acc


def _xor_reduce(x):
    import torch

    x = x.reshape(-1)
    while x.numel() > 1:
        if x.numel() % 2:
            x = torch.cat([x, x.new_zeros(1)])
        x = x[0::2] ^ x[1::2]
    return x[0] if x.numel() else x.new_zeros(())


def verify(out_blocks, before, world: int, rank: int, dev):
    """Checks of a sorted result, on the device: order inside every block and across blocks / ranks (10-byte key,
    ties by record number = input order, ranks in rank order), the payload still carries the tag of ITS key, and
    the key / record-number checksums and the row count equal those of the input."""
    import torch
    import torch.distributed as dist

    from ray_project_ray_b200 import ops
    from ray_project_ray_b200.block import _contig

    ok = {"sorted_within_rank": True, "tie_order": True, "payload_matches_key": True}
    ties = 0
    edges = []          # (hi, lo, recno) of every block's first and last row, in order
    for b in out_blocks:
        if b.num_rows == 0:
            continue
        hi_u = ops.pack_key_bytes(b.columns["key"], 0, 8).view(torch.int64)
        hi = hi_u ^ (-0x8000000000000000)                                    # unsigned order as signed
        lo = ops.pack_key_bytes(b.columns["key"], 8, 2).view(torch.int16).to(torch.int32) & 0xFFFF
        head = _contig(b.columns["payload"][:, 0:16]).view(torch.int64).reshape(-1, 2)
        rec, tag = head[:, 0], head[:, 1]
        same = (hi[1:] == hi[:-1]) & (lo[1:] == lo[:-1])
        in_order = (hi[1:] > hi[:-1]) | ((hi[1:] == hi[:-1]) & (lo[1:] > lo[:-1])) | same
        ok["sorted_within_rank"] &= bool(in_order.all())
        ok["tie_order"] &= bool((rec[1:] > rec[:-1])[same].all())
        ties += int(same.sum())
        ok["payload_matches_key"] &= bool((tag == (hi_u ^ (rec * TAG_MULT))).all())
        edges.append([int(hi[0]), int(lo[0]), int(rec[0])])
        edges.append([int(hi[-1]), int(lo[-1]), int(rec[-1])])
    for a, c in zip(edges[1::2], edges[2::2]):                                 # last of a block vs first of the next
        ok["sorted_within_rank"] &= tuple(c[:2]) >= tuple(a[:2])
        if tuple(c[:2]) == tuple(a[:2]):
            ok["tie_order"] &= c[2] > a[2]
            ties += 1
    after = checksums(out_blocks, dev)
    across = True
    if world > 1:
        mine = torch.tensor((edges[0] + edges[-1]) if edges else [0] * 6, dtype=torch.int64, device=dev)
        flag = torch.tensor([1 if edges else 0], dtype=torch.int64, device=dev)
        all_e = [torch.empty_like(mine) for _ in range(world)]
        all_f = [torch.empty_like(flag) for _ in range(world)]
        dist.all_gather(all_e, mine)
        dist.all_gather(all_f, flag)
        prev = None
        for e, f in zip(all_e, all_f):
            if not int(f.item()):
                continue
            e = [int(v) for v in e.tolist()]
            if prev is not None:
                across &= tuple(e[:2]) >= tuple(prev[3:5])
                if tuple(e[:2]) == tuple(prev[3:5]):
                    ok["tie_order"] &= e[2] > prev[5]
                    ties += 1
            prev = e
        # sums add up over ranks; xors are folded by gathering
        tot = torch.stack([before, after])
        gathered = [torch.empty_like(tot) for _ in range(world)]
        dist.all_gather(gathered, tot)
        g = torch.stack(gathered)                                              # [world, 2, 6]
        before, after = g[0, 0].clone(), g[0, 1].clone()
        for r in range(1, world):
            for k in (0, 1, 3, 4):
                before[k] += g[r, 0, k]
                after[k] += g[r, 1, k]
            for k in (2, 5):
                before[k] ^= g[r, 0, k]
                after[k] ^= g[r, 1, k]
        flags = torch.tensor([int(ok[k]) for k in ("sorted_within_rank", "tie_order", "payload_matches_key")] + [ties],
                             dtype=torch.int64, device=dev)
        allf = [torch.empty_like(flags) for _ in range(world)]
        dist.all_gather(allf, flags)
        st = torch.stack(allf)
        for j, k in enumerate(("sorted_within_rank", "tie_order", "payload_matches_key")):
            ok[k] = bool(st[:, j].min().item())
        ties = int(st[:, 3].sum().item())
    b, a = before.tolist(), after.tolist()
    ok["sorted_across_ranks"] = bool(across)
    ok["rows"] = int(a[0])
    ok["row_count"] = a[0] == b[0]
    ok["key_multiset_checksum"] = a[1:4] == b[1:4]
    ok["record_number_checksum"] = a[4:6] == b[4:6]
    ok["equal_key_pairs_checked"] = ties
    ok["all"] = all(v for k, v in ok.items() if isinstance(v, bool)) and ties > 0
    return ok


def run_ours(args):
    import ctypes

    import torch
    import torch.distributed as dist

    import ray_project_ray_b200.data as rd
    from ray_project_ray_b200 import _lib
    from ray_project_ray_b200.block import _contig

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the device path has no CPU fallback")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)
    lib = _lib.load()

    n = args.rows
    nblocks = 8
    # synthetic input, generated once on the device (seeded per rank); a pinned host copy feeds the e2e leg
    d_key, d_pay = make_records(n, rank, dev)
    h_rec = None
    if not args.no_e2e:      # the e2e leg's input on the host: row-oriented 100-byte records, as a gensort file holds them
        h_rec = torch.empty((n, ROW_BYTES), dtype=torch.uint8).pin_memory()
        h_rec[:, :KEY_BYTES].copy_(d_key)
        h_rec[:, KEY_BYTES:].copy_(d_pay)
    ds = rd.from_torch_columns({"key": d_key, "payload": d_pay}, override_num_blocks=nblocks, records=True, key_prefix=["key"])
    ds_noprefix = None
    if world == 1 and not args.no_extras:   # the same records without the key prefix column (extras)
        ds_noprefix = rd.from_torch_columns({"key": d_key, "payload": d_pay}, override_num_blocks=nblocks, records=True)
    del d_key, d_pay
    torch.cuda.synchronize()
    before = checksums(ds.get_internal_block_refs(), dev)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def step_resident():
        return ds.sort("key")

    # ---------------- device-resident: `value`
    out = None
    for _ in range(args.warmup):
        out = step_resident()
    barrier()
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    launches0 = lib.b200x_launch_count()
    lib.b200x_profile_enable(1)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(args.steps):
        out = step_resident()
    e1.record()
    barrier()
    ms_total = e0.elapsed_time(e1)
    launches = int(lib.b200x_launch_count() - launches0)

    kernel_times = {}
    for kid in KERNELS:
        k_ms, k_cnt = ctypes.c_double(0), ctypes.c_int(0)
        lib.b200x_profile_read(kid, ctypes.byref(k_ms), ctypes.byref(k_cnt))
        if k_cnt.value:
            kernel_times[kid] = (k_ms.value, k_cnt.value)
    lib.b200x_profile_enable(0)
    # correctness of what was timed (outside the timed region): DESIGN.md §6 "verified"
    verified = verify(out.get_internal_block_refs(), before, world, rank, dev)
    local_out_rows = out.count()
    if world > 1:
        t = torch.tensor([ms_total], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms_total = float(t.item())
        rows_all = [torch.zeros(1, dtype=torch.int64, device=dev) for _ in range(world)]
        dist.all_gather(rows_all, torch.tensor([local_out_rows], dtype=torch.int64, device=dev))
        rows_per_rank = [int(x.item()) for x in rows_all]
    else:
        rows_per_rank = [local_out_rows]
    ms_step = ms_total / args.steps
    total_rows = n * world
    value = total_rows / (ms_step / 1e3)
    del out

    # ---------------- end to end through the public API from pinned host buffers: `e2e`
    e2e = None
    pinned_ok = not args.no_e2e
    if not args.no_e2e:
        cap = int(n * (1.0 if world == 1 else 1.25))
        try:
            o_rec = torch.empty((cap, ROW_BYTES), dtype=torch.uint8).pin_memory()
        except RuntimeError as err:      # the host cannot pin another 14 GB per rank: no end-to-end leg (every rank agrees)
            pinned_ok, e2e = False, {"unavailable": f"pinned host allocation failed: {err}"[:200]}
        if world > 1:
            flag = torch.tensor([int(pinned_ok)], dtype=torch.int64, device=dev)
            dist.all_reduce(flag, op=dist.ReduceOp.MIN)
            if pinned_ok and not int(flag.item()):
                e2e = {"unavailable": "pinned host allocation failed on another rank"}
            pinned_ok = bool(flag.item())
    if pinned_ok:
        copy_stream = torch.cuda.Stream(device=dev)
        prev = {"objs": None, "done": None}  # the previous step's result while its D2H may still be in flight

        def step_e2e():
            # H2D of this step's inputs + ingest into records + sort on the main stream; egress (records -> dense
            # columns) + D2H of the result on a second stream, so that the D2H of step i overlaps the H2D of step
            # i+1 (PCIe is full duplex). Every step copies all of its inputs in and all of its result out.
            d = h_rec.to(dev, non_blocking=True)
            res = rd.from_torch_columns({"key": d[:, :KEY_BYTES], "payload": d[:, KEY_BYTES:]}, override_num_blocks=nblocks,
                                        records=True, key_prefix=["key"]).sort("key")
            del d
            done = torch.cuda.Event()
            done.record()
            if prev["done"] is not None:      # the buffers the previous D2H read from are released only once it is over
                prev["done"].synchronize()
            prev["objs"] = None
            copy_stream.wait_event(done)
            lo, objs = 0, [res]
            with torch.cuda.stream(copy_stream):
                for b in res.get_internal_block_refs():
                    m = min(b.num_rows, cap - lo)      # (a rank never receives 25 % more than its share; guarded anyway)
                    if m > 0:
                        if m < b.num_rows:
                            b = b.slice(0, m)
                        dense = b.rows_as_bytes(["key", "payload"])       # records -> dense 100-byte rows
                        o_rec[lo:lo + m].copy_(dense, non_blocking=True)
                        objs.append(dense)
                        lo += m
                ev = torch.cuda.Event()
                ev.record()
            prev["objs"], prev["done"] = objs, ev
            return lo

        e_steps = max(1, min(args.steps, 4))
        for _ in range(min(args.warmup, 1)):
            step_e2e()
        copy_stream.synchronize()
        barrier()
        f0, f1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        f0.record()
        got = 0
        for _ in range(e_steps):
            got = step_e2e()
        with torch.cuda.stream(copy_stream):
            f1.record()
        copy_stream.synchronize()
        barrier()
        e_ms = f0.elapsed_time(f1)
        prev["objs"] = None
        m = min(got, 200_000)
        hk = o_rec[:m, :KEY_BYTES].numpy()
        order_ok = all(bytes(hk[i]) <= bytes(hk[i + 1]) for i in range(0, m - 1, max(1, m // 2000)))
        assert order_ok, "e2e result is not sorted"
        if world > 1:
            t = torch.tensor([e_ms], dtype=torch.float64, device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            e_ms = float(t.item())
        e2e = {"value": total_rows / (e_ms / e_steps / 1e3), "unit": "rows/s", "h2d_bytes_per_step": n * ROW_BYTES,
               "d2h_bytes_per_step": int(got * ROW_BYTES), "ms_per_step": e_ms / e_steps, "steps": e_steps,
               "note": "per rank and step: H2D of the 100-byte records (row-oriented, as a gensort file holds them), ingest "
                       "into 128-byte records + key prefix, Dataset.sort, egress to dense 100-byte rows, D2H of the whole "
                       "sorted result"}
        del o_rec, h_rec

    # ---------------- side measurements (reported under "extras", not the metric)
    extras = None
    if world == 1 and not args.no_extras:
        extras = run_extras(rd, dev, nblocks, ds_noprefix, n)
    elif world > 1 and not args.no_extras:
        del ds
        extras = run_extras_distributed(rd, dev, world, rank, nblocks)
    del ds_noprefix

    clocks = sampler.stop() if rank == 0 else None
    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return

    peaks = {}
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            peaks = json.load(f)
    except OSError:
        pass
    peak = float(peaks.get("hbm_gbs", 6650.0))
    peak_src = "measured (MEASURED_PEAKS.json hbm_gbs)" if "hbm_gbs" in peaks else "fallback 6650 GB/s"
    breakdown = {}
    for kid, (k_ms, k_cnt) in kernel_times.items():
        kname, alg_b, traffic_b = KERNELS[kid]
        per_launch_ms = k_ms / k_cnt
        entry = {"launches": k_cnt, "ms_per_launch": per_launch_ms, "share_of_step": k_ms / ms_total}
        if alg_b is not None:
            rows_k = n if world == 1 else rows_per_rank[0]
            entry["achieved_gbs"] = rows_k * alg_b / (per_launch_ms / 1e3) / 1e9
            entry["frac"] = entry["achieved_gbs"] / peak
        breakdown[kname] = entry
    op_gbs = total_rows * 2 * ROW_BYTES / (ms_step / 1e3) / 1e9          # 2 x row_bytes x rows
    t_hbm = 2 * ROW_BYTES * n / (peak * 1e9)
    t_nvl = ROW_BYTES * n * (world - 1) / world / (NVLINK_GBS * 1e9)
    roofline_op = {"bound": "hbm", "achieved": op_gbs / world, "peak": peak, "unit": "GB/s", "frac": op_gbs / world / peak,
                   "algorithmic_bytes": "2 x row_bytes x rows (SURVEY.md 8d)", "peak_source": peak_src}
    roof = None
    if world == 1 and kernel_times:
        cand = {k: v for k, v in kernel_times.items() if KERNELS[k][1] is not None}
        kid = max(cand, key=lambda k: cand[k][0])   # dominant = largest share of the timed region
        kname, alg_b, traffic_b = KERNELS[kid]
        k_ms, k_cnt = kernel_times[kid]
        per_launch_ms = k_ms / k_cnt
        ach = n * alg_b / (per_launch_ms / 1e3) / 1e9
        roof = {"bound": "hbm", "kernel": kname, "achieved": ach, "peak": peak, "unit": "GB/s", "frac": ach / peak,
                "traffic": (n * traffic_b if traffic_b else None), "peak_source": peak_src, "launches_timed": k_cnt,
                "ms_per_launch": per_launch_ms, "share_of_step": k_ms / ms_total,
                "algorithmic_bytes_per_launch": n * alg_b}
    elif world > 1:
        # the step against the slower of the two rooflines of SURVEY.md 8d: NVLink carries row_bytes x rows x (G-1)/G
        # per GPU and direction, HBM 2 x row_bytes x rows
        nvl_bytes = ROW_BYTES * n * (world - 1) / world
        bound = "nvlink" if t_nvl >= t_hbm else "hbm"
        if bound == "nvlink":
            ach, pk, alg = nvl_bytes / (ms_step / 1e3) / 1e9, NVLINK_GBS, nvl_bytes
            pk_src = "nominal NVLink 5, 900 GB/s per direction and GPU"
        else:
            ach, pk, alg, pk_src = 2 * ROW_BYTES * n / (ms_step / 1e3) / 1e9, peak, 2 * ROW_BYTES * n, peak_src
        # bytes this GPU really stores into its peers per step: 96 bytes of every record (the record without its
        # padding and without the 8-byte key prefix) + the 8-byte key prefix column, for (G-1)/G of its rows
        traffic = (96 + 8) * n * (world - 1) / world
        roof = {"bound": bound, "achieved": ach, "peak": pk, "unit": "GB/s", "frac": ach / pk, "traffic": traffic,
                "traffic_note": "bytes one GPU stores into peer memory per step (96 of every record's 128 bytes + the 8-byte "
                                "key prefix, for (G-1)/G of its rows), computed; not a link counter",
                "peak_source": pk_src, "algorithmic_bytes_per_step": alg, "ms_per_step": ms_step,
                "roofline_ms": max(t_hbm, t_nvl) * 1e3, "hbm_roofline_ms": t_hbm * 1e3, "nvlink_roofline_ms": t_nvl * 1e3,
                "note": "op-level: algorithmic bytes / step time against the slower roofline (SURVEY.md 8d); rank 0's "
                        "kernels are under `kernels` (the key sort overlaps the payload exchange)"}
        roofline_op = {"bound": bound, "achieved": ach, "peak": pk, "unit": "GB/s", "frac": ach / pk,
                       "algorithmic_bytes": "row_bytes x rows x (G-1)/G per GPU and direction (SURVEY.md 8d)" if bound == "nvlink"
                       else "2 x row_bytes x rows (SURVEY.md 8d)", "peak_source": pk_src}
        if 7 in kernel_times:
            x_ms = kernel_times[7][0] / args.steps
            roof["exchange_kernels"] = {"ms_per_step": x_ms, "stored_gbs_per_direction": traffic / (x_ms / 1e3) / 1e9,
                                        "frac_of_nvlink": traffic / (x_ms / 1e3) / 1e9 / NVLINK_GBS}
    line = {
        "metric": "Dataset.sort rows/s", "value": value, "unit": "rows/s", "n_gpus": world, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": ms_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "u8", "data": "synthetic", "gb_per_s": op_gbs, "gigabytes_sorted_per_step": total_rows * ROW_BYTES / 1e9,
        "config": workload_config(n, world, nblocks),
        "roofline": roof, "roofline_op": roofline_op, "verified": verified, "rows_per_rank_after": rows_per_rank,
        "imbalance_max_over_mean": max(rows_per_rank) / (sum(rows_per_rank) / len(rows_per_rank)),
        "kernels": breakdown, "e2e": e2e, "extras": extras, "gpu_launches": launches, "clocks": clocks,
    }
    if not args.no_cpu_baseline and world == 1:
        import pyarrow

        sec, threads, nb = cpu_arm(args.cpu_rows, 2, 1)
        line["cpu_baseline"] = {"value": args.cpu_rows / sec, "unit": "rows/s", "cores": threads, "kind": "port",
                                "sample": f"{args.cpu_rows} gensort records x {ROW_BYTES} B (sort(['hi','lo'])), {nb} blocks, "
                                          f"pull-based map/reduce on a {threads}-thread pool, pyarrow {pyarrow.__version__}, "
                                          f"{sec:.2f} s per sort"}
    print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


def run_extras(rd, dev, nblocks, ds_noprefix, n_main):
    """The other operators / shapes of the path on one GPU: round 1's 72-byte sort, random_shuffle (configs[1]) and
    groupby-aggregate (configs[2], uniform and Zipf(1.1) keys)."""
    import torch

    from ray_project_ray_b200.aggregate import Count, Mean, Sum

    def timed(fn, steps=3, warm=2):
        for _ in range(warm):
            fn()
        torch.cuda.synchronize()
        g0, g1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        g0.record()
        for _ in range(steps):
            r = fn()
        g1.record()
        torch.cuda.synchronize()
        return g0.elapsed_time(g1) / steps, r

    peak_x = 6548.8
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            peak_x = float(json.load(f).get("hbm_gbs", peak_x))
    except OSError:
        pass
    extras = {}
    if ds_noprefix is not None:
        ms, r = timed(lambda: ds_noprefix.sort("key"))
        assert r.count() == n_main
        gbs = n_main * 2 * ROW_BYTES / (ms / 1e3) / 1e9
        extras["gensort_without_key_prefix_column"] = {
            "workload": "the metric's workload with the blocks holding the records only: every sort first reads the 10-byte "
                        "keys out of the 128-byte records (one pass over all records)",
            "ms_per_step": ms, "rows_per_s": n_main / (ms / 1e3), "gb_per_s": gbs, "frac_of_hbm_peak": gbs / peak_x}
        del r
    n = 100_000_000
    gen = torch.Generator(device=dev).manual_seed(3)
    ids = torch.randint(0, 2**63 - 1, (n,), dtype=torch.int64, device=dev, generator=gen)
    pay = torch.randint(0, 256, (n, 64), dtype=torch.uint8, device=dev, generator=gen)
    ds72 = rd.from_torch_columns({"id": ids, "payload": pay}, override_num_blocks=nblocks)
    ms, r = timed(lambda: ds72.sort("id"))
    assert r.count() == n
    gbs = n * 2 * 72 / (ms / 1e3) / 1e9
    extras["sort_100Mx72B"] = {"workload": "round 1's bench workload: 100 M rows x (int64 key uniform in [0, 2^63) + 64 B payload), dense columns",
                               "ms_per_step": ms, "rows_per_s": n / (ms / 1e3), "gb_per_s": gbs, "frac_of_hbm_peak": gbs / peak_x}
    del r
    ms, r = timed(lambda: ds72.random_shuffle(seed=0))
    assert r.count() == n
    gbs = n * 2 * 72 / (ms / 1e3) / 1e9
    extras["random_shuffle"] = {"workload": f"configs[1]: {n} rows x (8 B id + 64 B payload), seed=0, one global bijection, dense columns",
                                "ms_per_step": ms, "rows_per_s": n / (ms / 1e3), "gb_per_s": gbs, "frac_of_hbm_peak": gbs / peak_x}
    del r
    ds72r = rd.from_torch_columns({"id": ids, "payload": pay}, override_num_blocks=nblocks, records=True)
    del ds72, ids, pay
    ms, r = timed(lambda: ds72r.random_shuffle(seed=0))
    assert r.count() == n
    gbs = n * 2 * 72 / (ms / 1e3) / 1e9
    extras["random_shuffle_records"] = {"workload": f"the same rows as 128-byte records (id and payload move together)",
                                        "ms_per_step": ms, "rows_per_s": n / (ms / 1e3), "gb_per_s": gbs, "frac_of_hbm_peak": gbs / peak_x}
    del r, ds72r
    gn, groups = 200_000_000, 1_000_000
    gen = torch.Generator(device=dev).manual_seed(4)
    for tag in ("uniform", "zipf1.1"):
        if tag == "uniform":
            gk = torch.randint(0, groups, (gn,), dtype=torch.int64, device=dev, generator=gen)
        else:
            # Zipf(1.1) over `groups` ranks by inverse CDF, ranks spread over the key range by a fixed permutation
            w = torch.arange(1, groups + 1, dtype=torch.float64, device=dev).pow(-1.1)
            cdf = torch.cumsum(w / w.sum(), 0)
            u = torch.rand(gn, dtype=torch.float64, device=dev, generator=gen)
            ranks = torch.searchsorted(cdf, u).clamp_(max=groups - 1)
            spread = torch.randperm(groups, device=dev, generator=gen)
            gk = spread[ranks]
            del w, cdf, u, ranks, spread
        gv = torch.randint(-2**31, 2**31, (gn,), dtype=torch.int64, device=dev, generator=gen)
        gds = rd.from_torch_columns({"k": gk, "v": gv}, override_num_blocks=nblocks)
        ms, r = timed(lambda: gds.groupby("k").aggregate(Sum("v"), Mean("v"), Count()))
        distinct = r.count()
        assert int(r.sum("count()")) == gn and int(r.sum("sum(v)")) == int(gv.sum().item())
        gbs = (gn * 16 + distinct * 32) / (ms / 1e3) / 1e9
        extras[f"groupby_aggregate_{tag}"] = {
            "workload": f"configs[2]: {gn} rows, int64 keys ({tag}, {distinct} distinct), Sum+Mean+Count of one int64 column",
            "ms_per_step": ms, "rows_per_s": gn / (ms / 1e3), "gb_per_s": gbs, "frac_of_hbm_peak": gbs / peak_x,
            "checked": "row count and grand total of the group sums equal the input's"}
        del r, gds, gk, gv
    return extras


def run_extras_distributed(rd, dev, world, rank, nblocks):
    """The other two operators of the path across the ranks, each checked on the device (all ranks take part, rank 0
    reports): random_shuffle of 50 M rows x 72 B per GPU, groupby-aggregate of 100 M rows per GPU over 1 M keys."""
    import torch
    import torch.distributed as dist

    from ray_project_ray_b200.aggregate import Count, Mean, Sum

    def timed(fn, steps=2, warm=1):
        for _ in range(warm):
            r = fn()
        dist.barrier()
        torch.cuda.synchronize()
        g0, g1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        g0.record()
        for _ in range(steps):
            r = fn()
        g1.record()
        dist.barrier()
        torch.cuda.synchronize()
        t = torch.tensor([g0.elapsed_time(g1) / steps], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item()), r

    def total(x):
        t = torch.as_tensor(x, dtype=torch.int64, device=dev).reshape(1).clone()
        dist.all_reduce(t)
        return int(t.item())

    extras = {}
    n = 50_000_000
    ids = torch.arange(rank * n, (rank + 1) * n, dtype=torch.int64, device=dev)
    pay = torch.randint(0, 256, (n, 64), dtype=torch.uint8, device=dev, generator=torch.Generator(device=dev).manual_seed(5 + rank))
    pay[:, :8] = (ids * TAG_MULT).view(torch.uint8).reshape(n, 8)
    ds = rd.from_torch_columns({"id": ids, "payload": pay}, override_num_blocks=nblocks)
    ms, r = timed(lambda: ds.random_shuffle(seed=0))
    blocks = [b for b in r.get_internal_block_refs() if b.num_rows]
    oid = torch.cat([b["id"] for b in blocks])
    opay = torch.cat([b["payload"][:, :8].contiguous().view(torch.int64).reshape(-1) for b in blocks])
    again = torch.cat([b["id"] for b in ds.random_shuffle(seed=0).get_internal_block_refs() if b.num_rows])
    other = torch.cat([b["id"] for b in ds.random_shuffle(seed=1).get_internal_block_refs() if b.num_rows])
    home = (oid // n == rank).sum()
    ok = {"rows": total(oid.numel()) == n * world,
          "id_multiset_checksum": total(oid.sum()) == total(ids.sum()) and _all_equal_xor(_xor_reduce(oid), _xor_reduce(ids), dev),
          "payload_travels_with_id": bool((opay == oid * TAG_MULT).all()),
          "same_seed_same_order": oid.numel() == again.numel() and bool((oid == again).all()),
          "other_seed_other_order": oid.numel() != other.numel() or float((oid != other).double().mean()) > 0.9,
          "rows_mix_across_ranks": abs(total(home) / (n * world) - 1.0 / world) < 0.01}
    ok = {k: bool(total(int(v)) == world) for k, v in ok.items()}
    gbs = n * world * 2 * 72 / (ms / 1e3) / 1e9
    extras["random_shuffle"] = {"workload": f"{n} rows x (8 B id + 64 B payload) per GPU, seed=0, across {world} GPUs",
                                "ms_per_step": ms, "rows_per_s": n * world / (ms / 1e3), "gb_per_s": gbs, "verified": ok}
    del ds, ids, pay, r, blocks, oid, opay, again, other
    gn, groups = 100_000_000, 1_000_000
    gen = torch.Generator(device=dev).manual_seed(40 + rank)
    gk = torch.randint(0, groups, (gn,), dtype=torch.int64, device=dev, generator=gen)
    gv = torch.randint(-2**31, 2**31, (gn,), dtype=torch.int64, device=dev, generator=gen)
    gds = rd.from_torch_columns({"k": gk, "v": gv}, override_num_blocks=nblocks)
    ms, r = timed(lambda: gds.groupby("k").aggregate(Sum("v"), Mean("v"), Count()))
    blocks = [b for b in r.get_internal_block_refs() if b.num_rows]
    k = torch.cat([b["k"] for b in blocks]) if blocks else gk[:0]
    cnt = torch.cat([b["count()"] for b in blocks]) if blocks else gk[:0]
    sm = torch.cat([b["sum(v)"] for b in blocks]) if blocks else gk[:0]
    mean = torch.cat([b["mean(v)"] for b in blocks]) if blocks else gk[:0].double()
    edge = torch.tensor([int(k[0]), int(k[-1])] if k.numel() else [2**62, -1], dtype=torch.int64, device=dev)
    edges = [torch.empty_like(edge) for _ in range(world)]
    dist.all_gather(edges, edge)
    e = [x.tolist() for x in edges if x[1] >= 0]
    ok = {"keys_ascending_and_distinct": bool((k[1:] > k[:-1]).all()) and all(b[0] > a[1] for a, b in zip(e[:-1], e[1:])),
          "row_count": total(cnt.sum()) == gn * world, "grand_total": total(sm.sum()) == total(gv.sum()),
          "groups": total(k.numel()) == groups,
          "mean_is_sum_over_count": bool(torch.allclose(mean, sm.double() / cnt.double(), rtol=1e-12, atol=0))}
    ok = {kk: bool(total(int(v)) == world) for kk, v in ok.items()}
    extras["groupby_aggregate"] = {"workload": f"{gn} rows per GPU, {groups} distinct int64 keys (uniform), Sum+Mean+Count, across {world} GPUs",
                                   "ms_per_step": ms, "rows_per_s": gn * world / (ms / 1e3), "verified": ok}
    return extras


def _all_equal_xor(mine, theirs, dev):
    """XOR of a value over all ranks == XOR of another value over all ranks."""
    import torch
    import torch.distributed as dist

    world = dist.get_world_size()
    pair = torch.stack([mine.reshape(()), theirs.reshape(())]).to(dev)
    parts = [torch.empty_like(pair) for _ in range(world)]
    dist.all_gather(parts, pair)
    a = b = 0
    for p in parts:
        a ^= int(p[0])
        b ^= int(p[1])
    return a == b


if __name__ == "__main__":
    a = parse()
    if a.impl == "reference":
        run_reference(a)
    else:
        run_ours(a)
