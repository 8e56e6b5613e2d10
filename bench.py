#!/usr/bin/env python
"""Headline benchmark: Dataset.sort throughput on B200 (BASELINE.json `metric`).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...

Workload (config.workload = "sort_100Mx72B"): per GPU 100 M rows of (int64 key `id` drawn uniformly
from [0, 2^63) as the reference's own sort benchmark draws them,
release/nightly_tests/dataset/sort.py:39-47, + one 64-byte fixed-width `payload` column) — the
row shape of BASELINE.json configs[1], the largest configuration that fits one GPU — sorted by
`Dataset.sort("id")`. A step = one full sort of the resident dataset. N > 1: every rank holds
100 M rows (weak scaling) and the sort is global (range-partitioned all-to-all over NCCL).

Printed JSON (rank 0, one line): see DESIGN.md §Measurement for every key.
  value      rows/s, inputs resident in HBM, CUDA events around K steps, max over ranks
  e2e        same metric through the public API from pinned HOST buffers (H2D + sort + D2H per step)
  roofline   the kernel with the largest share of the step, CUDA events on its launching stream (all kernels: `kernels`)
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

ROWS_PER_GPU = 100_000_000
PAYLOAD_BYTES = 64
ROW_BYTES = 8 + PAYLOAD_BYTES
CPU_SAMPLE_ROWS = 100_000_000         # the CPU arm sorts the full single-GPU workload (a few seconds on 16+ cores)
# Algorithmic bytes per row of one launch of each timed kernel class, and the DRAM traffic
# per row ncu measured for it (profiles/, dram__bytes_read.sum + dram__bytes_write.sum; None = not captured).
KERNELS = {
    0: ("onesweep_pass_kernel", 22.0, 26.0),         # HYBRID: pass A reads 8 writes 12, pass B reads 12 writes 12
    1: ("gather_rows_kernel<uint4>", 4 + 2 * PAYLOAD_BYTES, 197.6),  # index + read row + write row (64 B rows cost 128 B of DRAM reads)
    2: ("finish_runs_kernel", 12 + 4 + 8, 23.5),      # read pair, write row index + decoded key
    3: ("digit2_hist_kernel", 8.0, 8.0),              # read key column
    6: ("hist_kernel", 8.0, None),
    # peer exchange (N > 1): the key column (8 B rows) and the payload (64 B rows) are each read and stored once,
    # in one launch each: (2 * 8 + 2 * 64) / 2 bytes per row and launch on average
    7: ("scatter_rows_staged_kernel (peer exchange)", float(8 + PAYLOAD_BYTES), None),
}


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--rows", type=int, default=ROWS_PER_GPU, help="rows per GPU")
    ap.add_argument("--cpu-rows", type=int, default=CPU_SAMPLE_ROWS)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-extras", action="store_true", help="skip the random_shuffle / groupby side measurements")
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


def cpu_arm(rows: int, steps: int, warmup: int):
    """The reference's CPU path (oracle port) on the host cores: seconds per sort of `rows` rows."""
    from oracle import cpu_baseline as cb

    threads = cb.host_cores()
    blocks = cb.make_blocks(rows, PAYLOAD_BYTES, max(threads, 8))
    times = []
    for i in range(warmup + steps):
        t0 = time.perf_counter()
        outs = cb.sort_blocks(blocks, "id", False, threads)
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
    steps, warmup = max(1, min(args.steps, 3)), min(args.warmup, 1)
    sec, threads, nblocks = cpu_arm(rows, steps, warmup)
    value = rows / sec
    sample = (f"{rows} rows x {ROW_BYTES} B (same key/payload generator, {nblocks} blocks), pull-based map/reduce on a "
              f"{threads}-thread pool, pyarrow {pyarrow.__version__}; rows/s is size-normalised")
    line = {
        "impl": "reference", "metric": "Dataset.sort rows/s", "value": value, "unit": "rows/s", "n_gpus": args.gpus,
        "steps": steps, "warmup": warmup, "ms_per_step": sec * 1e3, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "int64", "data": "synthetic",
        "gb_per_s": value * 2 * ROW_BYTES / 1e9,
        "config": {"workload": "sort_100Mx72B", "rows_per_gpu": args.rows, "row_bytes": ROW_BYTES, "key": "int64 uniform [0,2^63)",
                   "payload_bytes": PAYLOAD_BYTES, "cpu_sample_rows": rows},
        "cpu_baseline": {"value": value, "unit": "rows/s", "cores": threads, "kind": "port", "sample": sample},
        "e2e": {"value": value, "unit": "rows/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    print(json.dumps(line), flush=True)


def run_ours(args):
    import numpy as np
    import torch
    import torch.distributed as dist

    import ray_project_ray_b200.data as rd
    from ray_project_ray_b200 import _lib
    from ray_project_ray_b200.block import DeviceBlock

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
    # synthetic input, generated once on the host (seeded per rank), kept in pinned memory for the e2e leg
    rng = np.random.default_rng(1000 + rank)
    h_keys = torch.empty(n, dtype=torch.int64).pin_memory()
    h_pay = torch.empty((n, PAYLOAD_BYTES), dtype=torch.uint8).pin_memory()
    chunk = 10_000_000
    for lo in range(0, n, chunk):
        hi = min(n, lo + chunk)
        h_keys[lo:hi] = torch.from_numpy(rng.integers(0, np.iinfo(np.int64).max, hi - lo, dtype=np.int64))
        h_pay[lo:hi] = torch.from_numpy(rng.integers(0, 256, (hi - lo, PAYLOAD_BYTES), dtype=np.uint8))
    d_keys = h_keys.to(dev, non_blocking=True)
    d_pay = h_pay.to(dev, non_blocking=True)
    torch.cuda.synchronize()
    nblocks = 8
    ds = rd.from_torch_columns({"id": d_keys, "payload": d_pay}, override_num_blocks=nblocks)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def step_resident():
        return ds.sort("id")

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
    import ctypes

    kernel_times = {}
    for kid, (kname, _, _) in KERNELS.items():
        k_ms, k_cnt = ctypes.c_double(0), ctypes.c_int(0)
        lib.b200x_profile_read(kid, ctypes.byref(k_ms), ctypes.byref(k_cnt))
        if k_cnt.value:
            kernel_times[kid] = (k_ms.value, k_cnt.value)
    lib.b200x_profile_enable(0)
    # correctness of what was timed: global order + multiset checksum (cheap, outside the timed region)
    host_first = out.get_internal_block_refs()[0]["id"][:4096].cpu().numpy()
    assert (np.diff(host_first) >= 0).all(), "output is not sorted"
    assert out.count() > 0 if world > 1 else out.count() == n
    if world > 1:
        t = torch.tensor([ms_total], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms_total = float(t.item())
        cnt = torch.tensor([out.count()], dtype=torch.int64, device=dev)
        dist.all_reduce(cnt)
        assert int(cnt.item()) == n * world, "rows lost in the exchange"
    ms_step = ms_total / args.steps
    total_rows = n * world
    value = total_rows / (ms_step / 1e3)
    del out

    # ---------------- end to end through the public API from pinned host buffers: `e2e`
    e2e = None
    if not args.no_e2e:
        o_keys = torch.empty(int(n * (1.0 if world == 1 else 1.6)), dtype=torch.int64).pin_memory()
        o_pay = torch.empty((o_keys.numel(), PAYLOAD_BYTES), dtype=torch.uint8).pin_memory()

        copy_stream = torch.cuda.Stream(device=dev)
        keep = []  # results whose D2H is still in flight

        def step_e2e():
            # H2D of this step's inputs + sort on the main stream; D2H of the result on a second stream, so
            # that the D2H of step i overlaps the H2D of step i+1 (PCIe is full duplex). Every step still
            # copies all of its inputs in and all of its result out.
            k = h_keys.to(dev, non_blocking=True)
            p = h_pay.to(dev, non_blocking=True)
            res = rd.from_torch_columns({"id": k, "payload": p}, override_num_blocks=nblocks).sort("id")
            done = torch.cuda.Event()
            done.record()
            copy_stream.wait_event(done)
            lo = 0
            with torch.cuda.stream(copy_stream):
                for b in res.get_internal_block_refs():
                    m = b.num_rows
                    if m:
                        o_keys[lo:lo + m].copy_(b["id"], non_blocking=True)
                        o_pay[lo:lo + m].copy_(b["payload"], non_blocking=True)
                        lo += m
            keep.append(res)
            if len(keep) > 2:
                keep.pop(0)
            return lo

        e_steps = max(1, min(args.steps, 5))
        for _ in range(min(args.warmup, 2)):
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
        keep.clear()
        assert (np.diff(o_keys[:min(got, 100_000)].numpy()) >= 0).all()
        d2h = got * ROW_BYTES
        if world > 1:
            t = torch.tensor([e_ms], dtype=torch.float64, device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            e_ms = float(t.item())
        e2e = {"value": total_rows / (e_ms / e_steps / 1e3), "unit": "rows/s", "h2d_bytes_per_step": n * ROW_BYTES,
               "d2h_bytes_per_step": int(d2h), "ms_per_step": e_ms / e_steps, "steps": e_steps}
        del o_keys, o_pay

    # ---------------- the other two operators of the path (N = 1 only; reported under "extras", not the metric)
    extras = None
    if world == 1 and not args.no_extras:
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
        ms, r = timed(lambda: ds.random_shuffle(seed=0))
        assert r.count() == n
        gbs = n * 2 * ROW_BYTES / (ms / 1e3) / 1e9
        extras["random_shuffle"] = {"workload": f"configs[1]: {n} rows x (8 B id + {PAYLOAD_BYTES} B payload), seed=0, one global bijection",
                                    "ms_per_step": ms, "rows_per_s": n / (ms / 1e3), "gb_per_s": gbs, "frac_of_hbm_peak": gbs / peak_x}
        del r
        gn, groups = 2 * n, 1_000_000
        gen = torch.Generator(device=dev).manual_seed(4)
        gk = torch.randint(0, groups, (gn,), dtype=torch.int64, device=dev, generator=gen)
        gv = torch.randint(-2**31, 2**31, (gn,), dtype=torch.int64, device=dev, generator=gen)
        gds = rd.from_torch_columns({"k": gk, "v": gv}, override_num_blocks=nblocks)
        ms, r = timed(lambda: gds.groupby("k").aggregate(Sum("v"), Mean("v"), Count()))
        assert r.count() == groups and int(r.sum("count()")) == gn
        gbs = (gn * 16 + groups * 32) / (ms / 1e3) / 1e9
        extras["groupby_aggregate"] = {"workload": f"configs[2]: {gn} rows, {groups} distinct int64 keys (uniform), Sum+Mean+Count of one int64 column",
                                       "ms_per_step": ms, "rows_per_s": gn / (ms / 1e3), "gb_per_s": gbs, "frac_of_hbm_peak": gbs / peak_x}
        del r, gds, gk, gv

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
    roof, breakdown = None, {}
    for kid, (k_ms, k_cnt) in kernel_times.items():
        kname, alg_b, traffic_b = KERNELS[kid]
        per_launch_ms = k_ms / k_cnt
        ach = n * alg_b / (per_launch_ms / 1e3) / 1e9
        breakdown[kname] = {"launches": k_cnt, "ms_per_launch": per_launch_ms, "share_of_step": k_ms / ms_total,
                            "achieved_gbs": ach, "frac": ach / peak}
    if kernel_times:
        kid = max(kernel_times, key=lambda k: kernel_times[k][0])   # dominant = largest share of the timed region
        kname, alg_b, traffic_b = KERNELS[kid]
        k_ms, k_cnt = kernel_times[kid]
        per_launch_ms = k_ms / k_cnt
        ach = n * alg_b / (per_launch_ms / 1e3) / 1e9
        roof = {"bound": "hbm", "kernel": kname, "achieved": ach, "peak": peak, "unit": "GB/s", "frac": ach / peak,
                "traffic": (n * traffic_b if traffic_b else None), "peak_source": peak_src, "launches_timed": k_cnt,
                "ms_per_launch": per_launch_ms, "share_of_step": k_ms / ms_total,
                "algorithmic_bytes_per_launch": n * alg_b}
        if world > 1:
            roof["note"] = ("rank 0's kernels, timed one by one; the key sort overlaps the payload exchange, so shares "
                            "add up to more than the step")
            if 7 in kernel_times:
                # the exchange is link-bound: remote bytes this GPU stores per step against NVLink
                x_ms = kernel_times[7][0] / args.steps
                remote = n * ROW_BYTES * (world - 1) / world
                roof["nvlink"] = {"bytes_per_gpu_per_step": remote, "exchange_ms_per_step": x_ms,
                                  "achieved": remote / (x_ms / 1e3) / 1e9, "peak": 900.0, "unit": "GB/s",
                                  "frac": remote / (x_ms / 1e3) / 1e9 / 900.0,
                                  "peak_source": "nominal NVLink 5 per direction"}
    op_gbs = total_rows * 2 * ROW_BYTES / (ms_step / 1e3) / 1e9
    line = {
        "metric": "Dataset.sort rows/s", "value": value, "unit": "rows/s", "n_gpus": world, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": ms_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "int64", "data": "synthetic", "gb_per_s": op_gbs,
        "config": {"workload": "sort_100Mx72B", "rows_per_gpu": n, "row_bytes": ROW_BYTES,
                   "key": "int64 uniform [0,2^63)", "payload_bytes": PAYLOAD_BYTES, "blocks_per_gpu": nblocks,
                   "l2": "inputs (7.2 GB per GPU) exceed the 126 MB L2; no flush needed",
                   "parallelism": f"range-partitioned x{world}" if world > 1 else "single GPU"},
        "roofline": roof,
        "roofline_op": {"bound": "hbm", "achieved": op_gbs / world, "peak": peak, "unit": "GB/s", "frac": op_gbs / world / peak,
                        "algorithmic_bytes": "2 x row_bytes x rows (SURVEY.md 8d)", "peak_source": peak_src},
        "kernels": breakdown, "e2e": e2e, "extras": extras, "gpu_launches": launches, "clocks": clocks,
    }
    if not args.no_cpu_baseline and world == 1:
        import pyarrow

        sec, threads, nb = cpu_arm(args.cpu_rows, 1, 0)
        line["cpu_baseline"] = {"value": args.cpu_rows / sec, "unit": "rows/s", "cores": threads, "kind": "port",
                                "sample": f"{args.cpu_rows} rows x {ROW_BYTES} B, {nb} blocks, pull-based map/reduce on a "
                                          f"{threads}-thread pool, pyarrow {pyarrow.__version__}, {sec:.2f} s"}
    print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    a = parse()
    if a.impl == "reference":
        run_reference(a)
    else:
        run_ours(a)
