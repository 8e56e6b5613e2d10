"""Timed CPU baseline: a port of the reference's Dataset.sort path that calls the SAME third-party
kernels the reference calls (pyarrow.compute.sort_indices + Table.take, numpy.searchsorted), in
the reference's task structure, on the host cores.

TEST / BENCH INFRASTRUCTURE ONLY (see oracle/exchange_oracle.py). Used by bench.py's
`cpu_baseline` leg and by `bench.py --impl reference`; the reference itself (ray) cannot be
installed or imported on the GPU box, so kind = "port".

Structure restated (citations relative to /root/reference/python/ray/data/):
  sample       SortTaskSpec.sample_boundaries          _internal/planner/exchange/sort_task_spec.py:149-199
  map task     SortTaskSpec.map -> sort_and_partition   sort_task_spec.py:124-137, _internal/arrow_block.py:449-464
                 transform_pyarrow.sort                 _internal/arrow_ops/transform_pyarrow.py:17-21
                 find_partitions                        _internal/util.py:719-779
  reduce task  SortTaskSpec.reduce -> merge_sorted_blocks  sort_task_spec.py:139-147, arrow_block.py:554-567
                 transform_pyarrow.concat_and_sort      transform_pyarrow.py:269-276
  scheduler    PullBasedShuffleTaskScheduler            exchange/pull_based_shuffle_task_scheduler.py:86-124
Tasks run on a thread pool (pyarrow releases the GIL), one thread per host core; Ray's task,
object-store and serialisation overheads are absent, which flatters the reference.
"""
from __future__ import annotations

import os
import time
from concurrent.futures import ThreadPoolExecutor
from typing import List, Sequence, Tuple

import numpy as np
import pyarrow as pa
import pyarrow.compute as pac


def host_cores() -> int:
    try:
        return len(os.sched_getaffinity(0))
    except AttributeError:  # pragma: no cover
        return os.cpu_count() or 1


def _sort_table(t: pa.Table, key, descending: bool) -> pa.Table:
    keys = [key] if isinstance(key, str) else list(key)
    idx = pac.sort_indices(t, sort_keys=[(k, "descending" if descending else "ascending") for k in keys])
    return t.take(idx)


def _first(key) -> str:
    return key if isinstance(key, str) else key[0]


def _map_task(t: pa.Table, key: str, descending: bool, boundaries: Sequence) -> List[pa.Table]:
    if t.num_rows == 0:
        return [t] * (len(boundaries) + 1)
    s = _sort_table(t, key, descending)
    col = s[_first(key)].to_numpy()   # multi-column keys are cut on their first column (equal first keys stay together)
    parts, last = [], 0
    for b in boundaries:
        if descending:
            idx = len(col) - int(np.searchsorted(col[::-1], b, side="left"))
        else:
            idx = int(np.searchsorted(col, b, side="left"))
        parts.append(s.slice(last, idx - last))
        last = idx
    parts.append(s.slice(last))
    return parts


def _reduce_task(parts: Sequence[pa.Table], key: str, descending: bool) -> pa.Table:
    parts = [p for p in parts if p.num_rows > 0]
    if not parts:
        return pa.table({})
    return _sort_table(pa.concat_tables(parts), key, descending)


def sort_blocks(blocks: List[pa.Table], key, descending: bool, threads: int, seed: int = 0) -> List[pa.Table]:
    """`key`: a column name or a list of them (the reference's multi-column SortKey, sort_task_spec.py:18-87)."""
    R = len(blocks)
    rng = np.random.default_rng(seed)
    n_samples = max(1, int(R * 10 / len(blocks)))
    samples = np.concatenate([b[_first(key)].to_numpy()[rng.integers(0, b.num_rows, min(n_samples, b.num_rows))]
                              for b in blocks if b.num_rows])
    samples.sort()
    boundaries = list(np.quantile(samples, np.linspace(0, 1, R), method="nearest")[1:])
    if descending:
        boundaries.reverse()
    with ThreadPoolExecutor(max_workers=threads) as pool:
        mapped = list(pool.map(lambda b: _map_task(b, key, descending, boundaries), blocks))
        outs = list(pool.map(lambda j: _reduce_task([m[j] for m in mapped], key, descending), range(R)))
    return outs


def make_blocks(n_rows: int, payload_bytes: int, num_blocks: int, seed: int = 0) -> List[pa.Table]:
    """The bench workload on the host: random int64 keys in [0, 2^63) (as the reference's own sort
    benchmark draws them, release/nightly_tests/dataset/sort.py:39-47) + a fixed-width payload."""
    rng = np.random.default_rng(seed)
    per = n_rows // num_blocks
    out = []
    for i in range(num_blocks):
        m = per if i < num_blocks - 1 else n_rows - per * (num_blocks - 1)
        keys = rng.integers(0, np.iinfo(np.int64).max, m, dtype=np.int64)
        cols = {"id": pa.array(keys)}
        if payload_bytes:
            pay = rng.integers(0, 256, (m, payload_bytes), dtype=np.uint8)
            cols["payload"] = pa.FixedSizeBinaryArray.from_buffers(pa.binary(payload_bytes), m, [None, pa.py_buffer(pay)])
        out.append(pa.table(cols))
    return out


def make_gensort_blocks(n_rows: int, num_blocks: int, seed: int = 0) -> List[pa.Table]:
    """gensort-shaped 100-byte records on the host (BASELINE.json configs[3]): the 10-byte key as the
    reference can sort it - (hi: uint64 = key bytes 0..7 big-endian, lo: uint16 = bytes 8..9), `sort(["hi", "lo"])`,
    the form SURVEY.md 8c ran through the reference's own sample_boundaries / find_partition_index - plus a 90-byte
    fixed_size_binary payload: 100 bytes per row."""
    rng = np.random.default_rng(seed)
    per = n_rows // num_blocks
    out = []
    for i in range(num_blocks):
        m = per if i < num_blocks - 1 else n_rows - per * (num_blocks - 1)
        hi = rng.integers(0, np.iinfo(np.uint64).max, m, dtype=np.uint64)
        lo = rng.integers(0, 1 << 16, m, dtype=np.uint16)
        pay = rng.integers(0, 256, (m, 90), dtype=np.uint8)
        out.append(pa.table({"hi": pa.array(hi), "lo": pa.array(lo),
                             "payload": pa.FixedSizeBinaryArray.from_buffers(pa.binary(90), m, [None, pa.py_buffer(pay)])}))
    return out


def time_sort(n_rows: int, payload_bytes: int, threads: int, repeats: int = 1, num_blocks: int = 0) -> Tuple[float, int]:
    """Seconds per Dataset.sort of n_rows rows (best of `repeats`), and the block count used."""
    num_blocks = num_blocks or max(threads, 8)
    blocks = make_blocks(n_rows, payload_bytes, num_blocks)
    best = float("inf")
    for _ in range(repeats):
        t0 = time.perf_counter()
        outs = sort_blocks(blocks, "id", False, threads)
        best = min(best, time.perf_counter() - t0)
        assert sum(o.num_rows for o in outs) == n_rows
    return best, num_blocks
