"""CPU oracle for the all-to-all hot path (sort / random_shuffle / groupby-aggregate).

TEST INFRASTRUCTURE ONLY. Nothing under `ray-project-ray_b200/` imports this module; only
`tests/`, `__graft_entry__.smoke()` and `bench.py`'s cpu_baseline / `--impl reference` legs
may use it, and there only as the checker or as the timed CPU baseline.

It restates, in numpy, the reference's algorithm for the path (all citations relative to
/root/reference/python/ray/data/). The arithmetic the reference delegates to third-party
wheels that are not under /root/reference is restated from their published behaviour:
  * pyarrow (CI pin 12.0.1, python/requirements_compiled.txt:1606; floor >= 6.0.1):
    compute.sort_indices is a STABLE sort in both directions, -0.0 == +0.0, NaN after all
    numbers (then nulls) in both directions; compute.sum on int64 wraps; Table.take gathers.
  * numpy (CI pin 1.24.4, :1216): searchsorted, lexsort, quantile(interpolation="nearest").
Parity status: PINNED. tests/test_oracle.py checks every function here against
  (a) the golden vectors the reference's own tests hold for this path
      (tests/test_sort.py:23-51, :79-94, :142-191; tests/test_all_to_all.py:187-317, :508-580, :720-775), and
  (b) tests/golden/*.npz: outputs of the reference's own code (SortTaskSpec, ShuffleTaskSpec,
      SortAggregateTaskSpec, Pull/PushBasedShuffleTaskScheduler, generate_*_fn) run unmodified
      in this container under a stub `ray` package by tests/golden/make_golden.py, and
  (c) pyarrow.compute.sort_indices / TableGroupBy.aggregate of the installed pyarrow.
"""
from __future__ import annotations

import math
from typing import Dict, List, Optional, Sequence, Tuple

import numpy as np

Table = Dict[str, np.ndarray]  # column name -> [n] or [n, w] array; all columns same length

_U64 = np.uint64


# --------------------------------------------------------------------------------------
# key order (what pyarrow.compute.sort_indices compares; call site
# _internal/arrow_ops/transform_pyarrow.py:17-21)
# --------------------------------------------------------------------------------------
def encode_keys(col: np.ndarray, descending: bool) -> np.ndarray:
    """Order-preserving uint64 codes: ascending code order == the sort order of the column."""
    col = np.ascontiguousarray(col)
    kind, size = col.dtype.kind, col.dtype.itemsize
    bits = size * 8
    udt = {1: np.uint8, 2: np.uint16, 4: np.uint32, 8: np.uint64}[size]
    full = udt(np.iinfo(udt).max)
    if kind == "i":
        c = col.view(udt) ^ udt(1 << (bits - 1))
        if descending:
            c = ~c
        return c.astype(_U64)
    if kind == "u":
        c = col.view(udt).copy()
        if descending:
            c = ~c
        return c.astype(_U64)
    if kind == "f":
        b = col.view(udt).copy()
        sign = udt(1 << (bits - 1))
        absb = b & ~sign
        exp_all = udt(0x7FF0000000000000) if size == 8 else udt(0x7F800000)
        nan = absb > exp_all
        b[absb == 0] = 0  # -0.0 ties with +0.0
        neg = (b & sign) != 0
        c = np.where(neg, ~b, b | sign).astype(udt)
        if descending:
            c = (~c).astype(udt)
        out = c.astype(_U64)
        out[nan] = np.iinfo(np.uint64).max  # NaN last in both directions
        return out
    raise TypeError(f"unsupported key dtype {col.dtype}")


def stable_argsort(col: np.ndarray, descending: bool) -> np.ndarray:
    """Row order of transform_pyarrow.sort (transform_pyarrow.py:17-21) for one key column."""
    return np.argsort(encode_keys(col, descending), kind="stable")


def stable_argsort_multi(cols: Sequence[np.ndarray], descending: bool) -> np.ndarray:
    """Multi-column stable sort (first column is the major key), one direction for all columns
    (SortKey rejects mixed directions: _internal/planner/exchange/sort_task_spec.py:42-43)."""
    codes = [encode_keys(c, descending) for c in cols]
    return np.lexsort(list(reversed(codes)))  # lexsort: last key is primary; it is stable


def stable_argsort_bytes(col: np.ndarray, descending: bool) -> np.ndarray:
    """Row order of pyarrow's sort_indices on a fixed_size_binary / binary key (transform_pyarrow.py:17-21):
    lexicographic by unsigned bytes, stable in both directions. `col` is [n, width] uint8."""
    assert col.ndim == 2 and col.dtype == np.uint8
    keys = [(255 - col[:, j]) if descending else col[:, j] for j in range(col.shape[1])]
    return np.lexsort(list(reversed(keys)))  # last key is primary; stable


def take(table: Table, idx: np.ndarray) -> Table:
    """transform_pyarrow.take_table (transform_pyarrow.py:24-50)."""
    return {k: v[idx] for k, v in table.items()}


def num_rows(table: Table) -> int:
    for v in table.values():
        return int(v.shape[0])
    return 0


def concat(tables: Sequence[Table]) -> Table:
    """transform_pyarrow.concat (transform_pyarrow.py:154-266) for same-schema fixed-width tables."""
    tables = [t for t in tables if t]
    if not tables:
        return {}
    return {k: np.concatenate([t[k] for t in tables]) for k in tables[0]}


def slice_table(table: Table, lo: int, hi: int) -> Table:
    return {k: v[lo:hi] for k, v in table.items()}


# --------------------------------------------------------------------------------------
# sort: map / reduce (SortTaskSpec, sort_task_spec.py:124-147)
# --------------------------------------------------------------------------------------
def sort_table(table: Table, columns: Sequence[str], descending: bool) -> Table:
    """transform_pyarrow.sort (transform_pyarrow.py:17-21): sort_indices + take."""
    if num_rows(table) == 0:
        return table
    return take(table, stable_argsort_multi([table[c] for c in columns], descending))


def find_partition_index(table: Table, desired: Sequence, columns: Sequence[str], descending: bool) -> int:
    """_internal/util.py:719-758, restated with explicit counting instead of sorter= tricks."""
    left, right = 0, num_rows(table)
    for i in range(len(desired)):
        if left == right:
            return right
        col = table[columns[i]][left:right]
        d = desired[i]
        prevleft = left
        if descending:
            # block is descending: rows > d come first, then rows == d
            left = prevleft + int(np.count_nonzero(col > d))
            right = prevleft + int(np.count_nonzero(col >= d))
        else:
            left = prevleft + int(np.searchsorted(col, d, side="left"))
            right = prevleft + int(np.searchsorted(col, d, side="right"))
    return right if descending else left


def find_partitions(table: Table, boundaries: Sequence[Sequence], columns, descending) -> List[Table]:
    """_internal/util.py:761-779."""
    bounds = [find_partition_index(table, b, columns, descending) for b in boundaries]
    parts, last = [], 0
    for idx in bounds:
        parts.append(slice_table(table, last, idx))
        last = idx
    parts.append(slice_table(table, last, num_rows(table)))
    return parts


def sort_and_partition(table: Table, boundaries, columns, descending) -> List[Table]:
    """ArrowBlockAccessor.sort_and_partition (_internal/arrow_block.py:449-464)."""
    if num_rows(table) == 0:
        return [{} for _ in range(len(boundaries) + 1)]
    t = sort_table(table, columns, descending)
    if len(boundaries) == 0:
        return [t]
    return find_partitions(t, boundaries, columns, descending)


def merge_sorted_blocks(blocks: Sequence[Table], columns, descending) -> Table:
    """ArrowBlockAccessor.merge_sorted_blocks (_internal/arrow_block.py:554-567): drop empties,
    concat in mapper order, full stable re-sort (transform_pyarrow.concat_and_sort :269-276)."""
    blocks = [b for b in blocks if num_rows(b) > 0]
    if not blocks:
        return {}
    return sort_table(concat(blocks), columns, descending)


def boundaries_from_samples(samples: Table, columns: Sequence[str], num_reducers: int) -> List[tuple]:
    """The deterministic half of SortTaskSpec.sample_boundaries (sort_task_spec.py:176-199):
    lexsort the samples, per-column nearest-rank quantiles at linspace(0,1,R)[1:]."""
    if num_rows(samples) == 0:
        return [None] * (num_reducers - 1)
    vals = [samples[c] for c in columns]
    idx = np.lexsort(list(reversed(vals)))
    out_cols = []
    for v in vals:
        sv = v[idx]
        q = np.quantile(sv, np.linspace(0, 1, num_reducers), method="nearest")[1:]
        out_cols.append(list(q))
    return [tuple(c[i] for c in out_cols) for i in range(num_reducers - 1)]


def sort_dataset(blocks: Sequence[Table], key, descending: bool, boundaries: List[tuple]) -> List[Table]:
    """generate_sort_fn + PullBasedShuffleTaskScheduler (planner/sort.py:27-69,
    exchange/pull_based_shuffle_task_scheduler.py:86-124): `boundaries` ascending tuples;
    reversed for descending (sort.py:52-54); M map tasks x R reduce tasks."""
    columns = [key] if isinstance(key, str) else list(key)
    b = list(boundaries)
    if descending:
        b.reverse()
    mapped = [sort_and_partition(blk, b, columns, descending) for blk in blocks]
    R = len(b) + 1
    return [merge_sorted_blocks([m[j] for m in mapped], columns, descending) for j in range(R)]


def user_boundaries(bounds: Sequence[float]) -> List[tuple]:
    """SortKey boundary normalisation (sort_task_spec.py:46-54) + sort.py:50."""
    for item in bounds:
        if not isinstance(item, (int, float)):
            raise ValueError("The type of items in boundaries must be int or float.")
    return [(x,) for x in sorted(set(bounds))]


# --------------------------------------------------------------------------------------
# random_shuffle (ShuffleTaskSpec, shuffle_task_spec.py:42-132). The reference draws from
# numpy's MT19937 (RandomState.permutation / shuffle); the device path replaces the stream
# by a Feistel bijection, restated here bit for bit so that per-seed outputs can be checked
# exactly. The STRUCTURE (per-block permutation, ceil(n/R) slices, shuffled slice->reducer
# assignment, concat in mapper order, per-reducer permutation with the same seed) follows
# the reference.
# --------------------------------------------------------------------------------------
FEISTEL_ROUNDS = 6
_M64 = (1 << 64) - 1


def _splitmix64(x: int) -> int:
    x = (x + 0x9E3779B97F4A7C15) & _M64
    x = ((x ^ (x >> 30)) * 0xBF58476D1CE4E5B9) & _M64
    x = ((x ^ (x >> 27)) * 0x94D049BB133111EB) & _M64
    return x ^ (x >> 31)


def _fmix32(h: np.ndarray) -> np.ndarray:
    h = h.astype(np.uint32)
    h ^= h >> np.uint32(16)
    h = (h * np.uint32(0x85EBCA6B)).astype(np.uint32)
    h ^= h >> np.uint32(13)
    h = (h * np.uint32(0xC2B2AE35)).astype(np.uint32)
    h ^= h >> np.uint32(16)
    return h


def _feistel_spec(n: int, seed: int):
    bits = 0
    while n > 1 and (1 << bits) < n:
        bits += 1
    bits = max(bits, 2)
    lo = (bits + 1) // 2
    hi = bits - lo
    keys, st = [], seed & _M64
    for r in range(FEISTEL_ROUNDS):
        st = _splitmix64((st + r) & _M64)
        keys.append(st >> 32)
    return lo, (1 << lo) - 1, (1 << hi) - 1, keys


def feistel_permutation(n: int, seed: int, inverse: bool = False) -> np.ndarray:
    """pi(0..n) of csrc/shuffle.cu as a uint32 array: alternating Feistel rounds on the
    (hi | lo) split of a ceil(log2 n)-bit value, cycle-walked while the result is >= n."""
    if n <= 1:
        return np.arange(n, dtype=np.uint32)
    lo_bits, lo_mask, hi_mask, keys = _feistel_spec(n, seed)
    x = np.arange(n, dtype=np.uint32)
    pending = np.ones(n, dtype=bool)
    mul = np.uint32(0x9E3779B1)
    with np.errstate(over="ignore"):
        while pending.any():
            cur = x[pending]
            L = (cur >> np.uint32(lo_bits)) & np.uint32(hi_mask)
            R = cur & np.uint32(lo_mask)
            order = range(FEISTEL_ROUNDS - 1, -1, -1) if inverse else range(FEISTEL_ROUNDS)
            for r in order:
                if r % 2 == 0:
                    L = L ^ (_fmix32((R * mul + np.uint32(keys[r])).astype(np.uint32)) & np.uint32(hi_mask))
                else:
                    R = R ^ (_fmix32((L * mul + np.uint32(keys[r])).astype(np.uint32)) & np.uint32(lo_mask))
            nxt = ((L << np.uint32(lo_bits)) | R).astype(np.uint32)
            x[pending] = nxt
            pending[pending] = nxt >= n
    return x


def random_shuffle_block(table: Table, seed: Optional[int]) -> Table:
    """ArrowBlockAccessor.random_shuffle (_internal/arrow_block.py:249-255) with the Feistel stream."""
    n = num_rows(table)
    return take(table, feistel_permutation(n, 0 if seed is None else seed))


def shuffle_map(idx: int, table: Table, output_num_blocks: int, seed: Optional[int]) -> List[Table]:
    """ShuffleTaskSpec.map with random_shuffle=True (shuffle_task_spec.py:81-104)."""
    seed_i = (seed + idx) if seed is not None else idx
    block = random_shuffle_block(table, seed_i)
    n = num_rows(block)
    slice_sz = max(1, math.ceil(n / output_num_blocks))
    slices = [slice_table(block, i * slice_sz, (i + 1) * slice_sz) for i in range(output_num_blocks)]
    order = feistel_permutation(output_num_blocks, _splitmix64(seed_i & _M64))
    return [slices[int(j)] for j in order]


def shuffle_reduce(seed: Optional[int], mapper_outputs: Sequence[Table]) -> Table:
    """ShuffleTaskSpec.reduce (shuffle_task_spec.py:107-132)."""
    return random_shuffle_block(concat(mapper_outputs), seed)


def shuffle_dataset(blocks: Sequence[Table], seed: Optional[int], num_outputs: Optional[int] = None) -> List[Table]:
    """generate_random_shuffle_fn + pull-based scheduler (planner/random_shuffle.py:28-82)."""
    R = num_outputs or len(blocks)
    mapped = [shuffle_map(i, b, R, seed) for i, b in enumerate(blocks)]
    return [shuffle_reduce(seed, [m[j] for m in mapped]) for j in range(R)]


# --------------------------------------------------------------------------------------
# groupby-aggregate (SortAggregateTaskSpec, aggregate_task_spec.py:39-72;
# combine / aggregate_combined_blocks, _internal/arrow_block.py:466-680;
# Sum / Min / Max / Mean / Std / Count, aggregate/_aggregate.py:90-319).
# Aggregations are ("sum"|"min"|"max"|"mean"|"std"|"count", column or None).
# Non-null fixed-width columns only (the device block format), so the null wrappers
# (_internal/null_aggregate.py) reduce to has_data=True everywhere.
# --------------------------------------------------------------------------------------
def agg_name(kind: str, on: Optional[str]) -> str:
    return "count()" if kind == "count" else f"{kind}({on})"


def _group_runs(keys: np.ndarray) -> Tuple[np.ndarray, np.ndarray]:
    """Starts and lengths of runs of equal keys in a sorted column (arrow_block.py:490-515)."""
    n = len(keys)
    if n == 0:
        return np.zeros(0, np.int64), np.zeros(0, np.int64)
    head = np.ones(n, dtype=bool)
    head[1:] = keys[1:] != keys[:-1]
    starts = np.flatnonzero(head)
    lens = np.diff(np.append(starts, n))
    return starts, lens


def combine(table: Table, key: str, aggs: Sequence[Tuple[str, Optional[str]]]) -> Table:
    """ArrowBlockAccessor.combine on a block sorted by key: one partial-accumulator row per group.
    Accumulators (after the null wrapper): sum -> s; min/max -> m; count -> c;
    mean -> (s, c); std -> (M2, mean, c) (vectorized_std, _aggregate.py:284-296)."""
    keys = table[key]
    starts, lens = _group_runs(keys)
    out: Table = {key: keys[starts]}
    for i, (kind, on) in enumerate(aggs):
        name = f"acc{i}"
        if kind == "count":
            out[name] = lens.astype(np.int64)
            continue
        v = table[on]
        if kind in ("sum", "mean", "std"):
            s = np.add.reduceat(v, starts) if len(starts) else v[:0]
            if kind == "sum":
                out[name] = s
            elif kind == "mean":
                out[name + "_s"], out[name + "_c"] = s, lens.astype(np.int64)
            else:
                mean = s / lens
                d = v.astype(np.float64) - np.repeat(mean, lens)
                out[name + "_m2"] = np.add.reduceat(d * d, starts) if len(starts) else np.zeros(0)
                out[name + "_mean"], out[name + "_c"] = mean, lens.astype(np.int64)
        elif kind == "min":
            out[name] = np.minimum.reduceat(v, starts) if len(starts) else v[:0]
        elif kind == "max":
            out[name] = np.maximum.reduceat(v, starts) if len(starts) else v[:0]
        else:
            raise ValueError(kind)
    return out


def aggregate_combined_blocks(parts: Sequence[Table], key: str, aggs, finalize: bool = True, ddof: int = 1) -> Table:
    """ArrowBlockAccessor.aggregate_combined_blocks: merge partial rows with equal key in mapper
    order (heapq.merge is stable over its inputs), then finalize. Integer sums are merged as
    Python ints in the reference (unbounded); here object dtype keeps that exact."""
    parts = [p for p in parts if p and num_rows(p) > 0]
    if not parts:
        return {}
    allp = concat(parts)
    order = np.argsort(allp[key], kind="stable")
    allp = take(allp, order)
    starts, lens = _group_runs(allp[key])
    out: Table = {key: allp[key][starts]}
    for i, (kind, on) in enumerate(aggs):
        name = f"acc{i}"
        res = agg_name(kind, on)
        n_suffix, base = 2, res
        while res in out:  # _munge_conflict (arrow_block.py:550-552)
            res = f"{base}_{n_suffix}"
            n_suffix += 1
        if kind in ("sum", "count"):
            v = allp[name]
            if v.dtype.kind in "iu":
                tot = np.add.reduceat(v.astype(object), starts)
                out[res] = np.array([int(x) for x in tot], dtype=object) if (abs(tot) > 2**62).any() else tot.astype(np.int64)
            else:
                out[res] = np.add.reduceat(v, starts)
        elif kind == "min":
            out[res] = np.minimum.reduceat(allp[name], starts)
        elif kind == "max":
            out[res] = np.maximum.reduceat(allp[name], starts)
        elif kind == "mean":
            s = allp[name + "_s"]
            if s.dtype.kind in "iu":
                s = np.add.reduceat(s.astype(object), starts)
            else:
                s = np.add.reduceat(s, starts)
            c = np.add.reduceat(allp[name + "_c"], starts)
            out[res] = np.array([x / y for x, y in zip(s, c)], dtype=np.float64) if finalize else s
            if not finalize:
                out[res + "_c"] = c
        elif kind == "std":
            # pairwise merge in order (Std.merge, _aggregate.py:264-283)
            m2a, meana, ca = allp[name + "_m2"], allp[name + "_mean"], allp[name + "_c"]
            res_v = np.empty(len(starts))
            for g, (s0, ln) in enumerate(zip(starts, lens)):
                M2, mean, cnt = m2a[s0], meana[s0], ca[s0]
                for j in range(s0 + 1, s0 + ln):
                    delta = meana[j] - mean
                    tot = cnt + ca[j]
                    mean = (mean * cnt + meana[j] * ca[j]) / tot
                    M2 = M2 + m2a[j] + (delta**2) * cnt * ca[j] / tot
                    cnt = tot
                res_v[g] = 0.0 if cnt < 2 else math.sqrt(M2 / (cnt - ddof))
            out[res] = res_v
        else:
            raise ValueError(kind)
    return out


def aggregate_dataset(blocks: Sequence[Table], key: str, aggs, boundaries: List[tuple]) -> List[Table]:
    """generate_aggregate_fn + pull-based scheduler (planner/aggregate.py:35-83):
    map = prune columns, ascending sort_and_partition, combine each partition; reduce = merge."""
    used = {key} | {on for _, on in aggs if on is not None}
    R = len(boundaries) + 1
    mapped = []
    for blk in blocks:
        blk = {k: v for k, v in blk.items() if k in used}
        parts = sort_and_partition(blk, boundaries, [key], False)
        mapped.append([combine(p, key, aggs) if num_rows(p) else {} for p in parts])
    return [aggregate_combined_blocks([m[j] for m in mapped], key, aggs) for j in range(R)]


# --------------------------------------------------------------------------------------
# Quantile (aggregate/_aggregate.py:360-431). The reference's accumulator is the list of every
# value of the group (merge = list concatenation, :378-398), so the result does not depend on how
# rows were split into blocks; finalize is `percentile` (:406-417).
# --------------------------------------------------------------------------------------
def percentile(values: Sequence, q: float):
    """_aggregate.py:406-417, step by step (Python floats, Python round)."""
    if len(values) == 0:
        return None
    vals = sorted(values)
    k = (len(vals) - 1) * q
    f = math.floor(k)
    c = math.ceil(k)
    if f == c:
        return vals[int(k)]
    d0 = vals[int(f)] * (c - k)
    d1 = vals[int(c)] * (k - f)
    return round(d0 + d1, 5)


def quantile_dataset(blocks: Sequence[Table], key: Optional[str], on: str, q: float = 0.5) -> Table:
    """groupby(key).aggregate(Quantile(on, q)) over the whole dataset: keys ascending, one value per
    group (Python int/float as the reference returns them). key=None: one global group."""
    allp = concat([b for b in blocks if b and num_rows(b) > 0])
    v = allp[on]
    if key is None:
        return {f"quantile({on})": np.array([percentile(v.tolist(), q)], dtype=object)}
    order = np.argsort(allp[key], kind="stable")
    ks, vs = allp[key][order], v[order]
    starts, lens = _group_runs(ks)
    out = [percentile(vs[s:s + ln].tolist(), q) for s, ln in zip(starts, lens)]
    return {key: ks[starts], f"quantile({on})": np.array(out, dtype=object)}


def key_boundaries(key_columns: Sequence[np.ndarray]) -> List[int]:
    """End position of every group of a block already sorted by `key_columns`: GroupedData.map_groups'
    get_key_boundaries (grouped_data.py:181-200: searchsorted(..., side="right") over the key, for several key
    columns over `_MultiColumnSortedKey` tuples, :13-33) - a group ends where any key column changes."""
    n = int(key_columns[0].shape[0])
    if n == 0:
        return []
    change = np.zeros(n - 1, dtype=bool)
    for col in key_columns:
        change |= col[1:] != col[:-1]
    return (np.flatnonzero(change) + 1).tolist() + [n]


def null_group_aggregate(keys: Optional[np.ndarray], values: np.ndarray, is_null: np.ndarray, kind: str, ignore_nulls: bool,
                         ddof: int = 1) -> List[Optional[float]]:
    """One aggregation of a nullable value column per group (ascending keys; `keys=None`: one global group) under the
    reference's null policy (_internal/null_aggregate.py:11-17): ignore_nulls -> aggregate the non-null values, None
    if there are none; otherwise one null makes the group's result None. Mean / Std count non-null values only
    (arrow_block.py:384-386 + aggregate/_aggregate.py:205-302)."""
    groups = [None] if keys is None else np.unique(keys).tolist()
    out: List[Optional[float]] = []
    for g in groups:
        sel = np.ones(len(values), dtype=bool) if g is None else (keys == g)
        v, nul = values[sel], is_null[sel]
        if nul.all() or (nul.any() and not ignore_nulls) or len(v) == 0:
            out.append(None)
            continue
        v = v[~nul]
        if kind == "sum":
            out.append(v.sum().item())
        elif kind == "min":
            out.append(v.min().item())
        elif kind == "max":
            out.append(v.max().item())
        elif kind == "abs_max":
            out.append(np.abs(v).max().item())
        elif kind == "mean":
            out.append(float(v.sum()) / len(v))
        elif kind == "std":
            m = float(v.astype(np.float64).sum()) / len(v)
            m2 = float(((v.astype(np.float64) - m) ** 2).sum())
            out.append(0.0 if len(v) - ddof <= 0 else math.sqrt(m2 / (len(v) - ddof)))
        else:
            raise ValueError(kind)
    return out
