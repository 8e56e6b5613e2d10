"""Aggregations: the AggregateFn library and its device implementation.

API mirror of python/ray/data/aggregate/_aggregate.py:20-319 (AggregateFn, Count, Sum, Min, Max,
Mean, Std; same constructor arguments, same output column names `sum(x)`, `mean(x)`, `count()`,
same `_2` suffix for duplicate names, arrow_block.py:537-552).

The reference evaluates an aggregation with Python callables (init / accumulate_block / merge /
finalize) per group. On the device a built-in aggregation is instead a (kind, column) pair that
selects outputs of the segmented-reduce kernel, and a partially combined block (the output of
`combine`, the input of `aggregate_combined_blocks`) stores each accumulator in flat columns:

  count          name                      int64
  sum/min/max    name                      column dtype
  mean           name (= sum), name::count
  std            name (= M2),  name::sum, name::count      (M2 about the partial's own mean)
  quantile       (holistic) the accumulator is the list of every value, as in the reference: when a
                 Quantile is among the aggregations the partial block is the pruned rows themselves
                 (sorted by key) and every aggregation is evaluated at the final merge

User-defined AggregateFn callables cannot run on the device and raise NotImplementedError.

Nulls (the reference's policy, _internal/null_aggregate.py:11-17, wrappers :28-250): a nullable value column carries
a hidden validity column (block.py). Per group, with ignore_nulls=True the aggregation covers the non-null values and
is null when there are none; with ignore_nulls=False one null makes it null. Count() counts rows, nulls included;
Mean's and Std's denominators count non-null values (pyarrow's count, arrow_block.py:384-386). On the device a null row
enters a reduction as that reduction's identity (`b200x_mask_values`); the partial accumulators of a nullable column
gain `name::nvalid` (non-null values) and `name::rows` (rows), which are summed at the merge, and the finished column
gets its own hidden validity column (`b200x_group_validity`). Quantile over a nullable column is not implemented.
"""
from __future__ import annotations

from typing import Callable, Dict, List, Optional, Sequence, Tuple

import torch

from . import ops
from .block import DeviceBlock, _contig, valid_name

SEP = "::"


class AggregateFn:
    def __init__(self, init: Callable = None, merge: Callable = None, accumulate_row: Callable = None,
                 accumulate_block: Callable = None, finalize: Callable = lambda a: a, name: Optional[str] = None):
        if (accumulate_row is None and accumulate_block is None) or (
            accumulate_row is not None and accumulate_block is not None
        ):
            raise ValueError("Exactly one of accumulate_row or accumulate_block must be provided.")
        self.init, self.merge, self.finalize, self.name = init, merge, finalize, name
        self.accumulate_row, self.accumulate_block = accumulate_row, accumulate_block
        self._device_kind: Optional[str] = None

    def _validate(self, schema) -> None:
        pass


class _AggregateOnKeyBase(AggregateFn):
    def _setup(self, kind: str, on: Optional[str], alias_name: Optional[str], ignore_nulls: bool = True):
        self._key_fn = on
        self._device_kind = kind
        self._ignore_nulls = ignore_nulls
        self.name = alias_name if alias_name else f"{kind}({str(on)})"
        self._rs_name = self.name
        self.init = self.merge = self.accumulate_block = self.accumulate_row = None
        self.finalize = lambda a: a

    def _validate(self, schema) -> None:
        from .sort_task_spec import SortKey

        SortKey(self._key_fn).validate_schema(schema)


class Count(AggregateFn):
    """Rows per group (counts every row: _aggregate.py:90-102)."""

    def __init__(self):
        self.name = "count()"
        self._device_kind = "count"
        self._key_fn = None
        self.init = self.merge = self.accumulate_block = self.accumulate_row = None
        self.finalize = lambda a: a


class Sum(_AggregateOnKeyBase):
    def __init__(self, on: Optional[str] = None, ignore_nulls: bool = True, alias_name: Optional[str] = None):
        self._setup("sum", on, alias_name, ignore_nulls)


class Min(_AggregateOnKeyBase):
    def __init__(self, on: Optional[str] = None, ignore_nulls: bool = True, alias_name: Optional[str] = None):
        self._setup("min", on, alias_name, ignore_nulls)


class Max(_AggregateOnKeyBase):
    def __init__(self, on: Optional[str] = None, ignore_nulls: bool = True, alias_name: Optional[str] = None):
        self._setup("max", on, alias_name, ignore_nulls)


class Mean(_AggregateOnKeyBase):
    def __init__(self, on: Optional[str] = None, ignore_nulls: bool = True, alias_name: Optional[str] = None):
        self._setup("mean", on, alias_name, ignore_nulls)


class Std(_AggregateOnKeyBase):
    def __init__(self, on: Optional[str] = None, ddof: int = 1, ignore_nulls: bool = True,
                 alias_name: Optional[str] = None):
        self._setup("std", on, alias_name, ignore_nulls)
        self._ddof = ddof


class AbsMax(_AggregateOnKeyBase):
    """Largest absolute value per group (_aggregate.py:323-357)."""

    def __init__(self, on: Optional[str] = None, ignore_nulls: bool = True, alias_name: Optional[str] = None):
        self._setup("abs_max", on, alias_name, ignore_nulls)
        self._device_kind = "absmax"


class Quantile(_AggregateOnKeyBase):
    """q-quantile per group by linear interpolation, rounded to 5 decimals between rows
    (_aggregate.py:360-431; the result column is float64)."""

    def __init__(self, on: Optional[str] = None, q: float = 0.5, ignore_nulls: bool = True,
                 alias_name: Optional[str] = None):
        self._setup("quantile", on, alias_name, ignore_nulls)
        self._q = q


def _holistic(aggs: Sequence[AggregateFn]) -> bool:
    return any(getattr(a, "_device_kind", None) == "quantile" for a in aggs)


def _device_kind(agg: AggregateFn) -> str:
    kind = getattr(agg, "_device_kind", None)
    if kind is None:
        raise NotImplementedError(
            "user-defined AggregateFn callables cannot run on device blocks; use Count/Sum/Min/Max/Mean/Std/AbsMax/Quantile")
    return kind


def resolved_names(aggs: Sequence[AggregateFn]) -> List[str]:
    """Output column names with the reference's conflict rule (arrow_block.py:537-552)."""
    count: Dict[str, int] = {}
    out = []
    for agg in aggs:
        name = agg.name
        c = count.get(name, 0)
        out.append(name if c == 0 else f"{name}_{c + 1}")
        count[name] = c + 1
    return out


def _value_column(block: DeviceBlock, on: str) -> torch.Tensor:
    if on not in block.columns:
        raise ValueError(f"The column '{on}' does not exist in the schema '{block.schema()}'.")
    v = _contig(block.columns[on])
    if v.dim() != 1:
        raise TypeError(f"cannot aggregate the multi-byte payload column {on!r}")
    if v.dtype in (torch.int32, torch.int16, torch.int8, torch.uint8):
        raise TypeError(f"aggregations take int64 or float64 columns; cast {on!r} ({v.dtype}) first")
    if v.dtype == torch.float32:
        raise TypeError(f"aggregations take int64 or float64 columns; cast {on!r} (float32) first")
    return v


def _groups_of_sorted(block: DeviceBlock, key: Optional[str]):
    """(group_starts, num_groups tensor, G) of a block already sorted by `key` (None: one group)."""
    n = block.num_rows
    dev = block.device
    if key is None:
        starts = torch.zeros(1, dtype=torch.int32, device=dev)
        ng = torch.ones(1, dtype=torch.int32, device=dev)
        return starts, ng, 1
    codes = ops.encode_keys(_contig(block.columns[key]), False)
    starts, ng = ops.group_boundaries(codes)
    return starts, ng, int(ng.item())


def _reduce_partials(kind: str, agg: AggregateFn, name: str, src: DeviceBlock, starts, ng, G: int, perm, n: int,
                     raw: bool, cache: Optional[dict] = None) -> Dict[str, torch.Tensor]:
    """Accumulator columns of one aggregation for G groups.
    raw=True: `src` holds data rows (combine). raw=False: `src` holds partial rows (merge).
    `cache` (raw only) shares one pass over a value column between aggregations of that column:
    the first request computes sum, min and max of the column in a single launch."""
    out: Dict[str, torch.Tensor] = {}
    if raw and cache is not None:
        def counts():
            if "count" not in cache:
                cache["count"] = ops.group_counts(starts, ng, G, n)
            return cache["count"]

        def validity(on):
            m = src.columns.get(valid_name(on))
            return None if m is None else _contig(m)

        def nvalid(on):
            """Non-null values per group (= rows per group for a column without nulls)."""
            m = validity(on)
            if m is None:
                return counts()
            if ("nvalid", on) not in cache:
                cache[("valid64", on)] = ops.valid_as_int64(m)
                cache[("nvalid", on)] = ops.segmented_reduce(starts, ng, G, perm, cache[("valid64", on)], n)[0]
            return cache[("nvalid", on)]

        def smm(on):
            if ("smm", on) not in cache:
                v, m = _value_column(src, on), validity(on)
                if m is None:
                    cache[("smm", on)] = ops.segmented_reduce(starts, ng, G, perm, v, n, True, True, True)
                else:   # a null row enters every reduction as that reduction's identity
                    s_ = ops.segmented_reduce(starts, ng, G, perm, ops.mask_values(v, m, 0), n)[0]
                    mn = ops.segmented_reduce(starts, ng, G, perm, ops.mask_values(v, m, 1), n, False, True, False)[1]
                    mx = ops.segmented_reduce(starts, ng, G, perm, ops.mask_values(v, m, 2), n, False, False, True)[2]
                    cache[("smm", on)] = (s_, mn, mx)
            return cache[("smm", on)]

        def with_null_counts(acc, on):
            """Accumulators of a nullable column also say how many of the group's rows were non-null / there at all."""
            if validity(on) is not None:
                acc[name + SEP + "nvalid"], acc[name + SEP + "rows"] = nvalid(on), counts()
            return acc

        if kind == "count":
            return {name: counts()}
        if kind == "quantile":
            on = agg._key_fn
            if validity(on) is not None:
                raise NotImplementedError(f"Quantile over the nullable column {on!r} is not implemented on device blocks")
            if ("qperm", on) not in cache:  # rows ordered by (keys, value): same group positions as `starts`
                cache[("qperm", on)] = _argsort_keys(src, list(cache.get("keys", [])) + [on], want_codes=False)[0]
            return {name: ops.group_quantile(starts, ng, G, cache[("qperm", on)], _value_column(src, on), n, agg._q)}
        on = agg._key_fn
        if kind in ("sum", "min", "max"):
            return with_null_counts({name: smm(on)[("sum", "min", "max").index(kind)]}, on)
        if kind == "absmax":
            t = smm(on)
            return with_null_counts({name: ops.absmax_groups(t[1], t[2])}, on)
        if kind == "mean":      # the denominator counts the non-null values
            return with_null_counts({name: smm(on)[0], name + SEP + "count": nvalid(on)}, on)
        if kind == "std":
            s_, cnt = smm(on)[0], nvalid(on)
            if ("mean", on) not in cache:
                cache[("mean", on)] = ops.finalize_groups(s_, cnt, 0)
            if validity(on) is None:
                m2 = ops.segmented_reduce(starts, ng, G, perm, _value_column(src, on), n, center=cache[("mean", on)])[0]
            else:               # null rows get weight 0 in the sum of squared differences
                nvalid(on)
                m2 = ops.segmented_reduce(starts, ng, G, perm, ops.mask_values(_value_column(src, on), validity(on), 0), n,
                                          center=cache[("mean", on)], weights=cache[("valid64", on)])[0]
            return with_null_counts({name: m2, name + SEP + "sum": s_, name + SEP + "count": cnt}, on)
    if kind == "count":
        if raw:
            out[name] = ops.group_counts(starts, ng, G, n)
        else:
            out[name] = ops.segmented_reduce(starts, ng, G, perm, _contig(src.columns[name]), n)[0]
        return out
    if not raw and name + SEP + "nvalid" in src.columns:    # partials of a nullable column: both counts add up
        for extra in ("nvalid", "rows"):
            out[name + SEP + extra] = ops.segmented_reduce(starts, ng, G, perm, _contig(src.columns[name + SEP + extra]), n)[0]
    if raw and src.columns.get(valid_name(agg._key_fn)) is not None:
        raise NotImplementedError("nullable value columns are reduced through the shared-pass path (cache) only")
    col = _value_column(src, agg._key_fn) if raw else _contig(src.columns[name])
    if kind == "sum":
        out[name] = ops.segmented_reduce(starts, ng, G, perm, col, n)[0]
    elif kind == "min":
        out[name] = ops.segmented_reduce(starts, ng, G, perm, col, n, want_sum=False, want_min=True)[1]
    elif kind in ("max", "absmax"):
        if raw and kind == "absmax":
            t = ops.segmented_reduce(starts, ng, G, perm, col, n, False, True, True)
            out[name] = ops.absmax_groups(t[1], t[2])
        else:  # partial abs-max values are >= 0: merging them is a max
            out[name] = ops.segmented_reduce(starts, ng, G, perm, col, n, want_sum=False, want_max=True)[2]
    elif kind == "mean":
        out[name] = ops.segmented_reduce(starts, ng, G, perm, col, n)[0]
        if raw:
            out[name + SEP + "count"] = ops.group_counts(starts, ng, G, n)
        else:
            out[name + SEP + "count"] = ops.segmented_reduce(starts, ng, G, perm,
                                                             _contig(src.columns[name + SEP + "count"]), n)[0]
    elif kind == "std":
        if raw:
            s = ops.segmented_reduce(starts, ng, G, perm, col, n)[0]
            cnt = ops.group_counts(starts, ng, G, n)
            mean = ops.finalize_groups(s, cnt, 0)
            m2 = ops.segmented_reduce(starts, ng, G, perm, col, n, center=mean)[0]
        else:
            # Chan et al. merge of (M2, sum, count) partials (Std.merge, _aggregate.py:264-283):
            # M2 = sum_p M2_p + sum_p count_p * (mean_p - mean)^2
            ps, pc = _contig(src.columns[name + SEP + "sum"]), _contig(src.columns[name + SEP + "count"])
            s = ops.segmented_reduce(starts, ng, G, perm, ps, n)[0]
            cnt = ops.segmented_reduce(starts, ng, G, perm, pc, n)[0]
            mean = ops.finalize_groups(s, cnt, 0)
            part_mean = ops.finalize_groups(ps, pc, 0)
            m2 = ops.segmented_reduce(starts, ng, G, perm, part_mean, n, center=mean, weights=pc, addend=col)[0]
        out[name], out[name + SEP + "sum"], out[name + SEP + "count"] = m2, s, cnt
    else:
        raise ValueError(kind)
    return out


def _finalize(kind: str, agg: AggregateFn, name: str, acc: Dict[str, torch.Tensor]) -> torch.Tensor:
    if kind in ("count", "sum", "min", "max", "absmax", "quantile"):
        return acc[name]
    if kind == "mean":
        return ops.finalize_groups(acc[name], acc[name + SEP + "count"], 0)
    if kind == "std":
        return ops.finalize_groups(acc[name], acc[name + SEP + "count"], 1, getattr(agg, "_ddof", 1))
    raise ValueError(kind)


def _finalize_validity(agg: AggregateFn, name: str, acc: Dict[str, torch.Tensor], hidden: Dict[str, torch.Tensor]) -> None:
    """The finished column of a nullable input is nullable itself: null_aggregate.py:11-17 decides per group."""
    if name + SEP + "nvalid" in acc:
        hidden[valid_name(name)] = ops.group_validity(acc[name + SEP + "nvalid"], acc[name + SEP + "rows"],
                                                      bool(getattr(agg, "_ignore_nulls", True)))


def _check_key(key):
    """None (global aggregation), one column name, or a list of >= 2 column names (at most 4)."""
    if isinstance(key, list):
        if len(key) == 1:
            return key[0]
        if len(key) == 0:
            return None
        if len(key) > 4:
            raise NotImplementedError("grouping by more than four key columns is not implemented on device blocks")
        return list(key)
    return key


def _key_list(key) -> List[str]:
    return [] if key is None else ([key] if isinstance(key, str) else list(key))


def _sorted_key_codes(block: DeviceBlock, keys: List[str], perm) -> List[torch.Tensor]:
    """Ascending codes of every key column, in the sorted row order."""
    out = []
    for k in keys:
        col = _contig(block.columns[k])
        if perm is not None:
            col = ops.gather(col, perm)
        out.append(ops.encode_keys(col, False))
    return out


def _argsort_keys(block: DeviceBlock, keys: List[str], want_codes: bool):
    """Stable multi-column argsort (last key column first). Returns (perm, codes of the first column or None)."""
    perm = codes = None
    for i, k in enumerate(reversed(keys)):
        last = i == len(keys) - 1
        perm, codes, _ = ops.argsort(_contig(block.columns[k]), False, perm_in=perm, want_codes=want_codes and last)
    return perm, codes


def _refuse_nullable_keys(block: DeviceBlock, keys: Sequence[str]) -> None:
    for k in keys:
        if valid_name(k) in block.columns:
            raise NotImplementedError(f"groupby key column {k!r} is nullable: null keys are not supported on device blocks")


def aggregate_sorted(block: DeviceBlock, key, aggs: Sequence[AggregateFn], perm, codes, finalize: bool, n: int) -> DeviceBlock:
    """One-shot groupby over rows visited in sorted order `perm`. `codes`: the ascending codes of the
    (single) key column in that order, or None to derive them from the block."""
    key = _check_key(key)
    keys = _key_list(key)
    _refuse_nullable_keys(block, keys)
    dev = block.device
    names = resolved_names(aggs)
    if not finalize and _holistic(aggs):
        for agg in aggs:
            if getattr(agg, "_key_fn", None) is not None:
                _value_column(block, agg._key_fn)
        if perm is None:
            return block
        from .block import DeviceBlockAccessor

        return DeviceBlockAccessor(block).take(perm)
    if not keys:
        starts = torch.zeros(1, dtype=torch.int32, device=dev)
        ng = torch.ones(1, dtype=torch.int32, device=dev)
        G = 1
    else:
        if len(keys) == 1 and codes is not None:
            codes_list = [codes]
        else:
            codes_list = _sorted_key_codes(block, keys, perm)
        starts, ng = ops.group_boundaries_multi(codes_list)
        G = int(ng.item())
    cols: Dict[str, torch.Tensor] = {}
    if keys:
        first_rows = ops.gather(perm, starts[:G].contiguous()) if perm is not None else starts[:G].contiguous()
        for k in keys:
            cols[k] = ops.gather(_contig(block.columns[k]), first_rows)
    cache: dict = {"keys": keys}
    hidden: Dict[str, torch.Tensor] = {}
    for agg, name in zip(aggs, names):
        kind = _device_kind(agg)
        acc = _reduce_partials(kind, agg, name, block, starts, ng, G, perm, n, raw=True, cache=cache)
        if finalize:
            cols[name] = _finalize(kind, agg, name, acc)
            _finalize_validity(agg, name, acc, hidden)
        else:
            cols.update(acc)
    cols.update(hidden)
    return DeviceBlock(cols, num_rows=G, device=dev)


def combine_block(block: DeviceBlock, key, aggs: Sequence[AggregateFn], presorted: bool = True) -> DeviceBlock:
    """ArrowBlockAccessor.combine (arrow_block.py:466-548): one partial-accumulator row per group of
    a block sorted by key."""
    n = block.num_rows
    if n == 0:
        return DeviceBlock.empty(block.device)
    return aggregate_sorted(block, key, aggs, None, None, finalize=False, n=n)


def merge_combined_blocks(blocks: Sequence[DeviceBlock], key, aggs: Sequence[AggregateFn], finalize: bool) -> DeviceBlock:
    """ArrowBlockAccessor.aggregate_combined_blocks (arrow_block.py:569-680): merge partial rows with
    equal keys (mapper order within a key, as heapq.merge gives), optionally finalize."""
    key = _check_key(key)
    keys = _key_list(key)
    blocks = [b for b in blocks if b.num_rows > 0]
    if not blocks:
        return DeviceBlock.empty()
    allp = DeviceBlock.concat(blocks)
    n = allp.num_rows
    dev = allp.device
    names = resolved_names(aggs)
    if _holistic(aggs):  # the partials are rows: aggregate them as data (stable by key: mapper order within a key)
        perm = codes = None
        if keys:
            perm, codes = _argsort_keys(allp, keys, want_codes=len(keys) == 1)
        return aggregate_sorted(allp, key, aggs, perm, codes, finalize=finalize, n=n)
    if not keys:
        perm = None
        starts = torch.zeros(1, dtype=torch.int32, device=dev)
        ng = torch.ones(1, dtype=torch.int32, device=dev)
        G = 1
    else:
        perm, codes = _argsort_keys(allp, keys, want_codes=len(keys) == 1)
        codes_list = [codes] if len(keys) == 1 else _sorted_key_codes(allp, keys, perm)
        starts, ng = ops.group_boundaries_multi(codes_list)
        G = int(ng.item())
    cols: Dict[str, torch.Tensor] = {}
    if keys:
        first_rows = ops.gather(perm, starts[:G].contiguous())
        for k in keys:
            cols[k] = ops.gather(_contig(allp.columns[k]), first_rows)
    hidden: Dict[str, torch.Tensor] = {}
    for agg, name in zip(aggs, names):
        kind = _device_kind(agg)
        acc = _reduce_partials(kind, agg, name, allp, starts, ng, G, perm, n, raw=False)
        if finalize:
            cols[name] = _finalize(kind, agg, name, acc)
            _finalize_validity(agg, name, acc, hidden)
        else:
            cols.update(acc)
    cols.update(hidden)
    return DeviceBlock(cols, num_rows=G, device=dev)


def _global_reduce(col: torch.Tensor, want: str, center: Optional[float] = None):
    """Whole-column sum / min / max / sum of squared differences, returned as a Python scalar
    (BlockAccessor.sum/min/max/sum_of_squared_diffs_from_mean, arrow_block.py:404-447)."""
    col = _contig(col)
    n = col.numel()
    dev = col.device
    starts = torch.zeros(1, dtype=torch.int32, device=dev)
    ng = torch.ones(1, dtype=torch.int32, device=dev)
    if want == "m2":
        c = torch.full((1,), float(center), dtype=torch.float64, device=dev)
        return ops.segmented_reduce(starts, ng, 1, None, col, n, center=c)[0].item()
    s, mn, mx = ops.segmented_reduce(starts, ng, 1, None, col, n, want_sum=want == "sum", want_min=want == "min",
                                     want_max=want == "max")
    return {"sum": s, "min": mn, "max": mx}[want].item()


# ------------------------------------------------------------------ dense-range fast path
DENSE_MAX_RANGE = 1 << 22      # table entries: count + one to three int64 tables of this length stay inside the 126 MB L2
DENSE_MIN_ROWS = 1 << 16
_DENSE_KINDS = ("count", "sum", "min", "max", "mean", "absmax")


def dense_aggregate(block: DeviceBlock, key, aggs: Sequence[AggregateFn], finalize: bool) -> Optional[DeviceBlock]:
    """groupby(key).aggregate(aggs) without a sort, or None when it does not apply: one integer key column whose
    codes span at most DENSE_MAX_RANGE values (ids, categories, dates), and only Count / Sum / Min / Max / Mean /
    AbsMax over int64 columns - order-independent integer arithmetic, so the rows can be added straight into
    accumulator tables indexed by `key - min(key)` (csrc/groupby.cu, "Dense-range aggregation") and the result is
    bit-identical to the sorted path. Costs one read of the key column to find its range (and the decision)."""
    if not isinstance(key, str) or block.num_rows < DENSE_MIN_ROWS or valid_name(key) in block.columns:
        return None
    kcol = block.columns[key]
    if kcol.dim() != 1 or kcol.dtype.is_floating_point or kcol.dtype not in ops._KEY_TYPES:
        return None
    kinds = []
    for agg in aggs:
        kind = getattr(agg, "_device_kind", None)
        if kind not in _DENSE_KINDS:
            return None
        if kind != "count":
            v = block.columns.get(agg._key_fn)
            if v is None or v.dim() != 1 or v.dtype != torch.int64 or valid_name(agg._key_fn) in block.columns:
                return None
        kinds.append(kind)
    kcol = _contig(kcol)
    lo, hi = ops.key_code_range(kcol)
    rng = hi - lo + 1
    if rng > DENSE_MAX_RANGE or rng > 4 * block.num_rows:
        return None
    dev = block.device
    count = torch.empty(rng, dtype=torch.int64, device=dev)
    need: Dict[str, set] = {}
    for agg, kind in zip(aggs, kinds):
        if kind != "count":
            need.setdefault(agg._key_fn, set()).update({"sum": {"sum"}, "mean": {"sum"}, "min": {"min"}, "max": {"max"},
                                                        "absmax": {"min", "max"}}[kind])
    tabs: Dict[Tuple[str, str], torch.Tensor] = {}
    first = True
    for on, wants in need.items():
        s_, mn, mx = ops.dense_group_accumulate(kcol, lo, rng, _contig(block.columns[on]), count if first else None,
                                                "sum" in wants, "min" in wants, "max" in wants)
        first = False
        for what, t in (("sum", s_), ("min", mn), ("max", mx)):
            if t is not None:
                tabs[(on, what)] = t
    if first:   # only counts were asked for
        ops.dense_group_accumulate(kcol, lo, rng, None, count, False, False, False)
    order = list(tabs)
    keys_out, count_out, outs, G = ops.dense_group_compact(lo, rng, kcol.dtype, count, [tabs[k] for k in order],
                                                           min(rng, block.num_rows))
    got = dict(zip(order, outs))
    cols: Dict[str, torch.Tensor] = {key: keys_out}
    for agg, kind, name in zip(aggs, kinds, resolved_names(aggs)):
        on = agg._key_fn
        if kind == "count":
            cols[name] = count_out
        elif kind in ("sum", "min", "max"):
            cols[name] = got[(on, kind)]
        elif kind == "absmax":
            cols[name] = ops.absmax_groups(got[(on, "min")], got[(on, "max")])
        elif finalize:   # mean
            cols[name] = ops.finalize_groups(got[(on, "sum")], count_out, 0)
        else:
            cols[name], cols[name + SEP + "count"] = got[(on, "sum")], count_out
    return DeviceBlock(cols, num_rows=G, device=dev)
