"""Dataset / GroupedData over device blocks: the user-facing surface of the path.

Keeps the ray.data entry points of the all-to-all path
  Dataset.sort            python/ray/data/dataset.py:2148-2209
  Dataset.random_shuffle  dataset.py:1001-1047
  Dataset.groupby         dataset.py:1788-1836
  Dataset.aggregate / sum / min / max / mean / std   dataset.py:1882-2146, :4805-4841
  Dataset.repartition     dataset.py:944-999
  Dataset.unique          dataset.py:1839-1879
  GroupedData             python/ray/data/grouped_data.py:36-460
plus the constructors and consumers the reference's own tests use to observe the path
(SURVEY.md §8b). Each all-to-all method builds the logical operator, plans it with
plan_all_to_all_op and runs the resulting AllToAllOperator over the dataset's blocks.
Execution is eager (the reference is lazy; results are the same). D2H copies happen only in
the consumer calls (take / iter_rows / to_pandas / ...).
"""
from __future__ import annotations

import builtins
import itertools
from typing import Any, Callable, Dict, Iterator, List, Optional, Tuple, Union

import numpy as np
import torch

from .aggregate import AggregateFn, Count, Max, Mean, Min, Std, Sum
from .block import Block, BlockAccessor, BlockExecStats, BlockMetadata, DeviceBlock, DeviceBlockAccessor, default_device
from .context import DataContext
from .interfaces import RefBundle, StatsDict
from .planner import Aggregate, RandomizeBlocks, RandomShuffle, Repartition, Sort, plan_all_to_all_op
from .sort_task_spec import SortKey

_brange = builtins.range  # this module defines the ray.data-style `range()` constructor below


class Dataset:
    def __init__(self, blocks: List[DeviceBlock], stats: Optional[List[Tuple[str, StatsDict]]] = None):
        self._blocks: List[DeviceBlock] = list(blocks)
        self._stats: List[Tuple[str, StatsDict]] = list(stats or [])

    # ------------------------------------------------------------------ internals
    def _bundles(self) -> List[RefBundle]:
        return [RefBundle([(b, DeviceBlockAccessor(b).get_metadata(None, None))], owns_blocks=True)
                for b in self._blocks]

    def _run_all_to_all(self, logical_op) -> "Dataset":
        op = plan_all_to_all_op(logical_op)
        op.initialize_sub_progress_bars()
        for bundle in self._bundles():
            op.add_input(bundle, 0)
        op.all_inputs_done()
        out_blocks = []
        while op.has_next():
            for block, _ in op.get_next().blocks:
                out_blocks.append(block)
        op.close_sub_progress_bars()
        return Dataset(out_blocks, self._stats + [(logical_op.name, op.get_stats())])

    # ------------------------------------------------------------------ all-to-all ops
    def sort(self, key: Union[str, List[str], None] = None, descending: Union[bool, List[bool]] = False,
             boundaries: List[Union[int, float]] = None) -> "Dataset":
        sort_key = SortKey(key, descending, boundaries)
        return self._run_all_to_all(Sort(None, sort_key=sort_key))

    def random_shuffle(self, *, seed: Optional[int] = None, num_blocks: Optional[int] = None,
                       **ray_remote_args) -> "Dataset":
        if num_blocks is not None:
            raise DeprecationWarning(
                "`num_blocks` parameter is deprecated in Ray 2.9. random_shuffle() does not support to change "
                "the number of output blocks. Use repartition() instead.")
        return self._run_all_to_all(RandomShuffle(None, seed=seed, ray_remote_args=ray_remote_args))

    def randomize_block_order(self, *, seed: Optional[int] = None) -> "Dataset":
        """Shuffle the blocks (not the rows inside them); dataset.py:1050-1083."""
        return self._run_all_to_all(RandomizeBlocks(None, seed=seed))

    def repartition(self, num_blocks: int, *, shuffle: bool = False) -> "Dataset":
        return self._run_all_to_all(Repartition(None, num_outputs=num_blocks, shuffle=shuffle))

    def groupby(self, key: Union[str, List[str], None]) -> "GroupedData":
        if key is not None:
            SortKey(key).validate_schema(self.schema())
        return GroupedData(self, key)

    def unique(self, column: str) -> List[Any]:
        ds = self.select_columns([column]).groupby(column).count()
        return [r[column] for r in ds.take_all()]

    def aggregate(self, *aggs: AggregateFn) -> Union[Any, Dict[str, Any]]:
        ret = self.groupby(None).aggregate(*aggs).take(1)
        return ret[0] if len(ret) > 0 else None

    def _build_multicolumn_aggs(self, agg_cls: type, on, ignore_nulls: bool, *args, skip_cols=None, **kwargs):
        if on is None:
            schema = self.schema()
            if schema is not None and len(schema.names) > 0:
                skip = skip_cols or []
                if isinstance(skip, str):
                    skip = [skip]
                on = [c for c in schema.names if c not in skip]
        if not isinstance(on, list):
            on = [on]
        return [agg_cls(on_, *args, ignore_nulls=ignore_nulls, **kwargs) for on_ in on]

    def _aggregate_on(self, agg_cls: type, on, *args, **kwargs):
        aggs = self._build_multicolumn_aggs(agg_cls, on, *args, **kwargs)
        return self.aggregate(*aggs)

    @staticmethod
    def _aggregate_result(result):
        if result is not None and len(result) == 1:
            return next(iter(result.values()))
        return result

    def sum(self, on=None, ignore_nulls: bool = True):
        return self._aggregate_result(self._aggregate_on(Sum, on, ignore_nulls))

    def min(self, on=None, ignore_nulls: bool = True):
        return self._aggregate_result(self._aggregate_on(Min, on, ignore_nulls))

    def max(self, on=None, ignore_nulls: bool = True):
        return self._aggregate_result(self._aggregate_on(Max, on, ignore_nulls))

    def mean(self, on=None, ignore_nulls: bool = True):
        return self._aggregate_result(self._aggregate_on(Mean, on, ignore_nulls))

    def std(self, on=None, ddof: int = 1, ignore_nulls: bool = True):
        return self._aggregate_result(self._aggregate_on(Std, on, ignore_nulls, ddof=ddof))

    # ------------------------------------------------------------------ simple transforms
    def select_columns(self, cols: List[str]) -> "Dataset":
        return Dataset([b.select(cols) if b.columns else b for b in self._blocks], self._stats)

    def map_batches(self, fn: Callable, *, batch_format: str = "numpy", batch_size=None, **_ignored) -> "Dataset":
        """Host-side map over whole blocks (each block makes a D2H + H2D round trip); present so
        that reference tests which interleave `map_batches(lambda t: t)` carry over."""
        out = []
        for b in self._blocks:
            if b.num_rows == 0:
                out.append(b)
                continue
            if batch_format in ("pyarrow", "arrow"):
                batch = b.to_arrow()
            elif batch_format == "pandas":
                batch = b.to_pandas()
            else:
                batch = b.to_numpy()
            out.append(BlockAccessor.for_block(fn(batch)).to_block())
        return Dataset(out, self._stats)

    def filter(self, fn: Callable[[Dict[str, Any]], bool]) -> "Dataset":
        """Host-side row filter (off the hot path)."""
        out = []
        for b in self._blocks:
            if b.num_rows == 0:
                out.append(b)
                continue
            host = b.to_numpy()
            keep = np.fromiter((bool(fn({k: v[i] for k, v in host.items()})) for i in _brange(b.num_rows)),
                               dtype=bool, count=b.num_rows)
            out.append(DeviceBlock.from_numpy({k: v[keep] for k, v in host.items()}, b.device))
        return Dataset(out, self._stats)

    def limit(self, n: int) -> "Dataset":
        out, left = [], n
        for b in self._blocks:
            if left <= 0:
                break
            take = min(left, b.num_rows)
            out.append(b.slice(0, take))
            left -= take
        return Dataset(out, self._stats)

    def materialize(self) -> "Dataset":
        return self

    # ------------------------------------------------------------------ consumers (D2H)
    def count(self) -> int:
        return sum(b.num_rows for b in self._blocks)

    def num_blocks(self) -> int:
        return len(self._blocks)

    def _block_num_rows(self) -> List[int]:
        return [b.num_rows for b in self._blocks]

    def size_bytes(self) -> int:
        return sum(b.size_bytes() for b in self._blocks)

    def schema(self, fetch_if_missing: bool = True):
        for b in self._blocks:
            s = b.schema()
            if s is not None:
                return s
        return None

    def columns(self) -> Optional[List[str]]:
        s = self.schema()
        return None if s is None else list(s.names)

    def get_internal_block_refs(self) -> List[DeviceBlock]:
        return list(self._blocks)

    def iter_internal_ref_bundles(self) -> Iterator[RefBundle]:
        return iter(self._bundles())

    def iter_rows(self) -> Iterator[Dict[str, Any]]:
        for b in self._blocks:
            yield from DeviceBlockAccessor(b).iter_rows()

    def _iter_batch_blocks(self, batch_size, drop_last, local_shuffle_buffer_size, local_shuffle_seed):
        from .batcher import blocks_to_batches

        return blocks_to_batches(iter(self._blocks), batch_size=batch_size, drop_last=drop_last,
                                 shuffle_buffer_min_size=local_shuffle_buffer_size, shuffle_seed=local_shuffle_seed)

    def iter_batches(self, *, batch_size: Optional[int] = 256, batch_format: str = "numpy", drop_last: bool = False,
                     local_shuffle_buffer_size: Optional[int] = None, local_shuffle_seed: Optional[int] = None,
                     **_ignored):
        """Batches of `batch_size` rows cut across block boundaries (batch_size=None: one batch per
        block), optionally drawn through a local shuffle buffer (dataset.py:3640-3708)."""
        for c in self._iter_batch_blocks(batch_size, drop_last, local_shuffle_buffer_size, local_shuffle_seed):
            if batch_format in ("pyarrow", "arrow"):
                yield c.to_arrow()
            elif batch_format == "pandas":
                yield c.to_pandas()
            else:
                yield c.to_numpy()

    def iter_torch_batches(self, *, batch_size: Optional[int] = 256, drop_last: bool = False,
                           local_shuffle_buffer_size: Optional[int] = None, local_shuffle_seed: Optional[int] = None,
                           **_ignored):
        """Batches as dicts of CUDA tensors (no D2H): views of the blocks where a batch lies inside one
        block, device concatenations where it spans two (dataset.py:3710-3800)."""
        for c in self._iter_batch_blocks(batch_size, drop_last, local_shuffle_buffer_size, local_shuffle_seed):
            yield dict(c.columns)

    def take(self, limit: int = 20) -> List[Dict[str, Any]]:
        return list(itertools.islice(self.iter_rows(), limit))

    def take_all(self, limit: Optional[int] = None) -> List[Dict[str, Any]]:
        out = list(self.iter_rows())
        if limit is not None and len(out) > limit:
            raise ValueError(f"The dataset has more than the given limit of {limit} records.")
        return out

    def to_pandas(self, limit: Optional[int] = None):
        import pandas as pd

        frames = [b.to_pandas() for b in self._blocks if b.num_rows > 0]
        return pd.concat(frames, ignore_index=True) if frames else pd.DataFrame()

    def to_arrow_refs(self):
        return [b.to_arrow() for b in self._blocks]

    def to_numpy(self) -> Dict[str, np.ndarray]:
        return DeviceBlock.concat(self._blocks).to_numpy()

    def stats(self) -> str:
        lines = []
        for i, (name, st) in enumerate(self._stats):
            for stage, metas in st.items():
                rows = sum(m.num_rows or 0 for m in metas)
                wall = [m.exec_stats.wall_time_s for m in metas if m.exec_stats and m.exec_stats.wall_time_s is not None]
                lines.append(f"Operator {i} {name}{stage.capitalize()}: {len(metas)} tasks executed, {rows} rows, "
                             f"wall time max {max(wall) if wall else 0:.6f}s")
        return "\n".join(lines)

    def __repr__(self):
        return f"Dataset(num_blocks={self.num_blocks()}, num_rows={self.count()}, schema={self.schema()})"


class GroupedData:
    """A dataset grouped by key (grouped_data.py:36-460)."""

    def __init__(self, dataset: Dataset, key: Union[str, List[str], None]):
        self._dataset = dataset
        self._key = key

    def __repr__(self) -> str:
        return f"{self.__class__.__name__}(dataset={self._dataset}, key={self._key!r})"

    def aggregate(self, *aggs: AggregateFn) -> Dataset:
        return self._dataset._run_all_to_all(Aggregate(None, key=self._key, aggs=list(aggs)))

    def _aggregate_on(self, agg_cls: type, on, ignore_nulls: bool, *args, **kwargs) -> Dataset:
        aggs = self._dataset._build_multicolumn_aggs(agg_cls, on, ignore_nulls, *args, skip_cols=self._key, **kwargs)
        return self.aggregate(*aggs)

    def count(self) -> Dataset:
        return self.aggregate(Count())

    def sum(self, on=None, ignore_nulls: bool = True) -> Dataset:
        return self._aggregate_on(Sum, on, ignore_nulls)

    def min(self, on=None, ignore_nulls: bool = True) -> Dataset:
        return self._aggregate_on(Min, on, ignore_nulls)

    def max(self, on=None, ignore_nulls: bool = True) -> Dataset:
        return self._aggregate_on(Max, on, ignore_nulls)

    def mean(self, on=None, ignore_nulls: bool = True) -> Dataset:
        return self._aggregate_on(Mean, on, ignore_nulls)

    def std(self, on=None, ddof: int = 1, ignore_nulls: bool = True) -> Dataset:
        return self._aggregate_on(Std, on, ignore_nulls, ddof=ddof)

    def map_groups(self, fn: Callable, *, compute=None, batch_format: Optional[str] = "default", fn_args=None,
                   fn_kwargs=None, **_ray_remote_args) -> Dataset:
        """Apply `fn` to every group (grouped_data.py:105-236): sort by the key on the device (equal keys end up in
        one block), find each block's group boundaries on the device - a group starts where ANY key column changes,
        the reference's `get_key_boundaries` over `_MultiColumnSortedKey` tuples (:181-200) - then call `fn` per
        group on the host and build ONE output block per sorted block from the results, in group order."""
        from . import ops

        fn_args, fn_kwargs = tuple(fn_args or ()), dict(fn_kwargs or {})
        if self._key is None:
            sorted_ds = self._dataset.repartition(1)
            keys: List[str] = []
        else:
            sorted_ds = self._dataset.sort(self._key)
            keys = [self._key] if isinstance(self._key, str) else list(self._key)
        out = []
        for b in sorted_ds._blocks:
            if b.num_rows == 0:
                continue
            host = b.to_numpy()
            if not keys:
                bounds = [0, b.num_rows]
            elif len(keys) <= 4 and all(b.columns[k].dim() == 1 for k in keys):
                codes = [ops.encode_keys(b.key_column(k), False) for k in keys]
                starts, ng = ops.group_boundaries_multi(codes)
                bounds = starts[: int(ng.item())].cpu().numpy().astype(np.int64).tolist() + [b.num_rows]
            else:   # wide keys: the same rule on the host copy that `fn` needs anyway
                change = np.zeros(b.num_rows - 1, dtype=bool)
                for k in keys:
                    d = host[k][1:] != host[k][:-1]
                    change |= d.reshape(len(d), -1).any(axis=1)
                bounds = [0] + (np.flatnonzero(change) + 1).tolist() + [b.num_rows]
            results = []
            for lo, hi in zip(bounds[:-1], bounds[1:]):
                grp = {c: v[lo:hi] for c, v in host.items()}
                if batch_format == "pandas":
                    import pandas as pd

                    grp = pd.DataFrame(grp)
                elif batch_format in ("pyarrow", "arrow"):
                    import pyarrow as pa

                    grp = pa.table(grp)
                results.append(BlockAccessor.for_block(fn(grp, *fn_args, **fn_kwargs)).to_block())
            out.append(DeviceBlock.concat(results))
        return Dataset(out, sorted_ds._stats)


# --------------------------------------------------------------------------- read API
def _split_rows(n: int, parallelism: int) -> List[Tuple[int, int]]:
    """Row ranges of RangeDatasource's read tasks (range_datasource.py:45, :113-131)."""
    parallelism = max(1, min(parallelism, n)) if n > 0 else 1
    block_size = max(1, n // parallelism) if n else 0
    out, i = [], 0
    while i < n:
        out.append((i, min(n, i + block_size)))
        i += block_size
    return out or [(0, 0)]


def _default_parallelism() -> int:
    return max(torch.cuda.device_count(), 1) * 8 if torch.cuda.is_available() else 8


def _maybe_records(whole: DeviceBlock, records, key_prefix=None) -> DeviceBlock:
    """`records`: None/False = dense columns; True = all columns in one record per row (pitch chosen by
    block.auto_pitch); an int = that pitch; a list of column names = those columns only. `key_prefix`: byte-string
    columns of the record that also get a dense key prefix column (block.py module docstring)."""
    if not records:
        return whole
    if isinstance(records, (list, tuple)):
        return whole.to_records(list(records), key_prefix=key_prefix)
    return whole.to_records(pitch=None if records is True else int(records), key_prefix=key_prefix)


def from_numpy_columns(cols: Dict[str, np.ndarray], *, override_num_blocks: Optional[int] = None,
                       device: Optional[torch.device] = None, records=None, key_prefix=None) -> Dataset:
    """One H2D copy per column; blocks are views of the resulting HBM buffers. `records`: see `_maybe_records`
    (row-wise record layout for datasets of wide or oddly sized rows, block.py "Record layout")."""
    device = device or default_device()
    whole = _maybe_records(DeviceBlock.from_numpy(cols, device), records, key_prefix)
    n = whole.num_rows
    parts = _split_rows(n, override_num_blocks or _default_parallelism())
    return Dataset([whole.slice(lo, hi) for lo, hi in parts])


def range(n: int, *, parallelism: int = -1, override_num_blocks: Optional[int] = None,  # noqa: A001
          device: Optional[torch.device] = None) -> Dataset:
    """`id` = 0..n-1 as int64 (read_api.py:193-243, range_datasource.py:59-66), generated on the device."""
    from . import ops

    device = device or default_device()
    nb = override_num_blocks or (parallelism if parallelism and parallelism > 0 else _default_parallelism())
    ids = torch.arange(n, dtype=torch.int64, device=device)
    whole = DeviceBlock({"id": ids}, num_rows=n, device=device)
    return Dataset([whole.slice(lo, hi) for lo, hi in _split_rows(n, nb)])


def from_items(items: List[Any], *, parallelism: int = -1, override_num_blocks: Optional[int] = None) -> Dataset:
    """Rows given as dicts (or scalars under the column "item") (read_api.py:107-190)."""
    nb = override_num_blocks or (parallelism if parallelism and parallelism > 0 else _default_parallelism())
    if len(items) == 0:
        return Dataset([])
    if isinstance(items[0], dict):
        cols = {k: np.asarray([it[k] for it in items]) for k in items[0]}
    else:
        cols = {"item": np.asarray(items)}
    return from_numpy_columns(cols, override_num_blocks=nb)


def from_numpy(arr: Union[np.ndarray, List[np.ndarray]], *, override_num_blocks: Optional[int] = None) -> Dataset:
    """Column "data" from one ndarray or one block per ndarray (read_api.py:2469-2553)."""
    if isinstance(arr, list):
        return Dataset([DeviceBlock.from_numpy({"data": a}) for a in arr])
    return from_numpy_columns({"data": arr}, override_num_blocks=override_num_blocks or 1)


def from_arrow(tables, *, override_num_blocks: Optional[int] = None) -> Dataset:
    """One block per Arrow table (read_api.py:2556-2589)."""
    import pyarrow as pa

    if isinstance(tables, pa.Table):
        tables = [tables]
    return Dataset([DeviceBlock.from_arrow(t) for t in tables])


def from_pandas(dfs, *, override_num_blocks: Optional[int] = None) -> Dataset:
    """One block per DataFrame (read_api.py:2360-2406)."""
    import pandas as pd

    if isinstance(dfs, pd.DataFrame):
        dfs = [dfs]
    ds = Dataset([DeviceBlock.from_pandas(df) for df in dfs])
    if override_num_blocks:
        ds = ds.repartition(override_num_blocks)
    return ds


def from_torch_columns(cols: Dict[str, torch.Tensor], *, override_num_blocks: Optional[int] = None,
                       records=None, key_prefix=None) -> Dataset:
    """Adopt CUDA tensors that are already resident in HBM (no copy, unless `records` asks for the row-wise
    record layout: see `_maybe_records`)."""
    whole = _maybe_records(DeviceBlock(cols), records, key_prefix)
    parts = _split_rows(whole.num_rows, override_num_blocks or 1)
    return Dataset([whole.slice(lo, hi) for lo, hi in parts])
