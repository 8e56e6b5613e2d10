"""Logical all-to-all operators, their planner functions and the physical AllToAllOperator.

Mirrors
  logical ops      python/ray/data/_internal/logical/operators/all_to_all_operator.py:10-141
  planner fns      python/ray/data/_internal/planner/sort.py:21-74, random_shuffle.py:20-84,
                   aggregate.py:24-85, repartition.py:23-82, plan_all_to_all_op.py:21-92
  physical op      python/ray/data/_internal/execution/operators/base_physical_operator.py:40-132
"""
from __future__ import annotations

from functools import partial
from typing import Any, Dict, List, Optional, Tuple

import torch

from . import ops
from .aggregate import AggregateFn
from .aggregate_task_spec import SortAggregateTaskSpec
from .block import BYTE_KEY_SEP, DeviceBlock, DeviceBlockAccessor, _contig, _row_pitch
from .context import DataContext
from .interfaces import (AllToAllTransformFn, ExchangeTaskSpec, RefBundle, StatsDict, TaskContext, _NullProgressBar)
from .schedulers import (AutoBoundaries, GpuExchangeScheduler, PullBasedShuffleTaskScheduler,
                         PushBasedShuffleTaskScheduler)
from .shuffle_task_spec import ShuffleTaskSpec
from .sort_task_spec import SortKey, SortTaskSpec


# ------------------------------- logical operators -----------------------------------
class AbstractAllToAll:
    def __init__(self, name: str, input_op: Any = None, num_outputs: Optional[int] = None,
                 sub_progress_bar_names: Optional[List[str]] = None,
                 ray_remote_args: Optional[Dict[str, Any]] = None):
        self.name = name
        self._input_op = input_op
        self._num_outputs = num_outputs
        self._ray_remote_args = ray_remote_args or {}
        self._sub_progress_bar_names = sub_progress_bar_names


class RandomShuffle(AbstractAllToAll):
    def __init__(self, input_op: Any = None, seed: Optional[int] = None, num_outputs: Optional[int] = None,
                 ray_remote_args: Optional[Dict[str, Any]] = None):
        super().__init__("RandomShuffle", input_op, num_outputs=num_outputs,
                         sub_progress_bar_names=[ExchangeTaskSpec.MAP_SUB_PROGRESS_BAR_NAME,
                                                 ExchangeTaskSpec.REDUCE_SUB_PROGRESS_BAR_NAME],
                         ray_remote_args=ray_remote_args)
        self._seed = seed


class Repartition(AbstractAllToAll):
    def __init__(self, input_op: Any = None, num_outputs: int = 1, shuffle: bool = False):
        names = ([ExchangeTaskSpec.MAP_SUB_PROGRESS_BAR_NAME, ExchangeTaskSpec.REDUCE_SUB_PROGRESS_BAR_NAME]
                 if shuffle else [ShuffleTaskSpec.SPLIT_REPARTITION_SUB_PROGRESS_BAR_NAME])
        super().__init__("Repartition", input_op, num_outputs=num_outputs, sub_progress_bar_names=names)
        self._shuffle = shuffle


class RandomizeBlocks(AbstractAllToAll):
    """logical/operators/all_to_all_operator.py:37-57."""

    def __init__(self, input_op: Any = None, seed: Optional[int] = None):
        super().__init__("RandomizeBlockOrder", input_op)
        self._seed = seed
        self._name = self.name


class Sort(AbstractAllToAll):
    def __init__(self, input_op: Any = None, sort_key: SortKey = None):
        super().__init__("Sort", input_op,
                         sub_progress_bar_names=[SortTaskSpec.SORT_SAMPLE_SUB_PROGRESS_BAR_NAME,
                                                 ExchangeTaskSpec.MAP_SUB_PROGRESS_BAR_NAME,
                                                 ExchangeTaskSpec.REDUCE_SUB_PROGRESS_BAR_NAME])
        self._sort_key = sort_key


class Aggregate(AbstractAllToAll):
    def __init__(self, input_op: Any = None, key: Optional[str] = None, aggs: List[AggregateFn] = None):
        super().__init__("Aggregate", input_op,
                         sub_progress_bar_names=[SortTaskSpec.SORT_SAMPLE_SUB_PROGRESS_BAR_NAME,
                                                 ExchangeTaskSpec.MAP_SUB_PROGRESS_BAR_NAME,
                                                 ExchangeTaskSpec.REDUCE_SUB_PROGRESS_BAR_NAME])
        self._key = key
        self._aggs = aggs or []


# --------------------------------- planner functions ---------------------------------
def _pick_scheduler(spec: ExchangeTaskSpec):
    ctx = DataContext.get_current()
    if ctx.use_fused_device_exchange:
        return GpuExchangeScheduler(spec)
    if ctx.use_push_based_shuffle:
        return PushBasedShuffleTaskScheduler(spec)
    return PullBasedShuffleTaskScheduler(spec)


def _blocks_and_meta(refs: List[RefBundle]):
    blocks, metadata = [], []
    for ref_bundle in refs:
        for block, block_metadata in ref_bundle.blocks:
            blocks.append(block)
            metadata.append(block_metadata)
    return blocks, metadata


def _unified_schema(metadata):
    for m in metadata:
        if m.schema is not None and (m.num_rows is None or m.num_rows > 0):
            return m.schema
    return None




def _byte_key_columns(blocks: List[DeviceBlock], sort_key: SortKey) -> List[str]:
    """Key columns that are fixed-width byte strings ([n, w] uint8)."""
    for b in blocks:
        if b.columns:
            return [k for k in sort_key.get_columns()
                    if k in b.columns and b.columns[k].dim() == 2 and b.columns[k].dtype == torch.uint8]
    return []


def _expand_byte_keys(refs: List[RefBundle], sort_key: SortKey, byte_cols: List[str]):
    """A w-byte string key sorts like the tuple of its big-endian 8-byte chunks: replace the key column by
    those chunks as unsigned integer columns `<name>::be<i>` and sort by them (every scheduler, the boundary
    sampling and the multi-GPU exchange then see an ordinary multi-column integer key). The chunks hold
    every byte of the key, so the key column itself does not travel: it is rebuilt from the sorted chunks.

    A key column that is a record member travels inside its record anyway. Under the fused scheduler such a
    column is left where it is and only the SortKey is rewritten: the chunk names are then resolved on demand by
    `DeviceBlock.key_column`, and nothing is added to or rebuilt in the blocks."""
    if DataContext.get_current().use_fused_device_exchange and all(
            _row_pitch(b.columns[k]) is not None for ref in refs for b, _ in ref.blocks if b.num_rows for k in byte_cols):
        new_cols = []
        for k in sort_key.get_columns():
            if k in byte_cols:
                w = next(int(b.columns[k].shape[1]) for ref in refs for b, _ in ref.blocks if b.columns)
                new_cols += [f"{k}{BYTE_KEY_SEP}{i}" for i in range((w + 7) // 8)]
            else:
                new_cols.append(k)
        return refs, SortKey(new_cols, descending=[sort_key.get_descending()] * len(new_cols)), None
    out_refs, width, order = [], {}, None
    for ref in refs:
        pairs = []
        for block, _ in ref.blocks:
            if block.columns:
                order = order or list(block.columns)
                cols = {}
                for name, col in block.columns.items():
                    if name in byte_cols:
                        width[name] = int(col.shape[1])
                        for i, chunk in enumerate(ops.byte_key_chunks(_contig(col))):
                            cols[f"{name}{BYTE_KEY_SEP}{i}"] = chunk
                    else:
                        cols[name] = col
                block = DeviceBlock(cols, num_rows=block.num_rows, device=block.device)
            pairs.append((block, DeviceBlockAccessor(block).get_metadata(None, None)))
        out_refs.append(RefBundle(pairs, owns_blocks=ref.owns_blocks))
    new_cols = []
    for k in sort_key.get_columns():
        if k in byte_cols:
            new_cols += [f"{k}{BYTE_KEY_SEP}{i}" for i in range((width[k] + 7) // 8)]
        else:
            new_cols.append(k)
    new_key = SortKey(new_cols, descending=[sort_key.get_descending()] * len(new_cols))
    return out_refs, new_key, (width, order)


def _restore_byte_keys(refs: List[RefBundle], packed) -> List[RefBundle]:
    """Undo _expand_byte_keys on the output blocks: chunks -> key column, original column order."""
    if packed is None:   # the key columns travelled inside their records
        return refs
    width, order = packed
    out = []
    for ref in refs:
        pairs = []
        for block, meta in ref.blocks:
            if block.columns:
                cols = {}
                for name in order:
                    if name in width:
                        chunks = [_contig(block.columns[f"{name}{BYTE_KEY_SEP}{i}"]) for i in range((width[name] + 7) // 8)]
                        cols[name] = ops.unpack_key_bytes(chunks, width[name])
                    else:
                        cols[name] = block.columns[name]
                block = DeviceBlock(cols, num_rows=block.num_rows, device=block.device)
                meta = DeviceBlockAccessor(block).get_metadata(None, exec_stats=meta.exec_stats)
            pairs.append((block, meta))
        out.append(RefBundle(pairs, owns_blocks=ref.owns_blocks))
    return out


def generate_sort_fn(sort_key: SortKey,
                     _debug_limit_shuffle_execution_to_num_blocks: Optional[int] = None) -> AllToAllTransformFn:
    """Sort blocks by the key column(s) (planner/sort.py:21-74)."""

    def fn(sort_key: SortKey, refs: List[RefBundle], ctx: TaskContext, _expanded: bool = False) -> Tuple[List[RefBundle], StatsDict]:
        blocks, metadata = _blocks_and_meta(refs)
        if len(blocks) == 0:
            return (blocks, {})
        if not _expanded:
            sort_key.validate_schema(_unified_schema(metadata))
        byte_cols = [] if _expanded else _byte_key_columns(blocks, sort_key)
        if byte_cols:
            if sort_key.boundaries:
                raise ValueError("user-specified boundaries are numeric; they cannot cut a byte-string key")
            refs2, sort_key2, packed = _expand_byte_keys(refs, sort_key, byte_cols)
            out, stats = fn(sort_key2, refs2, ctx, True)
            return _restore_byte_keys(out, packed), stats
        num_outputs = len(blocks)  # same number of output partitions as mappers
        if not sort_key.boundaries and DataContext.get_current().use_fused_device_exchange:
            boundaries = AutoBoundaries(num_outputs)  # the fused scheduler cuts the sorted result itself
        elif not sort_key.boundaries:
            boundaries = SortTaskSpec.sample_boundaries(
                blocks, sort_key, num_outputs,
                samples_per_block=max(DataContext.get_current().sort_samples_per_block,
                                      int(num_outputs * 10 / len(blocks))))
        else:
            boundaries = [(b,) for b in sort_key.boundaries]
            num_outputs = len(boundaries) + 1
        _, ascending = sort_key.to_pandas_sort_args()
        if not ascending and not isinstance(boundaries, AutoBoundaries):
            boundaries.reverse()
        sort_spec = SortTaskSpec(boundaries=boundaries, sort_key=sort_key)
        return _pick_scheduler(sort_spec).execute(
            refs, num_outputs, ctx,
            _debug_limit_execution_to_num_blocks=_debug_limit_shuffle_execution_to_num_blocks)

    return partial(fn, sort_key)


def generate_random_shuffle_fn(seed: Optional[int], num_outputs: Optional[int] = None,
                               ray_remote_args: Optional[Dict[str, Any]] = None,
                               _debug_limit_shuffle_execution_to_num_blocks: Optional[int] = None) -> AllToAllTransformFn:
    """Randomly shuffle every row (planner/random_shuffle.py:20-84)."""

    def fn(refs: List[RefBundle], ctx: TaskContext) -> Tuple[List[RefBundle], StatsDict]:
        num_input_blocks = sum(len(r.blocks) for r in refs)
        map_transformer = ctx.upstream_map_transformer
        upstream_map_fn = None
        remote_args = ray_remote_args
        if map_transformer:
            map_transformer.set_target_max_block_size(float("inf"))

            def upstream_map_fn(blocks):
                return map_transformer.apply_transform(blocks, ctx)

            remote_args = ctx.upstream_map_ray_remote_args
        shuffle_spec = ShuffleTaskSpec(ctx.target_max_block_size, random_shuffle=True, random_seed=seed,
                                       upstream_map_fn=upstream_map_fn)
        dctx = DataContext.get_current()
        if dctx.use_push_based_shuffle and not dctx.use_fused_device_exchange and num_outputs is not None:
            raise NotImplementedError("Push-based shuffle doesn't support setting num_blocks yet.")
        return _pick_scheduler(shuffle_spec).execute(
            refs, num_outputs or num_input_blocks, task_ctx=ctx, map_ray_remote_args=remote_args,
            reduce_ray_remote_args=remote_args,
            _debug_limit_execution_to_num_blocks=_debug_limit_shuffle_execution_to_num_blocks)

    return fn


def generate_aggregate_fn(key: Optional[str], aggs: List[AggregateFn],
                          _debug_limit_shuffle_execution_to_num_blocks: Optional[int] = None) -> AllToAllTransformFn:
    """Aggregate blocks by key (planner/aggregate.py:24-85)."""
    if len(aggs) == 0:
        raise ValueError("Aggregate requires at least one aggregation")

    def fn(refs: List[RefBundle], ctx: TaskContext) -> Tuple[List[RefBundle], StatsDict]:
        blocks, metadata = _blocks_and_meta(refs)
        if len(blocks) == 0:
            return (blocks, {})
        unified_schema = _unified_schema(metadata)
        for agg_fn in aggs:
            agg_fn._validate(unified_schema)
        if key is None:
            num_outputs, boundaries = 1, []
        elif DataContext.get_current().use_fused_device_exchange:
            num_outputs = len(blocks)
            boundaries = AutoBoundaries(num_outputs)
        else:
            num_outputs = len(blocks)
            boundaries = SortTaskSpec.sample_boundaries(
                blocks, SortKey(key), num_outputs,
                samples_per_block=max(DataContext.get_current().sort_samples_per_block,
                                      int(num_outputs * 10 / len(blocks))))
        agg_spec = SortAggregateTaskSpec(boundaries=boundaries, key=key, aggs=aggs)
        return _pick_scheduler(agg_spec).execute(
            refs, num_outputs, ctx,
            _debug_limit_execution_to_num_blocks=_debug_limit_shuffle_execution_to_num_blocks)

    return fn


def generate_repartition_fn(num_outputs: int, shuffle: bool,
                            _debug_limit_shuffle_execution_to_num_blocks: Optional[int] = None) -> AllToAllTransformFn:
    """Repartition to num_outputs blocks keeping row order (planner/repartition.py:23-82). Both the
    shuffle and the split flavour keep order; on HBM both are re-cuts of block views."""

    def fn(refs: List[RefBundle], ctx: TaskContext) -> Tuple[List[RefBundle], StatsDict]:
        if shuffle:
            spec = ShuffleTaskSpec(ctx.target_max_block_size, random_shuffle=False)
            return _pick_scheduler(spec).execute(
                refs, num_outputs, ctx,
                _debug_limit_execution_to_num_blocks=_debug_limit_shuffle_execution_to_num_blocks)
        from .block import BlockExecStats, DeviceBlock, DeviceBlockAccessor
        from .schedulers import _bundle, _input_blocks

        stats = BlockExecStats.builder()
        table = DeviceBlock.concat(_input_blocks(refs))
        n = table.num_rows
        base, extra = divmod(n, num_outputs)
        outs, lo = [], 0
        for j in range(num_outputs):
            hi = lo + base + (1 if j < extra else 0)
            outs.append(table.slice(lo, hi) if n else DeviceBlock.empty(table.device))
            lo = hi
        st = stats.build()
        meta = [DeviceBlockAccessor(b).get_metadata(None, exec_stats=st) for b in outs]
        owned = all(b.owns_blocks for b in refs)
        return [_bundle(b, m, owned) for b, m in zip(outs, meta)], {"split": meta}

    return fn


def generate_randomize_blocks_fn(op: RandomizeBlocks) -> AllToAllTransformFn:
    """Shuffle the order of the blocks, not their rows (planner/randomize_blocks.py:12-57): a host-side
    `random.shuffle` of the (block, metadata) list, seeded with `random.seed(seed)` when a seed is given -
    the same calls as the reference, so a seed gives the same block order. No data moves."""

    def fn(refs: List[RefBundle], context: TaskContext) -> Tuple[List[RefBundle], StatsDict]:
        import random

        blocks_with_metadata = [(block, meta) for ref_bundle in refs for block, meta in ref_bundle.blocks]
        if len(blocks_with_metadata) == 0:
            return refs, {op._name: []}
        if op._seed is not None:
            random.seed(op._seed)
        input_owned = all(b.owns_blocks for b in refs)
        random.shuffle(blocks_with_metadata)
        output = [RefBundle([(block, meta)], owns_blocks=input_owned) for block, meta in blocks_with_metadata]
        return output, {op._name: [meta for _, meta in blocks_with_metadata]}

    return fn


class AllToAllOperator:
    """Blocking physical operator: buffers every input bundle, then runs `bulk_fn` once
    (base_physical_operator.py:40-132)."""

    def __init__(self, bulk_fn: AllToAllTransformFn, input_op: Any = None, target_max_block_size: Optional[int] = None,
                 num_outputs: Optional[int] = None, sub_progress_bar_names: Optional[List[str]] = None,
                 name: str = "AllToAll"):
        self._bulk_fn = bulk_fn
        self._next_task_index = 0
        self._num_outputs = num_outputs
        self._target_max_block_size = target_max_block_size
        self._output_rows = 0
        self._sub_progress_bar_names = sub_progress_bar_names
        self._sub_progress_bar_dict = None
        self._input_buffer: List[RefBundle] = []
        self._output_buffer: List[RefBundle] = []
        self._stats: StatsDict = {}
        self._inputs_complete = False
        self.name = name
        self.input_op = input_op

    def num_outputs_total(self) -> Optional[int]:
        return self._num_outputs

    def add_input(self, refs: RefBundle, input_index: int = 0) -> None:
        assert not self._inputs_complete
        assert input_index == 0, input_index
        self._input_buffer.append(refs)

    def all_inputs_done(self) -> None:
        ctx = TaskContext(task_idx=self._next_task_index, sub_progress_bar_dict=self._sub_progress_bar_dict,
                          target_max_block_size=self._target_max_block_size)
        self._output_buffer, self._stats = self._bulk_fn(self._input_buffer, ctx)
        self._next_task_index += 1
        self._input_buffer.clear()
        self._inputs_complete = True

    def has_next(self) -> bool:
        return len(self._output_buffer) > 0

    def get_next(self) -> RefBundle:
        bundle = self._output_buffer.pop(0)
        self._output_rows += bundle.num_rows() or 0
        return bundle

    def get_stats(self) -> StatsDict:
        return self._stats

    def initialize_sub_progress_bars(self, position: int = 0) -> int:
        self._sub_progress_bar_dict = {}
        for name in self._sub_progress_bar_names or []:
            self._sub_progress_bar_dict[name] = _NullProgressBar(name, self._num_outputs or 0)
        return len(self._sub_progress_bar_dict)

    def close_sub_progress_bars(self) -> None:
        for bar in (self._sub_progress_bar_dict or {}).values():
            bar.close()


def plan_all_to_all_op(op: AbstractAllToAll, input_physical_dag: Any = None) -> AllToAllOperator:
    """Logical all-to-all op -> AllToAllOperator (plan_all_to_all_op.py:21-92)."""
    ctx = DataContext.get_current()
    target_max_block_size = None
    if isinstance(op, RandomizeBlocks):
        fn = generate_randomize_blocks_fn(op)
    elif isinstance(op, RandomShuffle):
        debug = ctx.get_config("debug_limit_shuffle_execution_to_num_blocks", None)
        fn = generate_random_shuffle_fn(op._seed, op._num_outputs, op._ray_remote_args, debug)
        target_max_block_size = ctx.target_shuffle_max_block_size
    elif isinstance(op, Repartition):
        debug = None
        if op._shuffle:
            target_max_block_size = ctx.target_shuffle_max_block_size
            debug = ctx.get_config("debug_limit_shuffle_execution_to_num_blocks", None)
        fn = generate_repartition_fn(op._num_outputs, op._shuffle, debug)
    elif isinstance(op, Sort):
        debug = ctx.get_config("debug_limit_shuffle_execution_to_num_blocks", None)
        fn = generate_sort_fn(op._sort_key, debug)
        target_max_block_size = ctx.target_shuffle_max_block_size
    elif isinstance(op, Aggregate):
        debug = ctx.get_config("debug_limit_shuffle_execution_to_num_blocks", None)
        fn = generate_aggregate_fn(op._key, op._aggs, debug)
        target_max_block_size = ctx.target_shuffle_max_block_size
    else:
        raise ValueError(f"Found unknown logical operator during planning: {op}")
    return AllToAllOperator(fn, input_physical_dag, target_max_block_size=target_max_block_size,
                            num_outputs=op._num_outputs, sub_progress_bar_names=op._sub_progress_bar_names,
                            name=op.name)
