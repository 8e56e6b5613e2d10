"""SortAggregateTaskSpec (python/ray/data/_internal/planner/exchange/aggregate_task_spec.py:11-106)."""
from __future__ import annotations

from typing import List, Optional, Tuple, Union

from .aggregate import AggregateFn, Count, _AggregateOnKeyBase
from .block import Block, BlockAccessor, BlockExecStats, BlockMetadata, DeviceBlock, KeyType
from .interfaces import ExchangeTaskSpec
from .sort_task_spec import SortKey


class SortAggregateTaskSpec(ExchangeTaskSpec):
    """map: prune columns, sort + range-partition, combine each partition into partial accumulators;
    reduce: merge the partial rows of one key range and (unless partial_reduce) finalize."""

    def __init__(self, boundaries: List[KeyType], key: Optional[str], aggs: List[AggregateFn]):
        super().__init__(map_args=[boundaries, key, aggs], reduce_args=[key, aggs])

    @staticmethod
    def map(idx: int, block: Block, output_num_blocks: int, boundaries: List[KeyType],
            key: Union[str, List[str], None], aggs: List[AggregateFn]) -> List[Union[BlockMetadata, Block]]:
        stats = BlockExecStats.builder()
        block = SortAggregateTaskSpec._prune_unused_columns(block, key, aggs)
        if key is None:
            partitions = [block]
        else:
            partitions = BlockAccessor.for_block(block).sort_and_partition(boundaries, SortKey(key))
        parts = [BlockAccessor.for_block(p).combine(key, aggs) for p in partitions]
        meta = BlockAccessor.for_block(block).get_metadata(input_files=None, exec_stats=stats.build())
        return parts + [meta]

    @staticmethod
    def reduce(key: Optional[str], aggs: List[AggregateFn], *mapper_outputs: Block,
               partial_reduce: bool = False) -> Tuple[Block, BlockMetadata]:
        return BlockAccessor.for_block(mapper_outputs[0]).aggregate_combined_blocks(
            list(mapper_outputs), key, aggs, finalize=not partial_reduce)

    @staticmethod
    def _prune_unused_columns(block: Block, key: Union[str, List[str], None], aggs: Tuple[AggregateFn]) -> Block:
        prune_columns = True
        columns = []
        if isinstance(key, str):
            columns.append(key)
        elif isinstance(key, list):
            columns.extend(key)
        elif callable(key):
            prune_columns = False
        for agg in aggs:
            if isinstance(agg, _AggregateOnKeyBase) and isinstance(agg._key_fn, str):
                if agg._key_fn not in columns:
                    columns.append(agg._key_fn)
            elif not isinstance(agg, Count):
                prune_columns = False
        acc = BlockAccessor.for_block(block)
        if prune_columns and acc.num_rows() > 0:
            return acc.select(columns)
        return acc.to_block()
