"""ShuffleTaskSpec (python/ray/data/_internal/planner/exchange/shuffle_task_spec.py:15-132):
used by random_shuffle() and shuffle-repartition()."""
from __future__ import annotations

import logging
import math
from typing import Callable, Iterable, List, Optional, Tuple, Union

from . import ops
from .block import Block, BlockAccessor, BlockExecStats, BlockMetadata, DeviceBlock
from .builder import DeviceBlockBuilder
from .context import MAX_SAFE_BLOCK_SIZE_FACTOR
from .interfaces import ExchangeTaskSpec

logger = logging.getLogger("ray.data")

_M64 = (1 << 64) - 1


def slice_order_seed(seed_i: int) -> int:
    """Seed of the slice -> reducer assignment: an independent stream derived from the block seed
    (the reference reuses RandomState(seed_i) after the row permutation, shuffle_task_spec.py:97-99)."""
    x = (seed_i + 0x9E3779B97F4A7C15) & _M64
    x = ((x ^ (x >> 30)) * 0xBF58476D1CE4E5B9) & _M64
    x = ((x ^ (x >> 27)) * 0x94D049BB133111EB) & _M64
    return x ^ (x >> 31)


class ShuffleTaskSpec(ExchangeTaskSpec):
    SPLIT_REPARTITION_SUB_PROGRESS_BAR_NAME = "Split Repartition"

    def __init__(self, target_shuffle_max_block_size: int, random_shuffle: bool = False,
                 random_seed: Optional[int] = None,
                 upstream_map_fn: Optional[Callable[[Iterable[Block]], Iterable[Block]]] = None):
        super().__init__(
            map_args=[target_shuffle_max_block_size, upstream_map_fn, random_shuffle, random_seed],
            reduce_args=[random_shuffle, random_seed],
        )

    @staticmethod
    def map(idx: int, block: Block, output_num_blocks: int, target_shuffle_max_block_size: int,
            upstream_map_fn: Optional[Callable[[Iterable[Block]], Iterable[Block]]], random_shuffle: bool,
            random_seed: Optional[int]) -> List[Union[BlockMetadata, Block]]:
        stats = BlockExecStats.builder()
        if upstream_map_fn:
            builder = DeviceBlockBuilder()
            for mapped_block in upstream_map_fn([block]):
                builder.add_block(mapped_block)
            block = builder.build()
        acc = BlockAccessor.for_block(block)
        if target_shuffle_max_block_size and acc.size_bytes() > MAX_SAFE_BLOCK_SIZE_FACTOR * target_shuffle_max_block_size:
            logger.warning(
                "Input block to map task has size %dMiB, which exceeds "
                "DataContext.get_current().target_shuffle_max_block_size=%dMiB.",
                acc.size_bytes() // (1024 * 1024), target_shuffle_max_block_size // (1024 * 1024))

        seed_i = None
        if random_shuffle:
            # shuffle_task_spec.py:83: every block permutes with its own seed
            seed_i = random_seed + idx if random_seed is not None else None
            block = acc.random_shuffle(seed_i)
            acc = BlockAccessor.for_block(block)

        n = acc.num_rows()
        slice_sz = max(1, math.ceil(n / output_num_blocks))
        slices = [acc.slice(i * slice_sz, (i + 1) * slice_sz) for i in range(output_num_blocks)]

        if random_shuffle:
            # which reducer gets which slice is itself shuffled (prevents empty outputs for tiny blocks)
            s = slice_order_seed(seed_i if seed_i is not None else idx)
            order = ops.perm_indices_host(output_num_blocks, s)
            slices = [slices[int(j)] for j in order]

        num_rows = sum(s.num_rows for s in slices)
        assert num_rows == n, (num_rows, n)
        metadata = acc.get_metadata(input_files=None, exec_stats=stats.build())
        return slices + [metadata]

    @staticmethod
    def reduce(random_shuffle: bool, random_seed: Optional[int], *mapper_outputs: Block,
               partial_reduce: bool = False) -> Tuple[Block, BlockMetadata]:
        stats = BlockExecStats.builder()
        builder = DeviceBlockBuilder()
        for block in mapper_outputs:
            builder.add_block(block)
        new_block = builder.build()
        acc = BlockAccessor.for_block(new_block)
        if random_shuffle:
            new_block = acc.random_shuffle(random_seed if random_seed is not None else None)
            acc = BlockAccessor.for_block(new_block)
        new_metadata = BlockMetadata(num_rows=acc.num_rows(), size_bytes=acc.size_bytes(), schema=acc.schema(),
                                     input_files=None, exec_stats=stats.build())
        return new_block, new_metadata
