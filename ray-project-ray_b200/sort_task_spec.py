"""SortKey and SortTaskSpec (python/ray/data/_internal/planner/exchange/sort_task_spec.py:18-203)."""
from __future__ import annotations

from typing import Any, List, Optional, Tuple, TypeVar, Union

import numpy as np

from .block import Block, BlockAccessor, BlockExecStats, BlockMetadata, DeviceBlock
from .builder import DeviceBlockBuilder
from .interfaces import ExchangeTaskSpec

T = TypeVar("T")


class SortKey:
    """Normalised sort arguments: column list, one direction, optional numeric boundaries."""

    def __init__(self, key: Optional[Union[str, List[str]]] = None, descending: Union[bool, List[bool]] = False,
                 boundaries: Optional[list] = None):
        if key is None:
            key = []
        if isinstance(key, str):
            key = [key]
        if not (isinstance(key, list) and all(isinstance(k, str) for k in key)):
            raise ValueError(f"Key must be a string or a list of strings, but got {key}.")
        if isinstance(descending, bool):
            descending = [descending for _ in key]
        elif isinstance(descending, list):
            if len(descending) != len(key):
                raise ValueError("Length of `descending` does not match the length of the key.")
            if len(set(descending)) != 1:
                raise ValueError("Sorting with mixed key orders not supported yet.")
        self._columns = key
        self._descending = descending
        if boundaries:
            for item in boundaries:
                if not isinstance(item, (int, float)):
                    raise ValueError("The type of items in boundaries must be int or float.")
            boundaries = sorted(set(boundaries))
        self._boundaries = boundaries

    def get_columns(self) -> List[str]:
        return self._columns

    def get_descending(self) -> bool:
        return bool(self._descending[0]) if self._descending else False

    def to_arrow_sort_args(self) -> List[Tuple[str, str]]:
        order = "descending" if self.get_descending() else "ascending"
        return [(key, order) for key in self._columns]

    def to_pandas_sort_args(self) -> Tuple[List[str], bool]:
        return self._columns, not self.get_descending()

    def validate_schema(self, schema) -> None:
        if schema is None:
            return
        if self._columns and len(schema.names) > 0:
            for column in self._columns:
                if column not in schema.names:
                    raise ValueError("The column '{}' does not exist in the schema '{}'.".format(column, schema))

    @property
    def boundaries(self):
        return self._boundaries


class SortTaskSpec(ExchangeTaskSpec):
    """Distributed sort = sample boundaries, sort+partition each block (map), merge (reduce)."""

    SORT_SAMPLE_SUB_PROGRESS_BAR_NAME = "Sort Sample"

    def __init__(self, boundaries: List[T], sort_key: SortKey):
        super().__init__(map_args=[boundaries, sort_key], reduce_args=[sort_key])

    @staticmethod
    def map(idx: int, block: Block, output_num_blocks: int, boundaries: List[T],
            sort_key: SortKey) -> List[Union[BlockMetadata, Block]]:
        stats = BlockExecStats.builder()
        acc = BlockAccessor.for_block(block)
        out = acc.sort_and_partition(boundaries, sort_key)
        meta = acc.get_metadata(input_files=None, exec_stats=stats.build())
        return out + [meta]

    @staticmethod
    def reduce(sort_key: SortKey, *mapper_outputs: Block, partial_reduce: bool = False) -> Tuple[Block, BlockMetadata]:
        return BlockAccessor.for_block(mapper_outputs[0]).merge_sorted_blocks(list(mapper_outputs), sort_key)

    @staticmethod
    def sample_boundaries(blocks: List[Block], sort_key: SortKey, num_reducers: int,
                          samples_per_block: Optional[int] = None) -> List[T]:
        """num_reducers - 1 ascending boundary tuples that cut the key domain into ranges of about
        equal row count (sort_task_spec.py:149-199): sample every block, lexsort the samples,
        take per-column nearest-rank quantiles at linspace(0, 1, R)[1:]."""
        columns = sort_key.get_columns()
        if len(blocks) == 0:
            return [None] * (num_reducers - 1)
        n_samples = samples_per_block or int(num_reducers * 10 / len(blocks))
        n_samples = max(n_samples, 1)
        samples = [BlockAccessor.for_block(b).sample(n_samples, sort_key) for b in blocks]
        samples = [s for s in samples if s.num_rows > 0]
        if len(samples) == 0:
            return [None] * (num_reducers - 1)
        builder = DeviceBlockBuilder()
        for s in samples:
            builder.add_block(s)
        sample_dict = builder.build().to_numpy(columns)  # a few rows per block: host quantiles
        indices = np.lexsort(list(reversed(list(sample_dict.values()))))
        for k, v in sample_dict.items():
            sorted_v = v[indices]
            sample_dict[k] = list(np.quantile(sorted_v, np.linspace(0, 1, num_reducers), method="nearest")[1:])
        return [tuple(sample_dict[k][i] for k in sample_dict) for i in range(num_reducers - 1)]


def _sample_block(block: Block, n_samples: int, sort_key: SortKey) -> Block:
    return BlockAccessor.for_block(block).sample(n_samples, sort_key)
