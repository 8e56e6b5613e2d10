"""The plugin surface of the exchange: task spec, scheduler, and the execution-side types.

Mirrors python/ray/data/_internal/planner/exchange/interfaces.py:14-130 (ExchangeTaskSpec,
ExchangeTaskScheduler), execution/interfaces/ref_bundle.py:13-115 (RefBundle),
execution/interfaces/task_context.py:11-41 (TaskContext) and
execution/interfaces/transform_fn.py:8-10 (AllToAllTransformFn).

There is no object store here: a "block ref" is the DeviceBlock itself.
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import Any, Callable, Dict, List, Optional, Tuple, Union

from .block import Block, BlockMetadata, DeviceBlock
from .context import DataContext

StatsDict = Dict[str, List[BlockMetadata]]


@dataclass
class RefBundle:
    """A group of (block, metadata) pairs moved between operators as one unit."""

    blocks: Tuple[Tuple[DeviceBlock, BlockMetadata], ...]
    owns_blocks: bool

    def __post_init__(self):
        if not isinstance(self.blocks, tuple):
            object.__setattr__(self, "blocks", tuple(self.blocks))
        for b in self.blocks:
            assert len(b) == 2, b
            assert isinstance(b[1], BlockMetadata), b
            if b[1].size_bytes is None:
                raise ValueError("The size in bytes of the block must be known: {}".format(b))

    def __setattr__(self, key, value):
        if hasattr(self, key) and key in ("blocks", "owns_blocks"):
            raise ValueError(f"The `{key}` field of RefBundle cannot be updated.")
        object.__setattr__(self, key, value)

    @property
    def block_refs(self) -> List[DeviceBlock]:
        return [b for b, _ in self.blocks]

    @property
    def metadata(self) -> List[BlockMetadata]:
        return [m for _, m in self.blocks]

    def num_rows(self) -> Optional[int]:
        total = 0
        for _, m in self.blocks:
            if m.num_rows is None:
                return None
            total += m.num_rows
        return total

    def size_bytes(self) -> int:
        return sum(m.size_bytes for _, m in self.blocks)

    def destroy_if_owned(self) -> int:
        return self.size_bytes() if self.owns_blocks else 0

    def __eq__(self, other) -> bool:
        return self is other

    def __hash__(self) -> int:
        return id(self)

    def __len__(self) -> int:
        return len(self.blocks)


class _NullProgressBar:
    def __init__(self, name: str = "", total: int = 0):
        self.name, self.total, self.done = name, total, 0

    def fetch_until_complete(self, refs: List[Any]) -> List[Any]:
        self.done += len(refs)
        return list(refs)

    def update(self, n: int = 1, total: Optional[int] = None) -> None:
        self.done += n

    def set_description(self, name: str) -> None:
        self.name = name

    def close(self) -> None:
        pass


@dataclass
class TaskContext:
    """Per-operator execution context (task_context.py:11-41)."""

    task_idx: int
    sub_progress_bar_dict: Optional[Dict[str, Any]] = None
    upstream_map_transformer: Optional[Any] = None
    upstream_map_ray_remote_args: Optional[Dict[str, Any]] = None
    target_max_block_size: Optional[int] = None

    def sub_bar(self, name: str) -> Any:
        if self.sub_progress_bar_dict is None:
            self.sub_progress_bar_dict = {}
        return self.sub_progress_bar_dict.setdefault(name, _NullProgressBar(name))


AllToAllTransformFn = Callable[[List[RefBundle], TaskContext], Tuple[List[RefBundle], StatsDict]]


class ExchangeTaskSpec:
    """Map and reduce halves of an exchange; subclasses implement the static `map` / `reduce`
    and pass their extra arguments as `map_args` / `reduce_args` (interfaces.py:14-69)."""

    MAP_SUB_PROGRESS_BAR_NAME = "Shuffle Map"
    REDUCE_SUB_PROGRESS_BAR_NAME = "Shuffle Reduce"

    def __init__(self, map_args: List[Any] = None, reduce_args: List[Any] = None):
        self._map_args = map_args or []
        self._reduce_args = reduce_args or []
        assert isinstance(self._map_args, list)
        assert isinstance(self._reduce_args, list)

    @staticmethod
    def map(idx: int, block: Block, output_num_blocks: int) -> List[Union[BlockMetadata, Block]]:
        """One input block -> output_num_blocks blocks followed by the input's BlockMetadata."""
        raise NotImplementedError

    @staticmethod
    def reduce(*mapper_outputs: Block, partial_reduce: bool = False) -> Tuple[Block, BlockMetadata]:
        """The j-th output of every mapper -> one block and its metadata."""
        raise NotImplementedError


class ExchangeTaskScheduler:
    """Runs an ExchangeTaskSpec over all input blocks (interfaces.py:72-130)."""

    def __init__(self, exchange_spec: ExchangeTaskSpec):
        self._exchange_spec = exchange_spec
        self.warn_on_driver_memory_usage_bytes: Optional[int] = (
            DataContext.get_current().warn_on_driver_memory_usage_bytes
        )

    def execute(self, refs: List[RefBundle], output_num_blocks: int, task_ctx: TaskContext,
                map_ray_remote_args: Optional[Dict[str, Any]] = None,
                reduce_ray_remote_args: Optional[Dict[str, Any]] = None,
                **kwargs) -> Tuple[List[RefBundle], StatsDict]:
        raise NotImplementedError

    def warn_on_driver_memory_usage(self, memory_usage_bytes: int, log_str: str) -> None:
        import logging

        if self.warn_on_driver_memory_usage_bytes is None:
            return
        if memory_usage_bytes > self.warn_on_driver_memory_usage_bytes:
            logging.getLogger("ray.data").warning(log_str)
            self.warn_on_driver_memory_usage_bytes = memory_usage_bytes * 2

    def warn_on_high_local_memory_store_usage(self) -> None:
        # no driver-side object store exists on the device path
        return
