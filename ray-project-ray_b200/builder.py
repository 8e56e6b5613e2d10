"""DeviceBlockBuilder: concatenates device blocks (reference: DelegatingBlockBuilder,
python/ray/data/_internal/delegating_block_builder.py, as used by ShuffleTaskSpec.reduce and
SortTaskSpec.sample_boundaries)."""
from __future__ import annotations

from typing import List

from .block import DeviceBlock, _as_device_block


class DeviceBlockBuilder:
    def __init__(self):
        self._blocks: List[DeviceBlock] = []

    def add_block(self, block) -> None:
        self._blocks.append(_as_device_block(block))

    def num_rows(self) -> int:
        return sum(b.num_rows for b in self._blocks)

    def build(self) -> DeviceBlock:
        return DeviceBlock.concat(self._blocks)


DelegatingBlockBuilder = DeviceBlockBuilder
