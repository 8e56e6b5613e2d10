"""Batchers: re-chunk a stream of blocks into fixed-size batches, optionally through a local
shuffle buffer (python/ray/data/_internal/batcher.py:23-331; driven by
_internal/block_batching/util.py:74-139 `blocks_to_batches`).

Host bookkeeping follows the reference (same thresholds, same compaction rule, seed + 1 per
re-shuffle); the data operations are device ones: slices are views of HBM buffers, building a
batch is a device concat and the shuffle is the seeded bijection kernel (`b200x_permute_rows`).
"""
from __future__ import annotations

from typing import Iterator, List, Optional

from .block import DeviceBlock, DeviceBlockAccessor
from .builder import DeviceBlockBuilder

# has_batch() waits until this many times the minimum buffer is queued beyond one batch, so that
# one compaction (concat + shuffle) serves several batches (batcher.py:20)
SHUFFLE_BUFFER_COMPACTION_RATIO = 1.5


class Batcher:
    """Batches of exactly `batch_size` rows cut across block boundaries (batcher.py:53-164).
    batch_size=None yields every block as one batch."""

    def __init__(self, batch_size: Optional[int], ensure_copy: bool = False):
        self._batch_size = batch_size
        self._queue: List[DeviceBlock] = []
        self._queued_rows = 0
        self._done_adding = False
        self._ensure_copy = ensure_copy

    def add(self, block: DeviceBlock) -> None:
        if block.num_rows > 0:  # empty blocks never enter the buffer
            self._queue.append(block)
            self._queued_rows += block.num_rows

    def done_adding(self) -> None:
        self._done_adding = True

    def has_any(self) -> bool:
        return self._queued_rows > 0

    def has_batch(self) -> bool:
        return self.has_any() and (self._batch_size is None or self._queued_rows >= self._batch_size)

    def next_batch(self) -> DeviceBlock:
        assert self.has_batch() or (self._done_adding and self.has_any())
        if self._batch_size is None:
            assert len(self._queue) == 1
            block = self._queue.pop()
            self._queued_rows = 0
            return DeviceBlockAccessor(block).slice(0, block.num_rows, copy=True) if self._ensure_copy else block
        out = DeviceBlockBuilder()
        rest: List[DeviceBlock] = []
        want = self._batch_size
        pieces = 0
        for block in self._queue:
            rows = block.num_rows
            if want <= 0:
                rest.append(block)
            elif rows <= want:
                out.add_block(block)
                want -= rows
                pieces += 1
            else:
                out.add_block(block.slice(0, want))
                rest.append(block.slice(want, rows))  # consumed first by the next batch
                want = 0
                pieces += 1
        self._queue = rest
        self._queued_rows -= self._batch_size
        batch = out.build()
        if self._ensure_copy and pieces == 1:  # a single piece is still a view of its source block
            batch = DeviceBlockAccessor(batch).slice(0, batch.num_rows, copy=True)
        return batch


class ShufflingBatcher:
    """Batches drawn from a local shuffle buffer (batcher.py:167-331): queued blocks are
    concatenated with what is left of the buffer and the whole buffer is shuffled whenever the
    materialised part drops to `shuffle_buffer_min_size` rows (or no more blocks will come)."""

    def __init__(self, batch_size: Optional[int], shuffle_buffer_min_size: int, shuffle_seed: Optional[int] = None):
        if batch_size is None:
            raise ValueError("Must specify a batch_size if using a local shuffle.")
        self._batch_size = batch_size
        self._shuffle_seed = shuffle_seed
        shuffle_buffer_min_size = max(shuffle_buffer_min_size, batch_size)
        self._buffer_capacity = max(2 * shuffle_buffer_min_size, shuffle_buffer_min_size + batch_size)
        self._buffer_min_size = shuffle_buffer_min_size
        self._builder = DeviceBlockBuilder()
        self._shuffle_buffer: Optional[DeviceBlock] = None
        self._batch_head = 0
        self._done_adding = False

    def add(self, block: DeviceBlock) -> None:
        if block.num_rows > 0:
            self._builder.add_block(block)

    def done_adding(self) -> None:
        self._done_adding = True

    def has_any(self) -> bool:
        return self._buffer_size() > 0

    def has_batch(self) -> bool:
        size = self._buffer_size()
        if self._done_adding:
            return size >= self._batch_size
        return (self._materialized_buffer_size() >= self._buffer_min_size
                or size - self._batch_size >= self._buffer_min_size * SHUFFLE_BUFFER_COMPACTION_RATIO)

    def _buffer_size(self) -> int:
        return self._builder.num_rows() + self._materialized_buffer_size()

    def _materialized_buffer_size(self) -> int:
        if self._shuffle_buffer is None:
            return 0
        return max(0, self._shuffle_buffer.num_rows - self._batch_head)

    def next_batch(self) -> DeviceBlock:
        assert self.has_batch() or (self._done_adding and self.has_any())
        if self._builder.num_rows() > 0 and (self._done_adding
                                             or self._materialized_buffer_size() <= self._buffer_min_size):
            if self._shuffle_buffer is not None:
                if self._batch_head > 0:
                    self._shuffle_buffer = self._shuffle_buffer.slice(self._batch_head, self._shuffle_buffer.num_rows)
                self._builder.add_block(self._shuffle_buffer)
            merged = self._builder.build()
            self._shuffle_buffer = DeviceBlockAccessor(merged).random_shuffle(self._shuffle_seed)
            if self._shuffle_seed is not None:
                self._shuffle_seed += 1
            self._builder = DeviceBlockBuilder()
            self._batch_head = 0
        assert self._shuffle_buffer is not None
        take = min(self._batch_size, self._shuffle_buffer.num_rows)
        lo = self._batch_head
        self._batch_head += take
        return self._shuffle_buffer.slice(lo, self._batch_head)


def blocks_to_batches(blocks: Iterator[DeviceBlock], batch_size: Optional[int] = None, drop_last: bool = False,
                      shuffle_buffer_min_size: Optional[int] = None, shuffle_seed: Optional[int] = None,
                      ensure_copy: bool = False) -> Iterator[DeviceBlock]:
    """block_batching/util.py:74-139."""
    if shuffle_buffer_min_size is not None:
        batcher = ShufflingBatcher(batch_size, shuffle_buffer_min_size, shuffle_seed)
    else:
        batcher = Batcher(batch_size, ensure_copy)
    for block in blocks:
        batcher.add(block)
        while batcher.has_batch():
            yield batcher.next_batch()
    batcher.done_adding()
    while batcher.has_batch():
        yield batcher.next_batch()
    if not drop_last and batcher.has_any():
        yield batcher.next_batch()
