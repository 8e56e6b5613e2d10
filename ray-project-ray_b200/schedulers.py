"""Exchange schedulers over device blocks.

PullBasedShuffleTaskScheduler / PushBasedShuffleTaskScheduler keep the task structure of
python/ray/data/_internal/planner/exchange/pull_based_shuffle_task_scheduler.py:17-154 and
push_based_shuffle_task_scheduler.py:28-835 (M map tasks, merge rounds, R reduce tasks), with
each "task" a call of the spec's map / reduce on HBM-resident blocks in this process, in
submission order. They exist so that any ExchangeTaskSpec keeps working and so that the
reference's scheduler tests carry over.

GpuExchangeScheduler is the B200-first execution of the same exchange: all input blocks are
treated as one segmented HBM buffer and the whole map -> exchange -> reduce is ONE device pass
per operator (one stable radix argsort + one gather for sort; one bijection gather for
random_shuffle; one argsort + segmented reduce for aggregate); output blocks are views cut at
the partition boundaries. Results are identical to the task-structured schedulers; only the
number of passes over HBM differs (2x row bytes instead of 4-6x).
"""
from __future__ import annotations

import math
from typing import Any, Dict, List, Optional, Tuple

import numpy as np
import torch

from . import ops
from .aggregate import aggregate_sorted, dense_aggregate
from .aggregate_task_spec import SortAggregateTaskSpec
from .block import (Block, BlockAccessor, BlockExecStats, BlockMetadata, DeviceBlock, DeviceBlockAccessor,
                    _as_device_block, _contig, boundary_codes, default_device, find_partition_indices)
from .interfaces import ExchangeTaskScheduler, ExchangeTaskSpec, RefBundle, StatsDict, TaskContext
from .shuffle_task_spec import ShuffleTaskSpec
from .sort_task_spec import SortKey, SortTaskSpec

CALLER_MEMORY_USAGE_PER_OBJECT_REF = 3000


def _input_blocks(refs: List[RefBundle]) -> List[DeviceBlock]:
    out = []
    for ref_bundle in refs:
        for block, _ in ref_bundle.blocks:
            out.append(_as_device_block(block))
    return out


def _bundle(block: DeviceBlock, meta: BlockMetadata, owned: bool) -> RefBundle:
    return RefBundle([(block, meta)], owns_blocks=owned)


class PullBasedShuffleTaskScheduler(ExchangeTaskScheduler):
    """MapReduce: every map output is materialised, then reducer j takes column j of the M x R matrix."""

    def execute(self, refs: List[RefBundle], output_num_blocks: int, task_ctx: TaskContext,
                map_ray_remote_args: Optional[Dict[str, Any]] = None,
                reduce_ray_remote_args: Optional[Dict[str, Any]] = None,
                _debug_limit_execution_to_num_blocks: Optional[int] = None) -> Tuple[List[RefBundle], StatsDict]:
        input_blocks_list = _input_blocks(refs)
        input_num_blocks = len(input_blocks_list)
        input_owned = all(b.owns_blocks for b in refs)
        self.warn_on_driver_memory_usage(
            input_num_blocks * output_num_blocks * CALLER_MEMORY_USAGE_PER_OBJECT_REF,
            "Execution is estimated to use a large amount of driver memory for block references.")
        spec = self._exchange_spec
        map_bar = task_ctx.sub_bar(ExchangeTaskSpec.MAP_SUB_PROGRESS_BAR_NAME)
        reduce_bar = task_ctx.sub_bar(ExchangeTaskSpec.REDUCE_SUB_PROGRESS_BAR_NAME)

        if _debug_limit_execution_to_num_blocks is not None:
            input_blocks_list = input_blocks_list[:_debug_limit_execution_to_num_blocks]
        shuffle_map_out, shuffle_map_metadata = [], []
        for i, block in enumerate(input_blocks_list):
            res = spec.map(i, block, output_num_blocks, *spec._map_args)
            shuffle_map_metadata.append(res[-1])
            shuffle_map_out.append(res[:-1])
        if _debug_limit_execution_to_num_blocks is not None:
            while len(shuffle_map_out) < output_num_blocks:
                shuffle_map_out.append(shuffle_map_out[0][:])
        map_bar.fetch_until_complete(shuffle_map_metadata)

        if _debug_limit_execution_to_num_blocks is not None:
            output_num_blocks = _debug_limit_execution_to_num_blocks
        new_blocks, new_metadata = [], []
        for j in range(output_num_blocks):
            block, meta = spec.reduce(*spec._reduce_args, *[m[j] for m in shuffle_map_out])
            new_blocks.append(block)
            new_metadata.append(meta)
        del shuffle_map_out
        reduce_bar.fetch_until_complete(new_metadata)

        output = [_bundle(b, m, input_owned) for b, m in zip(new_blocks, new_metadata)]
        return output, {"map": shuffle_map_metadata, "reduce": new_metadata}


class _MergeTaskSchedule:
    """Which reducers each merge task of a round feeds: R reducers are dealt out contiguously over
    N mergers, the first R mod N mergers taking one extra (push_based...:28-109)."""

    def __init__(self, output_num_blocks: int, num_merge_tasks_per_round: int):
        self.output_num_blocks = output_num_blocks
        self.num_merge_tasks_per_round = num_merge_tasks_per_round
        self.num_reducers_per_merger, self._num_mergers_with_extra_reducer = divmod(
            output_num_blocks, num_merge_tasks_per_round)
        if self.num_reducers_per_merger == 0:
            # fewer reducers than mergers: one reducer each, drop the idle mergers
            self.num_merge_tasks_per_round = self._num_mergers_with_extra_reducer
            self.num_reducers_per_merger = 1
            self._num_mergers_with_extra_reducer = 0

    def __repr__(self):
        return (f"    num merge tasks per round: {self.num_merge_tasks_per_round}\n"
                f"    num reduce tasks per merge task: {self.num_reducers_per_merger}\n"
                f"    num merge tasks with extra reduce task: {self._num_mergers_with_extra_reducer}")

    def _first_reducer(self, merge_idx: int) -> int:
        extra = min(merge_idx, self._num_mergers_with_extra_reducer)
        return merge_idx * self.num_reducers_per_merger + extra

    def get_num_reducers_per_merge_idx(self, merge_idx: int) -> int:
        assert merge_idx < self.num_merge_tasks_per_round
        return self.num_reducers_per_merger + (1 if merge_idx < self._num_mergers_with_extra_reducer else 0)

    def get_merge_idx_for_reducer_idx(self, reducer_idx: int) -> int:
        big = self.num_reducers_per_merger + 1
        boundary = big * self._num_mergers_with_extra_reducer
        if reducer_idx < boundary:
            merge_idx = reducer_idx // big
        else:
            merge_idx = self._num_mergers_with_extra_reducer + (reducer_idx - boundary) // self.num_reducers_per_merger
        assert merge_idx < self.num_merge_tasks_per_round
        return merge_idx

    def round_robin_reduce_idx_iterator(self):
        """Reducer indices interleaved across mergers: first reducer of every merger, then the
        second of every merger, ... so consecutive submissions land on different mergers."""
        emitted, k = 0, 0
        while emitted < self.output_num_blocks:
            for merge_idx in range(self.num_merge_tasks_per_round):
                if k < self.get_num_reducers_per_merge_idx(merge_idx):
                    yield self._first_reducer(merge_idx) + k
                    emitted += 1
            k += 1


class _PushBasedShuffleStage:
    def __init__(self, output_num_blocks: int, num_rounds: int, num_map_tasks_per_round: int,
                 merge_task_placement: List[str]):
        self.num_rounds = num_rounds
        self.num_map_tasks_per_round = num_map_tasks_per_round
        self._merge_task_placement = list(merge_task_placement)
        self._merge_task_options = [{"node_id": node_id, "soft": True} for node_id in merge_task_placement]
        self.merge_schedule = _MergeTaskSchedule(output_num_blocks, len(merge_task_placement))

    def get_merge_task_options(self, merge_idx: int) -> Dict[str, Any]:
        return self._merge_task_options[merge_idx]

    def get_estimated_num_refs(self) -> int:
        inter = self.num_rounds * self.num_map_tasks_per_round * self.merge_schedule.num_merge_tasks_per_round
        return self.num_rounds * self.num_map_tasks_per_round + inter + self.merge_schedule.output_num_blocks

    def __repr__(self):
        return (f"num map tasks per round (num args per merge task): {self.num_map_tasks_per_round}\n"
                f"num rounds (num args per reduce task): {self.num_rounds}\n"
                f"merge task placement: \n{self.merge_schedule}")


def _get_num_cpus_per_node_map() -> Dict[str, int]:
    """The "nodes" of the device path are the GPUs of the box; each offers its SM-side task slots
    as 3 (one map+map+merge group per device at the default merge factor)."""
    n = torch.cuda.device_count() if torch.cuda.is_available() else 1
    return {f"gpu{i}": 3 for i in range(max(n, 1))}


class PushBasedShuffleTaskScheduler(ExchangeTaskScheduler):
    """Exoshuffle-style map -> merge -> reduce in rounds (push_based_shuffle_task_scheduler.py:400-657).

    Round r maps M consecutive input blocks; merger m partially reduces, for each of its reducers,
    the r-th round's M map outputs; reducer j finally reduces its one merged block per round.
    Mapper order is preserved inside a round and rounds are consecutive mapper ranges, so sort
    ties keep global (block, row) order.
    """

    def execute(self, refs: List[RefBundle], output_num_blocks: int, task_ctx: TaskContext,
                map_ray_remote_args: Optional[Dict[str, Any]] = None,
                reduce_ray_remote_args: Optional[Dict[str, Any]] = None, merge_factor: float = 2,
                _debug_limit_execution_to_num_blocks: Optional[int] = None) -> Tuple[List[RefBundle], StatsDict]:
        input_blocks_list = _input_blocks(refs)
        input_owned = all(b.owns_blocks for b in refs)
        spec = self._exchange_spec
        stage = self._compute_shuffle_schedule(_get_num_cpus_per_node_map(), len(input_blocks_list), merge_factor,
                                               output_num_blocks)
        sched = stage.merge_schedule
        self.warn_on_driver_memory_usage(
            stage.get_estimated_num_refs() * CALLER_MEMORY_USAGE_PER_OBJECT_REF,
            "Execution is estimated to use a large amount of driver memory for block references.")
        map_bar = task_ctx.sub_bar(ExchangeTaskSpec.MAP_SUB_PROGRESS_BAR_NAME)
        merge_bar = task_ctx.sub_bar("Shuffle Merge")
        reduce_bar = task_ctx.sub_bar(ExchangeTaskSpec.REDUCE_SUB_PROGRESS_BAR_NAME)

        if _debug_limit_execution_to_num_blocks is not None:
            input_blocks_list = input_blocks_list[:_debug_limit_execution_to_num_blocks]
            stage.num_rounds = min(stage.num_rounds, 1)

        map_meta: List[BlockMetadata] = []
        merge_meta: List[BlockMetadata] = []
        # merged[m][r] = list of P blocks (one per reducer of merger m) produced in round r
        merged: List[List[List[DeviceBlock]]] = [[] for _ in range(sched.num_merge_tasks_per_round)]
        M = stage.num_map_tasks_per_round
        for rnd in range(stage.num_rounds):
            lo = rnd * M
            round_inputs = input_blocks_list[lo:lo + M]
            if not round_inputs:
                break
            per_mapper = []  # per mapper: list (per merger) of lists (per reducer) of blocks
            for i, block in enumerate(round_inputs):
                parts = list(self._map_partition(spec.map, lo + i, block, output_num_blocks, sched, *spec._map_args))
                map_meta.append(parts[-1])
                per_mapper.append(parts[:-1])
            map_bar.update(len(round_inputs))
            for m in range(sched.num_merge_tasks_per_round):
                outs = list(self._merge(spec.reduce, *[pm[m] for pm in per_mapper], reduce_args=spec._reduce_args))
                merge_meta.append(outs[-1])
                merged[m].append(outs[:-1])
            merge_bar.update(sched.num_merge_tasks_per_round)

        new_blocks: Dict[int, DeviceBlock] = {}
        new_meta: Dict[int, BlockMetadata] = {}
        for reduce_idx in sched.round_robin_reduce_idx_iterator():
            m = sched.get_merge_idx_for_reducer_idx(reduce_idx)
            k = reduce_idx - sched._first_reducer(m)
            args = [round_out[k] for round_out in merged[m]]
            block, meta = spec.reduce(*spec._reduce_args, *args)
            new_blocks[reduce_idx], new_meta[reduce_idx] = block, meta
        reduce_bar.update(len(new_blocks))

        order = sorted(new_blocks)
        output = [_bundle(new_blocks[j], new_meta[j], input_owned) for j in order]
        stats = {"map": map_meta, "merge": merge_meta, "reduce": [new_meta[j] for j in order]}
        return output, stats

    @staticmethod
    def _map_partition(map_fn, idx: int, block: Block, output_num_blocks: int, schedule: _MergeTaskSchedule,
                       *map_args: Any):
        """Run the map and hand each merger the contiguous run of outputs its reducers own."""
        mapper_outputs = map_fn(idx, block, output_num_blocks, *map_args)
        meta = mapper_outputs[-1]
        outs = mapper_outputs[:-1]
        pos = 0
        for merge_idx in range(schedule.num_merge_tasks_per_round):
            p = schedule.get_num_reducers_per_merge_idx(merge_idx)
            yield outs[pos:pos + p]
            pos += p
        assert pos == len(outs), (pos, len(outs))
        assert isinstance(meta, BlockMetadata)
        yield meta

    @staticmethod
    def _merge(reduce_fn, *all_mapper_outputs: List[Block], reduce_args: Optional[List[Any]] = None):
        """Partial reduce of one round: for each reducer slot, reduce what every mapper sent."""
        assert len({len(mo) for mo in all_mapper_outputs}) == 1, "Received different number of map inputs"
        stats = BlockExecStats.builder()
        reduce_args = reduce_args or []
        num_rows = size_bytes = 0
        schema = None
        for mapper_outputs in zip(*all_mapper_outputs):
            block, _ = reduce_fn(*reduce_args, *mapper_outputs, partial_reduce=True)
            yield block
            acc = BlockAccessor.for_block(block)
            num_rows += acc.num_rows()
            size_bytes += acc.size_bytes()
            schema = acc.schema()
        yield BlockMetadata(num_rows=num_rows, size_bytes=size_bytes, schema=schema, input_files=None,
                            exec_stats=stats.build())

    @staticmethod
    def _compute_shuffle_schedule(num_cpus_per_node_map: Dict[str, int], num_input_blocks: int, merge_factor: float,
                                  num_output_blocks: int) -> _PushBasedShuffleStage:
        """Choose how many merge tasks run per round and where (push_based...:738-821).

        Task slots are grouped as `merge_factor` map tasks + 1 merge task. Each node gets
        round(slots / group) merge tasks (never more than the ceil(inputs / merge_factor) the job
        needs); slots left over on several nodes are pooled into extra groups; the remaining slots
        run map tasks."""
        num_cpus_total = sum(num_cpus_per_node_map.values())
        group = merge_factor + 1
        merges_needed = math.ceil(num_input_blocks / merge_factor)

        placement: List[str] = []
        leftover = 0
        inputs_left = num_input_blocks
        for node, cpus in num_cpus_per_node_map.items():
            slots = min(cpus, inputs_left)
            here = min(round(slots / group), merges_needed)
            for _ in range(here):
                placement.append(node)
                inputs_left -= merge_factor
                cpus -= group
            leftover += cpus
            if leftover >= group and len(placement) < merges_needed:
                placement.append(node)
                leftover -= group
                inputs_left -= merge_factor
            inputs_left = max(0, inputs_left)

        if not placement:
            for node, cpus in num_cpus_per_node_map.items():
                if cpus >= 1:
                    placement.append(node)
                    break
        assert len(placement) > 0, placement

        maps_per_round = max(min(num_cpus_total - len(placement), num_input_blocks), 1)
        num_rounds = math.ceil(num_input_blocks / maps_per_round)
        return _PushBasedShuffleStage(num_output_blocks, num_rounds, maps_per_round, placement)


class GpuExchangeScheduler(ExchangeTaskScheduler):
    """One fused device pass per all-to-all operator (see module docstring).

    Same `execute` signature and StatsDict shape as the reference schedulers; `merge_factor` is
    accepted for signature parity. Specs it does not recognise run through the pull-based
    task structure."""

    def execute(self, refs: List[RefBundle], output_num_blocks: int, task_ctx: TaskContext,
                map_ray_remote_args: Optional[Dict[str, Any]] = None,
                reduce_ray_remote_args: Optional[Dict[str, Any]] = None, merge_factor: float = 2,
                _debug_limit_execution_to_num_blocks: Optional[int] = None) -> Tuple[List[RefBundle], StatsDict]:
        spec = self._exchange_spec
        # One fused pass handles what one sort call and one GPU's memory hold: more rows than `fused_max_rows`
        # (default: the argsort's limit) go through the task-structured scheduler, which works block by block.
        from . import distributed as D
        from .context import DataContext

        total_rows = sum(m.num_rows or 0 for rb in refs for _, m in rb.blocks)
        too_large = total_rows > int(DataContext.get_current().get_config("fused_max_rows", ops.MAX_SORT_ROWS)) \
            and not D.is_distributed() and type(spec) in (SortTaskSpec, SortAggregateTaskSpec)
        if too_large or _debug_limit_execution_to_num_blocks is not None or type(spec) not in (
                SortTaskSpec, ShuffleTaskSpec, SortAggregateTaskSpec):
            if type(spec) in (SortTaskSpec, SortAggregateTaskSpec) and isinstance(spec._map_args[0], AutoBoundaries):
                # the task-structured path needs real splitters: sample them now
                key = spec._map_args[1]
                sort_key = key if isinstance(key, SortKey) else SortKey(key)
                bounds = SortTaskSpec.sample_boundaries(_input_blocks(refs), sort_key, output_num_blocks)
                if sort_key.get_descending():
                    bounds.reverse()
                spec._map_args[0] = bounds
            return PullBasedShuffleTaskScheduler(spec).execute(
                refs, output_num_blocks, task_ctx, map_ray_remote_args, reduce_ray_remote_args,
                _debug_limit_execution_to_num_blocks=_debug_limit_execution_to_num_blocks)
        blocks = _input_blocks(refs)
        input_owned = all(b.owns_blocks for b in refs)
        in_meta = [m for rb in refs for _, m in rb.blocks]
        stats = BlockExecStats.builder()
        if D.is_distributed():
            outs = _distributed_exchange(spec, blocks, output_num_blocks)
            st = stats.build()
            out_meta = [DeviceBlockAccessor(b).get_metadata(None, exec_stats=st) for b in outs]
            return [_bundle(b, m, input_owned) for b, m in zip(outs, out_meta)], {"map": in_meta, "reduce": out_meta}
        if type(spec) is SortTaskSpec:
            boundaries, sort_key = spec._map_args
            outs = fused_sort(blocks, boundaries, sort_key)
        elif type(spec) is ShuffleTaskSpec:
            _, upstream_map_fn, random_shuffle, seed = spec._map_args
            if upstream_map_fn is not None:
                mapped = []
                for b in blocks:
                    mapped.extend(_as_device_block(x) for x in upstream_map_fn([b]))
                blocks = mapped
            outs = fused_shuffle(blocks, output_num_blocks, random_shuffle, seed)
        else:
            boundaries, key, aggs = spec._map_args
            outs = fused_aggregate(blocks, boundaries, key, aggs)
        st = stats.build()
        out_meta = [DeviceBlockAccessor(b).get_metadata(None, exec_stats=st) for b in outs]
        output = [_bundle(b, m, input_owned) for b, m in zip(outs, out_meta)]
        return output, {"map": in_meta, "reduce": out_meta}


class AutoBoundaries(list):
    """Marker passed as `boundaries` when the scheduler may choose the partition cuts itself:
    `len()` is R-1 like a real boundary list. The fused scheduler then cuts the globally sorted
    result into R near-equal blocks at tie-respecting positions, which needs no sampling pass
    (block boundaries are load-balance hints in the reference, not part of its result contract)."""

    def __init__(self, num_outputs: int):
        super().__init__([None] * max(num_outputs - 1, 0))


def _cut_sorted_evenly(table: DeviceBlock, key: str, descending: bool, parts: int) -> List[DeviceBlock]:
    """Cut a sorted block into `parts` near-equal blocks without splitting a run of equal keys
    (equal keys must land in one block: util.py:756-758, tests/test_sort.py:79-94)."""
    n = table.num_rows
    if parts <= 1 or n == 0:
        return [table] + [DeviceBlock.empty(table.device) for _ in range(parts - 1)]
    cuts = ops.even_cuts(table.key_column(key), descending, parts).cpu().tolist()
    return _cut(table, [int(c) for c in cuts])


def _distributed_exchange(spec: ExchangeTaskSpec, blocks: List[DeviceBlock], output_num_blocks: int) -> List[DeviceBlock]:
    """One process per GPU (torch.distributed initialised): this rank's blocks are its shard; the
    operator runs across all ranks and this rank keeps its key range / slice, cut into as many
    output blocks as it had input blocks."""
    from . import distributed as D

    parts = max(len(blocks), 1)
    local = DeviceBlock.concat(blocks)
    dev = local.device if local.device is not None else default_device()
    local = D.adopt_schema(local, dev)   # a rank whose blocks are all empty takes part with zero rows
    if local is None:
        return [DeviceBlock.empty(dev) for _ in range(parts)]
    if type(spec) is SortTaskSpec:
        _, sort_key = spec._map_args
        if any(local.validity(c) is not None for c in sort_key.get_columns()):
            raise NotImplementedError("nullable sort keys are not supported across ranks")
        boundaries = spec._map_args[0]
        if not isinstance(boundaries, AutoBoundaries) and len(boundaries) > 0:
            # user boundaries (already in partition order): the partitions are dealt to the ranks in order, every rank
            # sorts its key range and cuts it at the boundaries that fall inside it
            import torch.distributed as dist

            first, desc = sort_key.get_columns()[0], sort_key.get_descending()
            codes = boundary_codes(boundaries, local.key_column(first), desc).cpu().numpy().view(np.uint64)
            splitters, edges = D.rank_splitters_from_boundaries(codes, dist.get_world_size(), desc)
            out = D.dist_sort(local, sort_key, splitter_codes=splitters)
            rank = dist.get_rank()
            mine = list(boundaries[edges[rank]:edges[rank + 1] - 1]) if edges[rank + 1] > edges[rank] else []
            if edges[rank + 1] == edges[rank]:
                return []
            if not mine or out.num_rows == 0:
                return [out] + [DeviceBlock.empty(dev) for _ in range(len(mine))]
            return _cut(out, find_partition_indices(out.key_column(first), mine, desc))
        out = D.dist_sort(local, sort_key)
        return _cut_sorted_evenly(out, sort_key.get_columns()[0], sort_key.get_descending(), parts)
    if type(spec) is ShuffleTaskSpec:
        _, upstream_map_fn, random_shuffle, seed = spec._map_args
        if upstream_map_fn is not None or not random_shuffle:
            raise NotImplementedError("only random_shuffle runs across ranks")
        out = D.dist_random_shuffle(local, seed)
        n = out.num_rows
        return _cut(out, [(i * n) // parts for i in range(1, parts)])
    _, key, aggs = spec._map_args
    k = key[0] if isinstance(key, list) and len(key) == 1 else key
    out = D.dist_aggregate(local, k, aggs)
    if k is None or (isinstance(k, list) and not k) or not out.columns:
        return [out]
    # output blocks are cut on the first key column (rows equal in it stay together)
    return _cut_sorted_evenly(out, k[0] if isinstance(k, list) else k, False, parts)


class _AsyncCuts:
    """even_cuts of a sorted key column with an asynchronous device -> host read of the positions."""

    _pinned: Dict[int, torch.Tensor] = {}

    def __init__(self, descending: bool, parts: int):
        self.descending, self.parts = descending, parts
        self._event = None
        self._host = None

    def start(self, sorted_keys: torch.Tensor) -> None:
        dev = sorted_keys.device
        cuts = ops.even_cuts(_contig(sorted_keys), self.descending, self.parts)
        host = self._pinned.get(self.parts)
        if host is None or host.numel() != cuts.numel():
            host = torch.empty(cuts.numel(), dtype=torch.int64).pin_memory()
            self._pinned[self.parts] = host
        host.copy_(cuts, non_blocking=True)
        self._event = torch.cuda.Event()
        self._event.record(torch.cuda.current_stream(dev))
        self._host, self._keep = host, cuts

    def result(self) -> List[int]:
        self._event.synchronize()
        return [int(c) for c in self._host.tolist()]


def _cut(table: DeviceBlock, cuts: List[int]) -> List[DeviceBlock]:
    parts, last = [], 0
    for idx in cuts:
        idx = max(idx, last)
        parts.append(table.slice(last, idx))
        last = idx
    parts.append(table.slice(last, table.num_rows))
    return parts


def fused_sort(blocks: List[DeviceBlock], boundaries: List[Any], sort_key: SortKey) -> List[DeviceBlock]:
    """Global stable sort of the concatenated blocks, cut at the (already direction-ordered)
    boundaries: equal to SortTaskSpec.map on every block + SortTaskSpec.reduce per range, because
    reducers concatenate in mapper order and re-sort stably (arrow_block.py:554-567)."""
    R = len(boundaries) + 1
    table = DeviceBlock.concat(blocks)
    if table.num_rows == 0:
        return [DeviceBlock.empty(table.device) for _ in range(R)]
    acc = DeviceBlockAccessor(table)
    first = sort_key.get_columns()[0]
    if table.validity(first) is not None:
        # nulls sort last whatever the direction, so the sorted key column is not monotone: no key-based cuts
        if not isinstance(boundaries, AutoBoundaries) and R > 1:
            raise NotImplementedError("user boundaries on a nullable sort key are not supported on device blocks")
        out, _ = acc.sort(sort_key)
        n = out.num_rows
        return _cut(out, [(i * n) // R for i in range(1, R)])
    if isinstance(boundaries, AutoBoundaries) and R > 1:
        # The cut positions only need the sorted keys: they are computed and read back while the payload
        # gather is still running, so the host is already enqueuing the next operator when the gather ends.
        cuts = _AsyncCuts(sort_key.get_descending(), R)
        out, _ = acc.sort(sort_key, on_sorted_key=cuts.start)
        return _cut(out, cuts.result())
    out, _ = acc.sort(sort_key)
    if R == 1:
        return [out]
    if isinstance(boundaries, AutoBoundaries):
        return _cut_sorted_evenly(out, first, sort_key.get_descending(), R)
    if any(b is None or (isinstance(b, tuple) and b[0] is None) for b in boundaries):
        return [out] + [DeviceBlock.empty(table.device) for _ in range(R - 1)]
    cuts = find_partition_indices(out.key_column(first), boundaries, sort_key.get_descending())
    return _cut(out, cuts)


def shuffle_output_sizes(block_rows: List[int], R: int) -> List[int]:
    """Rows each reducer receives under ShuffleTaskSpec.map's slicing (ceil(n/R)-row slices per
    mapper; the slice -> reducer assignment only permutes which reducer gets which slice, so the
    multiset of sizes per mapper is fixed). The fused path deals sizes out in slice order."""
    sizes = [0] * R
    for n in block_rows:
        sz = max(1, math.ceil(n / R))
        for j in range(R):
            sizes[j] += max(0, min(n, (j + 1) * sz) - j * sz)
    return sizes


def fused_shuffle(blocks: List[DeviceBlock], R: int, random_shuffle: bool, seed: Optional[int]) -> List[DeviceBlock]:
    """random_shuffle: ONE seeded bijection over all rows, out[j] = in[pi(j)] (a single gather
    pass, 2x row bytes of traffic), cut into R blocks of the sizes the reference's slicing gives.
    Non-random shuffle (repartition): rows keep their order and are re-cut."""
    table = DeviceBlock.concat(blocks)
    n = table.num_rows
    if n == 0:
        return [DeviceBlock.empty(table.device) for _ in range(R)]
    if random_shuffle:
        table = DeviceBlockAccessor(table).random_shuffle(seed)
        sizes = shuffle_output_sizes([b.num_rows for b in blocks], R)
    else:
        # ShuffleTaskSpec without random_shuffle: reducer j = concat over mappers of slice j
        return _repartition_slices(blocks, R)
    cuts, acc = [], 0
    for s in sizes[:-1]:
        acc += s
        cuts.append(acc)
    return _cut(table, cuts)


def _repartition_slices(blocks: List[DeviceBlock], R: int) -> List[DeviceBlock]:
    outs = []
    per_block = []
    for b in blocks:
        n = b.num_rows
        sz = max(1, math.ceil(n / R))
        per_block.append([b.slice(j * sz, (j + 1) * sz) for j in range(R)])
    for j in range(R):
        outs.append(DeviceBlock.concat([pb[j] for pb in per_block]))
    return outs


def fused_aggregate(blocks: List[DeviceBlock], boundaries: List[Any], key, aggs) -> List[DeviceBlock]:
    """One argsort by key + one segmented reduction per aggregated column over ALL rows, output
    cut at the boundaries: equal to combine per (block, range) + aggregate_combined_blocks per
    range, without materialising partial accumulators."""
    R = len(boundaries) + 1 if key is not None else 1
    used = SortAggregateTaskSpec._prune_unused_columns
    blocks = [used(b, key, aggs) for b in blocks]
    table = DeviceBlock.concat(blocks)
    n = table.num_rows
    if n == 0:
        return [DeviceBlock.empty(table.device) for _ in range(R)]
    k = key[0] if isinstance(key, list) and len(key) == 1 else key
    if k is None or (isinstance(k, list) and len(k) == 0):
        return [aggregate_sorted(table, None, aggs, None, None, finalize=True, n=n)]
    if isinstance(k, list):
        from .aggregate import _argsort_keys

        perm, _ = _argsort_keys(table, k, want_codes=False)
        out = aggregate_sorted(table, k, aggs, perm, None, finalize=True, n=n)
        k = k[0]  # output blocks are cut on the first key column (rows equal in it stay together)
    else:
        out = dense_aggregate(table, k, aggs, finalize=True)     # small-range integer keys: no sort needed
        if out is None:
            perm, codes, _ = ops.argsort(_contig(table.columns[k]), False, want_codes=True)
            out = aggregate_sorted(table, k, aggs, perm, codes, finalize=True, n=n)
    if R == 1:
        return [out]
    if isinstance(boundaries, AutoBoundaries):
        return _cut_sorted_evenly(out, k, False, R)
    if any(b is None or (isinstance(b, tuple) and b[0] is None) for b in boundaries):
        return [out] + [DeviceBlock.empty(table.device) for _ in range(R - 1)]
    cuts = find_partition_indices(out.columns[k], boundaries, False)
    return _cut(out, cuts)
