"""Multi-GPU exchange: one process per GPU, key-range (sort / groupby) or slice (shuffle)
sharded across the ranks of a torch.distributed group, with ONE all-to-all-v per operator.

This is what the reference's map -> (merge) -> reduce exchange through the object store
(pull_based_shuffle_task_scheduler.py:86-124, push_based_shuffle_task_scheduler.py:435-657;
blocks moved as plasma objects / object-manager gRPC chunks) becomes on one NVSwitch box:

  sort      sample keys -> all_gather -> W-1 splitters (every rank computes the same ones)
            local stable argsort; rows for rank d are a contiguous run of the sorted order
            counts all_gather; per column: gather into send order, all_to_all_single
            (ncclSend/ncclRecv pairs over NVLink, receive buffer in SOURCE-RANK order)
            local stable argsort of the received runs + gather.
            Equal keys never straddle ranks (util.py:756-758); ties keep (rank, row) order
            because the receive buffer is rank-ordered and both sorts are stable.
  groupby   local sort-based partial aggregation -> partial rows exchanged by key range ->
            merge of partials (aggregate_combined_blocks) on the owner rank.
  shuffle   the reference's two-level structure with ranks as mappers/reducers: local seeded
            permutation, equal slices dealt to ranks in a seeded order, second local permutation.

torch.distributed is the plumbing (rendezvous, NCCL communicator, stream ordering); all data
movement inside a rank is the C-ABI kernels.
"""
from __future__ import annotations

import math
from typing import Dict, List, Optional, Sequence, Tuple

import numpy as np
import torch
import torch.distributed as dist

from . import ops
from .block import DeviceBlock, DeviceBlockAccessor, _contig
from .sort_task_spec import SortKey


def is_distributed() -> bool:
    return dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1


# ----------------------------------------------------------------- host-side planning
def choose_splitters(all_sample_codes: np.ndarray, world: int) -> np.ndarray:
    """W-1 splitter codes from the union of every rank's sample codes (uint64): nearest-rank
    quantiles at i/W, like SortTaskSpec.sample_boundaries (sort_task_spec.py:184-193)."""
    codes = np.sort(np.asarray(all_sample_codes, dtype=np.uint64))
    if codes.size == 0 or world <= 1:
        return np.zeros(0, dtype=np.uint64)
    q = (np.arange(1, world) * (codes.size - 1) / world).round().astype(np.int64)
    return codes[q]


def split_counts_from_cuts(cuts: Sequence[int], n: int) -> List[int]:
    """Rows sent to each rank given the W-1 cut positions in the locally sorted order."""
    edges = [0] + [min(max(int(c), 0), n) for c in cuts] + [n]
    for i in range(1, len(edges)):
        edges[i] = max(edges[i], edges[i - 1])
    return [edges[i + 1] - edges[i] for i in range(len(edges) - 1)]


def all_gather_counts(send_counts: Sequence[int], group=None, device=None) -> np.ndarray:
    """The W x W count matrix C[src, dst] (8 x 8 int64 on one box: SURVEY.md §5 "counts first")."""
    world = dist.get_world_size(group)
    dev = device if device is not None else torch.device("cpu")
    mine = torch.tensor(list(send_counts), dtype=torch.int64, device=dev)
    out = [torch.empty_like(mine) for _ in range(world)]
    dist.all_gather(out, mine, group=group)
    return torch.stack(out).cpu().numpy()


def exchange_rows(send: torch.Tensor, send_counts: Sequence[int], recv_counts: Sequence[int], group=None) -> torch.Tensor:
    """all-to-all-v of rows (dim 0) of a contiguous [n, ...] tensor; result is ordered by source rank."""
    n_recv = int(sum(recv_counts))
    flat = send.reshape(send.shape[0], -1) if send.dim() > 1 else send.reshape(-1, 1)
    as_bytes = flat.view(torch.uint8) if flat.dtype != torch.uint8 else flat
    row_bytes = as_bytes.shape[1] if as_bytes.dim() == 2 else 1
    as_bytes = as_bytes.reshape(send.shape[0], row_bytes)
    out = torch.empty((n_recv, row_bytes), dtype=torch.uint8, device=send.device)
    dist.all_to_all_single(out, as_bytes, output_split_sizes=[int(c) for c in recv_counts],
                           input_split_sizes=[int(c) for c in send_counts], group=group)
    shape = (n_recv,) + tuple(send.shape[1:])
    return out.view(send.dtype).reshape(shape) if send.dtype != torch.uint8 else out.reshape(shape)


def exchange_block(block: DeviceBlock, send_counts: Sequence[int], group=None) -> DeviceBlock:
    """Exchange every column of a block already laid out in destination order."""
    rank = dist.get_rank(group)
    C = all_gather_counts(send_counts, group, block.device if block.device is not None and block.device.type == "cuda" else None)
    recv_counts = C[:, rank].tolist()
    cols = {k: exchange_rows(_contig(v), send_counts, recv_counts, group) for k, v in block.columns.items()}
    return DeviceBlock(cols, num_rows=int(sum(recv_counts)), device=block.device)


# ------------------------------------------------------------------------ device steps
def _sample_codes(codes_sorted: torch.Tensor, num_samples: int) -> torch.Tensor:
    n = codes_sorted.numel()
    if n == 0:
        return codes_sorted[:0]
    m = min(num_samples, n)
    pos = ((np.arange(m, dtype=np.float64) + 0.5) * (n / m)).astype(np.int64).astype(np.int32)
    return ops.gather(codes_sorted, torch.from_numpy(pos).to(codes_sorted.device))


def dist_sort(local: DeviceBlock, sort_key: SortKey, group=None, samples_per_rank: int = 4096) -> DeviceBlock:
    """Globally sort the rows held by all ranks; rank r ends up with the r-th key range."""
    world = dist.get_world_size(group)
    dev = local.device
    acc = DeviceBlockAccessor(local)
    n = local.num_rows
    first = sort_key.get_columns()[0]
    desc = sort_key.get_descending()
    if n > 0:
        sorted_local, codes = acc.sort(sort_key, want_codes=True)
        samples = _sample_codes(codes, samples_per_rank)
    else:
        sorted_local, codes = local, torch.empty(0, dtype=torch.int64, device=dev)
        samples = codes
    # splitters: every rank gathers every rank's samples and computes the same quantiles
    padded = torch.full((samples_per_rank + 1,), -1, dtype=torch.int64, device=dev)
    padded[0] = samples.numel()
    padded[1:1 + samples.numel()] = samples
    gathered = [torch.empty_like(padded) for _ in range(world)]
    dist.all_gather(gathered, padded, group=group)
    host = torch.stack(gathered).cpu().numpy()
    all_codes = np.concatenate([row[1:1 + int(row[0])].view(np.uint64) for row in host])
    splitters = choose_splitters(all_codes, world)
    if n > 0 and splitters.size:
        sp = torch.from_numpy(splitters.view(np.int64)).to(dev)
        cuts = ops.search_sorted_codes(codes, sp, upper=False).cpu().tolist()
    else:
        cuts = [0] * (world - 1) if n == 0 else [n] * (world - 1)
    send_counts = split_counts_from_cuts(cuts, n)
    received = exchange_block(sorted_local, send_counts, group)
    if received.num_rows == 0:
        return received
    out, _ = DeviceBlockAccessor(received).sort(sort_key)
    return out


def dist_random_shuffle(local: DeviceBlock, seed: Optional[int], group=None) -> DeviceBlock:
    """ShuffleTaskSpec.map / reduce with one mapper and one reducer per rank."""
    from .shuffle_task_spec import ShuffleTaskSpec

    world, rank = dist.get_world_size(group), dist.get_rank(group)
    parts = ShuffleTaskSpec.map(rank, local, world, 0, None, True, seed)[:-1]
    send = DeviceBlock.concat(parts) if any(p.num_rows for p in parts) else local.slice(0, 0)
    received = exchange_block(send if send.columns else local.slice(0, 0), [p.num_rows for p in parts], group)
    out, _ = ShuffleTaskSpec.reduce(True, seed, received)
    return out


def dist_aggregate(local: DeviceBlock, key: Optional[str], aggs, group=None, samples_per_rank: int = 4096) -> DeviceBlock:
    """Sort-based groupby across ranks: local partials -> range exchange of partial rows -> merge."""
    from .aggregate import aggregate_sorted, merge_combined_blocks
    from .aggregate_task_spec import SortAggregateTaskSpec

    world, rank = dist.get_world_size(group), dist.get_rank(group)
    dev = local.device
    local = SortAggregateTaskSpec._prune_unused_columns(local, key, aggs)
    n = local.num_rows
    if key is None:
        partial = aggregate_sorted(local, None, aggs, None, None, finalize=False, n=n) if n else DeviceBlock.empty(dev)
        # every rank sends its single partial row to rank 0
        counts = [partial.num_rows] + [0] * (world - 1)
        if not partial.columns:
            raise ValueError("global aggregation needs at least one row per rank")
        received = exchange_block(partial, counts, group)
        return merge_combined_blocks([received], None, aggs, finalize=True) if rank == 0 else DeviceBlock.empty(dev)
    if n == 0:
        raise ValueError("distributed groupby needs at least one row per rank")
    perm, codes, _ = ops.argsort(_contig(local.columns[key]), False, want_codes=True)
    partial = aggregate_sorted(local, key, aggs, perm, codes, finalize=False, n=n)  # ascending by key
    pcodes = ops.encode_keys(_contig(partial.columns[key]), False)
    samples = _sample_codes(pcodes, samples_per_rank)
    padded = torch.full((samples_per_rank + 1,), -1, dtype=torch.int64, device=dev)
    padded[0] = samples.numel()
    padded[1:1 + samples.numel()] = samples
    gathered = [torch.empty_like(padded) for _ in range(world)]
    dist.all_gather(gathered, padded, group=group)
    host = torch.stack(gathered).cpu().numpy()
    all_codes = np.concatenate([row[1:1 + int(row[0])].view(np.uint64) for row in host])
    splitters = choose_splitters(all_codes, world)
    sp = torch.from_numpy(splitters.view(np.int64)).to(dev)
    cuts = ops.search_sorted_codes(pcodes, sp, upper=False).cpu().tolist() if splitters.size else []
    send_counts = split_counts_from_cuts(cuts, partial.num_rows)
    received = exchange_block(partial, send_counts, group)
    return merge_combined_blocks([received], key, aggs, finalize=True)
