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

import os
from typing import Dict, List, Optional, Sequence, Tuple

import numpy as np
import torch
import torch.distributed as dist

from . import ops
from .block import BYTE_KEY_SEP, DeviceBlock, DeviceBlockAccessor, _contig, record_groups
from .sort_task_spec import SortKey


def _transport() -> str:
    """"auto" (peer-memory stores when every rank can map every other, else NCCL), "p2p" or "nccl"."""
    from .context import DataContext

    import os

    return str(DataContext.get_current().get_config("exchange_transport", os.environ.get("B200X_EXCHANGE_TRANSPORT", "auto")))


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
    row_bytes = send.element_size()
    for d in send.shape[1:]:
        row_bytes *= int(d)
    as_bytes = send.reshape(-1).view(torch.uint8).reshape(send.shape[0], row_bytes)   # zero rows included
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


# ------------------------------------------------------- peer-memory (NVLink P2P) transport
_symm_cache: Dict[int, Tuple[torch.Tensor, object]] = {}
_p2p_ok: Dict[str, Optional[bool]] = {"state": None}
_ROW_ROUND = 1 << 20


def _symm_buffer(slot: int, nbytes: int, dev: torch.device, group):
    """A receive buffer that every rank can store into: torch symmetric memory (cuMem + peer mapping). ONE buffer
    per slot (= position of the column in an exchange), reused by later exchanges and replaced by a larger one
    (at least 1.5x) when an exchange needs more: the memory held is bounded by the largest exchange seen, not
    by the number of different sizes. Every rank asks for the same sizes (they derive from the global count
    matrix), so the rendezvous of a replacement is collective."""
    import torch.distributed._symmetric_memory as symm

    have = _symm_cache.get(slot)
    if have is None or have[0].numel() < nbytes:
        size = nbytes if have is None else max(nbytes, have[0].numel() * 3 // 2)
        size = (size + (1 << 21) - 1) >> 21 << 21
        _symm_cache.pop(slot, None)
        t = symm.empty(size, dtype=torch.uint8, device=dev)
        h = symm.rendezvous(t, group if group is not None else dist.group.WORLD)
        _symm_cache[slot] = (t, h)
    return _symm_cache[slot]


def p2p_available(dev: torch.device, group=None) -> bool:
    """True when every rank can map every other rank's memory (one NVSwitch box). Decided once,
    collectively: if any rank fails to set up symmetric memory, all ranks use the NCCL transport."""
    if _p2p_ok["state"] is None:
        ok = 1
        try:
            _symm_buffer(-1, 4096, dev, group)
        except Exception as e:  # noqa: BLE001  (driver / container without peer mapping support)
            import logging

            logging.getLogger("ray.data").warning("peer-memory exchange unavailable (%s); using NCCL all-to-all", e)
            ok = 0
        flag = torch.tensor([ok], dtype=torch.int32, device=dev)
        dist.all_reduce(flag, op=dist.ReduceOp.MIN, group=group)
        _p2p_ok["state"] = bool(flag.item())
    return bool(_p2p_ok["state"])


def chunk_layout_from_tables(tables: np.ndarray, rank: int) -> Tuple[List[int], List[int]]:
    """Where the chunks of a chunked exchange land in rank `rank`'s receive buffers. `tables[r, c, d]` = rows among
    source rank r's first c chunks that go to rank d (`tables[r, K, d]` = all of r's rows for d). Buffers are filled
    in source-rank order, every source's rows in row order, so chunk c of source r fills the slots
    [base_r + tables[r, c, rank], base_r + tables[r, c + 1, rank]) with base_r = rows of the lower source ranks.
    Returns (starts, sizes): starts[r * K + c] = first slot of chunk c of source r (ascending); sizes[c] = rows of
    chunk c over all sources."""
    world, K = tables.shape[0], tables.shape[1] - 1
    starts, sizes = [], [0] * K
    for r in range(world):
        base = int(tables[:r, K, rank].sum())
        for c in range(K):
            starts.append(base + int(tables[r, c, rank]))
            sizes[c] += int(tables[r, c + 1, rank] - tables[r, c, rank])
    return starts, sizes


class P2PExchange:
    """Fused gather + exchange over peer memory. `columns` maps name -> (source column, row order or
    None). Row i of the send order (row order[i] of the source) goes to the rank that owns it, written
    by ONE kernel per column directly into that rank's receive buffer over NVLink (csrc/exchange.cu);
    no send buffer, no NCCL copy. Columns can be sent in groups (`send`) separated by `fence`, so
    that work on the first group (sorting the received keys) overlaps the transfer of the rest.
    `view` returns tensors that ALIAS the receive buffers: consume them before the next exchange."""

    def __init__(self, columns: Dict[str, Tuple[torch.Tensor, Optional[torch.Tensor]]], n: int,
                 send_counts: Optional[Sequence[int]], dev: torch.device, group=None, target: Optional[torch.Tensor] = None,
                 chunk_table: Optional[torch.Tensor] = None, recv_row_bytes: Optional[Dict[str, int]] = None):
        import ctypes

        self.columns, self.n, self.dev, self.group = columns, n, dev, group
        self.target = target  # per-row owner<<28|slot (partition-first mode): rows are read in order
        self.world, rank = dist.get_world_size(group), dist.get_rank(group)
        self.rank = rank
        if chunk_table is not None:
            # chunk_table[c, d] (int64, on the device) = rows among this rank's first chunk_rows[c] rows that go to rank
            # d (last row = all rows): gathered from every rank - the one device -> host read of the exchange - it
            # gives the count matrix and tells a receiver which slots of its buffer chunk c of rank r fills
            parts = [torch.empty_like(chunk_table) for _ in range(self.world)]
            dist.all_gather(parts, chunk_table.contiguous(), group=group)
            self.tables = torch.stack(parts).cpu().numpy()[:, :, :self.world]   # [source rank, chunk boundary, destination]
            C = self.tables[:, -1, :]
            send_counts = [int(x) for x in C[rank]]
        else:
            self.tables = None
            C = all_gather_counts(send_counts, group, dev)
        self.C = C
        self.n_recv = int(C[:, rank].sum())
        cap_rows = int(C.sum(axis=0).max())
        cap_rows = max(_ROW_ROUND, (cap_rows + _ROW_ROUND - 1) // _ROW_ROUND * _ROW_ROUND)
        send_off = np.concatenate(([0], np.cumsum(np.asarray(send_counts, dtype=np.int64))))
        row_off = C[:rank, :].sum(axis=0)  # rows that lower ranks put before mine, per destination
        self._so = (ctypes.c_ulonglong * (self.world + 1))(*[int(x) for x in send_off])
        self._ro = (ctypes.c_ulonglong * self.world)(*[int(x) for x in row_off])
        self._bufs = {}
        for slot, (name, (col, _)) in enumerate(columns.items()):
            row_bytes = col.element_size()
            for d in col.shape[1:]:
                row_bytes *= int(d)
            row_bytes = (recv_row_bytes or {}).get(name, row_bytes)   # records arrive without padding / key prefix
            self._bufs[name] = (_symm_buffer(slot, cap_rows * row_bytes, dev, group), row_bytes)
        self._handle = next(iter(self._bufs.values()))[0][1]

    def send_records(self, name: str, copy_bytes: int, src_offset: int, row0: int, nrows: int) -> None:
        """Rows [row0, row0 + nrows) of the record unit `name`: bytes [src_offset, src_offset + copy_bytes) of every
        record go to the row's slot in the owner's receive buffer. Slots lie copy_bytes apart (dense): the rows an
        owner receives from one sender are consecutive slots, so they leave as whole 128-byte lines - NVLink moved a
        line per store however few of its bytes were written (5.3 G lines/s measured at 8 GPUs, whether 96 or 128 bytes
        of each 128-byte slot were sent), so the padding must not exist at the destination either."""
        col, _ = self.columns[name]
        (buf, h), pitch = self._bufs[name]
        ops.scatter_records_by_target(col, self.target, [int(p) for p in h.buffer_ptrs], [int(x) for x in self._ro], pitch,
                                      copy_bytes=copy_bytes, src_offset=src_offset, row0=row0, nrows=nrows)

    def chunk_layout(self) -> Tuple[List[int], List[int]]:
        """(starts, sizes) of the chunks in this rank's receive buffers: see `chunk_layout_from_tables`."""
        return chunk_layout_from_tables(self.tables, self.rank)

    def fence(self) -> None:
        """Device-side barrier over all ranks on the current stream (symmetric-memory signal pads): before
        the first send = everyone is done reading the previous contents; after a send = all stores landed."""
        self._handle.barrier(channel=0)

    def send(self, names: Sequence[str]) -> None:
        import ctypes

        from . import _lib

        with torch.cuda.device(self.dev):
            st = torch.cuda.current_stream(self.dev).cuda_stream
            for name in names:
                col, order = self.columns[name]
                (buf, h), row_bytes = self._bufs[name]
                bases = (ctypes.c_void_p * self.world)(*[int(p) for p in h.buffer_ptrs])
                if self.target is not None:
                    _lib.call("b200x_scatter_rows_by_target", _contig(col).data_ptr(), row_bytes, self.target.data_ptr(),
                              self.n, bases, self._ro, self.world, st)
                    continue
                _lib.call("b200x_scatter_rows_to_peers", _contig(col).data_ptr(), row_bytes,
                          None if order is None else order.data_ptr(), self.n, self._so, bases, self._ro, self.world, st)

    def view(self, names: Sequence[str]) -> Dict[str, torch.Tensor]:
        out = {}
        for name in names:
            col, _ = self.columns[name]
            (buf, _), row_bytes = self._bufs[name]
            flat = buf[: self.n_recv * row_bytes]
            if col.dim() == 2 and col.dtype == torch.uint8 and row_bytes != col.shape[1]:   # records received dense
                out[name] = flat.reshape(self.n_recv, row_bytes)
                continue
            shape = (self.n_recv,) + tuple(col.shape[1:])
            out[name] = flat.view(col.dtype).reshape(shape) if col.dtype != torch.uint8 else flat.reshape(shape)
        return out


_side_streams: Dict[Tuple[int, int], torch.cuda.Stream] = {}


def _side_stream(dev: torch.device, which: int = 0) -> torch.cuda.Stream:
    idx = dev.index if dev.index is not None else torch.cuda.current_device()
    if (idx, which) not in _side_streams:
        _side_streams[(idx, which)] = torch.cuda.Stream(device=dev)
    return _side_streams[(idx, which)]


# ------------------------------------------------------------------------ device steps
def _sample_codes(codes_sorted: torch.Tensor, num_samples: int) -> torch.Tensor:
    n = codes_sorted.numel()
    if n == 0:
        return codes_sorted[:0]
    m = min(num_samples, n)
    pos = ((np.arange(m, dtype=np.float64) + 0.5) * (n / m)).astype(np.int64).astype(np.int32)
    return ops.gather(codes_sorted, torch.from_numpy(pos).to(codes_sorted.device))


def rank_splitters_from_boundaries(boundary_codes: np.ndarray, world: int, descending: bool) -> Tuple[np.ndarray, List[int]]:
    """User boundaries across ranks (planner/sort.py:47-54: the R = len(boundaries) + 1 output partitions are fixed by
    the user): rank r owns the partitions [edges[r], edges[r+1]), edges[j] = j * R // W, so the splitter between rank r
    and r + 1 is the boundary that separates partition edges[r+1] - 1 from edges[r+1]. `boundary_codes`: the
    boundaries' codes in partition order (ascending as codes for both directions). A row whose key equals a boundary
    belongs to the partition on its right when ascending (`x < b` goes left, util.py:745-758) and to the one on its left
    when descending (`x >= b` goes left): there the splitter is the boundary's code + 1. Returns (W-1 splitter codes,
    edges)."""
    R = len(boundary_codes) + 1
    edges = [(j * R) // world for j in range(world + 1)]
    out = np.zeros(world - 1, dtype=np.uint64)
    top = np.uint64(0xFFFFFFFFFFFFFFFF)
    for r in range(world - 1):
        i = edges[r + 1] - 1
        if i < 0:
            out[r] = 0                      # this rank and the ones before it own no partition: everything goes right
        else:
            c = np.uint64(boundary_codes[i])
            out[r] = (c + np.uint64(1) if c != top else top) if descending else c
    return out, edges


def dist_sort(local: DeviceBlock, sort_key: SortKey, group=None, samples_per_rank: int = 4096,
              splitter_codes: Optional[np.ndarray] = None) -> DeviceBlock:
    """Globally sort the rows held by all ranks; rank r ends up with the r-th key range. `splitter_codes` (W-1 ascending
    uint64 codes, the same on every rank) fixes the ranges instead of sampling them."""
    world = dist.get_world_size(group)
    dev = local.device
    acc = DeviceBlockAccessor(local)
    n = local.num_rows
    first = sort_key.get_columns()[0]
    desc = sort_key.get_descending()
    use_p2p = dev.type == "cuda" and _transport() != "nccl" and p2p_available(dev, group)
    if use_p2p and _transport() != "p2p_sort_first":
        return _p2p_partition_sort(local, sort_key, n, dev, group, samples_per_rank, splitter_codes)
    perm = skeys = None
    if n > 0 and use_p2p:
        perm, codes, skeys = acc._argsort(sort_key, want_codes=True)   # no local gather: rows leave in sort order
        sorted_local = None
        samples = _sample_codes(codes, samples_per_rank)
    elif n > 0:
        sorted_local, codes = acc.sort(sort_key, want_codes=True)
        samples = _sample_codes(codes, samples_per_rank)
    else:
        sorted_local, codes = local, torch.empty(0, dtype=torch.int64, device=dev)
        samples = codes
        if use_p2p:
            perm = torch.empty(0, dtype=torch.int32, device=dev)
    # splitters: every rank gathers every rank's samples and computes the same quantiles
    padded = torch.full((samples_per_rank + 1,), -1, dtype=torch.int64, device=dev)
    padded[0] = samples.numel()
    padded[1:1 + samples.numel()] = samples
    gathered = [torch.empty_like(padded) for _ in range(world)]
    dist.all_gather(gathered, padded, group=group)
    host = torch.stack(gathered).cpu().numpy()
    all_codes = np.concatenate([row[1:1 + int(row[0])].view(np.uint64) for row in host])
    splitters = choose_splitters(all_codes, world) if splitter_codes is None else np.asarray(splitter_codes, dtype=np.uint64)
    if n > 0 and splitters.size:
        sp = torch.from_numpy(splitters.view(np.int64)).to(dev)
        cuts = ops.search_sorted_codes(codes, sp, upper=False).cpu().tolist()
    else:
        cuts = [0] * (world - 1) if n == 0 else [n] * (world - 1)
    send_counts = split_counts_from_cuts(cuts, n)
    if use_p2p:
        return _p2p_sort_exchange(local, sort_key, perm, skeys, n, send_counts, dev, group)
    received = exchange_block(sorted_local, send_counts, group)
    if received.num_rows == 0:
        return received
    out, _ = DeviceBlockAccessor(received).sort(sort_key)
    return out


_HEADER = 8


def _gather_samples(sample_codes: torch.Tensor, samples_per_rank: int, world: int, dev, group, header: Sequence[int] = ()):
    """all_gather every rank's (possibly fewer than samples_per_rank) sample codes -> one uint64 host array, plus the
    ranks' `header` values (up to 8 int64 per rank, gathered in the same collective) as an int64 [world, 8] array."""
    padded = torch.full((_HEADER + samples_per_rank + 1,), -1, dtype=torch.int64, device=dev)
    hdr = list(header) + [0] * (_HEADER - len(header))
    padded[:_HEADER] = torch.tensor(hdr, dtype=torch.int64)
    padded[_HEADER] = sample_codes.numel()
    padded[_HEADER + 1:_HEADER + 1 + sample_codes.numel()] = sample_codes
    gathered = [torch.empty_like(padded) for _ in range(world)]
    dist.all_gather(gathered, padded, group=group)
    host = torch.stack(gathered).cpu().numpy()
    codes = np.concatenate([row[_HEADER + 1:_HEADER + 1 + int(row[_HEADER])].view(np.uint64) for row in host])
    return codes, host[:, :_HEADER]


def _densify(block: DeviceBlock) -> DeviceBlock:
    """The same block with every column dense (records dissolved, no key prefix columns)."""
    return DeviceBlock({k: _contig(v) for k, v in block.columns.items()}, num_rows=block.num_rows, device=block.device)


def _exchange_units(local: DeviceBlock, key_cols: Sequence[str]):
    """What travels in a sort exchange, in sending order: the key columns `key_cols` (dense; a `<col>::be<i>` chunk
    name is the block's key prefix column or is packed here), then every record group as ONE column of whole
    records, then the remaining columns. Returns (units: name -> dense tensor, groups, names of the non-key units)."""
    groups, _ = record_groups(local)
    in_group = {k for g in groups for k in g.members}
    units: Dict[str, torch.Tensor] = dict(zip(key_cols, local.key_columns(key_cols)))
    rest = []
    for i, g in enumerate(groups):
        units[f"\x00rec{i}"] = g.rec
        rest.append(f"\x00rec{i}")
    for c in local.column_names:
        if c not in in_group and c not in units:
            units[c] = _contig(local.columns[c])
            rest.append(c)
    return units, groups, rest


def _received_block(local: DeviceBlock, groups, recv: Dict[str, torch.Tensor], n_recv: int, dev) -> DeviceBlock:
    """The received rows (arrival order) as a block of the sender's columns: views of the receive buffers."""
    cols: Dict[str, torch.Tensor] = {}
    for i, g in enumerate(groups):
        cols.update(g.views(recv[f"\x00rec{i}"]))
    for c in local.column_names:
        if c not in cols and c in recv:
            cols[c] = recv[c]
    return DeviceBlock({c: cols[c] for c in local.column_names if c in cols}, num_rows=n_recv, device=dev)


def _assemble_sorted(local: DeviceBlock, groups, recv: Dict[str, torch.Tensor], perm2: torch.Tensor, first: str, skeys2,
                     n_recv: int, dev) -> DeviceBlock:
    """The received rows in sorted order: records are gathered whole and re-cut into their member columns."""
    in_group = {k for g in groups for k in g.members}
    out_cols: Dict[str, torch.Tensor] = {}
    for i, g in enumerate(groups):
        out_cols.update(g.views(ops.gather_records(recv[f"\x00rec{i}"], perm2)))
    decoded = skeys2 is not None and first in local.columns and first not in in_group
    names = [c for c in local.column_names if c not in in_group and not (decoded and c == first)]
    outs = ops.gather_columns([recv[c] for c in names], perm2)
    out_cols.update(zip(names, outs))
    if decoded:
        out_cols[first] = skeys2
    aux = {first: skeys2} if (skeys2 is not None and first not in local.columns and first.endswith(f"{BYTE_KEY_SEP}0")) else {}
    return DeviceBlock({c: out_cols[c] for c in local.column_names}, num_rows=n_recv, device=dev, aux=aux)


def _p2p_partition_sort(local: DeviceBlock, sort_key: SortKey, n: int, dev, group, samples_per_rank: int,
                        splitter_codes: Optional[np.ndarray] = None) -> DeviceBlock:
    """Partition-first distributed sort over peer memory:
      sample raw keys -> identical splitters on every rank -> per-row (owner, slot) targets (stable split,
      no sort) -> key columns stored into the owners' receive buffers -> while the payload columns follow
      over NVLink, the received keys are sorted on a second stream -> one gather of the received payload.
    The payload is read once sequentially on the sender and gathered once on the receiver.

    Later key columns that are chunks of a byte-string key held inside the records (gensort: `key::be1`) do not
    travel as columns of their own: the first column alone is exchanged and sorted early, and once the records
    have landed the few runs of equal first keys are put in order from the key bytes in the records (`fix_ties`)."""
    world = dist.get_world_size(group)
    key_cols = sort_key.get_columns()
    first, desc = key_cols[0], sort_key.get_descending()
    plan = _record_exchange_plan(local, key_cols, desc) if os.environ.get("B200X_RECORD_EXCHANGE", "1") != "0" else None
    has_groups = bool(record_groups(local)[0])
    keycol = local.key_column(first)
    if n > 0:
        m = min(samples_per_rank, n)
        pos = ((np.arange(m, dtype=np.float64) + 0.5) * (n / m)).astype(np.int64).astype(np.int32)
        sample_codes = ops.encode_keys(ops.gather(keycol, torch.from_numpy(pos).to(dev)), desc)
    else:
        sample_codes = torch.empty(0, dtype=torch.int64, device=dev)
    # every rank must walk through the same protocol: the layout-dependent choices are agreed on in the collective
    # that gathers the samples. A rank without rows cannot tell how the others lay out their records; if there is
    # one, or the layouts differ, all ranks exchange plain dense columns.
    header = [int(n > 0), int(plan is not None)] + ([plan["group"].pitch, plan["used"], plan["shift"], plan["copy"]] if plan else [0] * 4) \
        + [int(has_groups)]
    all_codes, hdrs = _gather_samples(sample_codes, samples_per_rank, world, dev, group, header)
    splitters = choose_splitters(all_codes, world) if splitter_codes is None else np.asarray(splitter_codes, dtype=np.uint64)
    if splitters.size != world - 1:   # no rank has any row
        splitters = np.full(world - 1, np.iinfo(np.uint64).max, dtype=np.uint64)
    everyone_has_rows = bool(hdrs[:, 0].all())
    if everyone_has_rows and bool(hdrs[:, 1].all()) and (hdrs[:, 2:6] == hdrs[0, 2:6]).all():
        return _p2p_partition_sort_records(local, sort_key, n, dev, group, keycol, splitters, plan)
    if hdrs[:, 6].any() and not (everyone_has_rows and hdrs[:, 6].all()):
        local = _densify(local)
    late = 1 < len(key_cols) <= 1 + ops.FIX_TIES_MAX_COLS and all(
        c not in local.columns and c not in local.aux and BYTE_KEY_SEP in c for c in key_cols[1:])
    early = [first] if late else list(key_cols)
    units, groups, rest = _exchange_units(local, early)
    keycol = units[first]
    sp = torch.from_numpy(splitters.view(np.int64)).to(dev)
    target, counts = ops.partition_targets(keycol, desc, sp)
    send_counts = [int(c) for c in counts.cpu().tolist()[:world]]
    cols = {name: (t, None) for name, t in units.items()}
    ex = P2PExchange(cols, n, send_counts, dev, group, target=target)
    main = torch.cuda.current_stream(dev)
    ex.fence()
    ex.send(early)
    ex.fence()
    if ex.n_recv == 0:
        ex.send(rest)
        ex.fence()
        return local.slice(0, 0)
    keys_block = DeviceBlock(ex.view(early), num_rows=ex.n_recv, device=dev)
    side = _side_stream(dev)
    landed, sorted_ev = torch.cuda.Event(), torch.cuda.Event()
    landed.record(main)
    # The payload transfer is link-bound: two blocks per SM keep NVLink busy and leave the rest of every SM to
    # the sort running next to it (measured at 2 GPUs: sort || send 6.3 ms, against 10.5 ms with one block per
    # tile and 9.4 ms back to back). It is enqueued first: the sort below may synchronise the host with its own
    # stream (multi-column keys probe whether the first column alone decides the order).
    from . import _lib

    _lib.call("b200x_set_tuning", 1, int(os.environ.get("B200X_PAYLOAD_BLOCKS_PER_SM", "2")))
    try:
        ex.send(rest)
    finally:
        _lib.call("b200x_set_tuning", 1, 0)
    ex.fence()
    with torch.cuda.stream(side):
        side.wait_event(landed)
        if late:
            perm2, _, skeys2 = ops.argsort(keys_block.columns[first], desc, want_keys=True)
        else:
            perm2, _, skeys2 = DeviceBlockAccessor(keys_block)._argsort(sort_key, False)
        sorted_ev.record(side)
    main.wait_event(sorted_ev)
    perm2.record_stream(main)
    if skeys2 is not None:
        skeys2.record_stream(main)
    recv = ex.view(list(cols))
    verdict = None
    if late:
        from .block import _TieVerdict

        received = _received_block(local, groups, recv, ex.n_recv, dev)
        verdict = _TieVerdict(ops.fix_ties(skeys2, perm2, [received.key_piece(c) for c in key_cols[1:]], desc))
    out = _assemble_sorted(local, groups, recv, perm2, first, skeys2, ex.n_recv, dev)
    if verdict is not None and not verdict.ok():     # long runs of equal first keys: sort the received rows by all columns
        received.aux[first] = keys_block.columns[first]
        perm2, _, skeys2 = DeviceBlockAccessor(received)._argsort_chain(sort_key, False)
        out = _assemble_sorted(local, groups, recv, perm2, first, skeys2, ex.n_recv, dev)
    return out


def _record_exchange_plan(local: DeviceBlock, key_cols: Sequence[str], desc: bool):
    """Can this sort use the record exchange (one record group that moves in chunks, the key prefix elided)?
    Needs: exactly one record group; every key column after the first is a chunk of a byte-string member of it
    (compared where it lies in the received rows). Returns None or a dict with the group, the bytes in use, and
    `shift` = 8 when the first 8 bytes of every record ARE the first key column as the sort writes it out (the
    big-endian prefix of a byte-string key at offset 0: those bytes then travel only once, as the key column)."""
    from .block import _inner

    groups, _ = record_groups(local)
    if len(groups) != 1 or local.num_rows == 0:
        return None
    g = groups[0]
    first = key_cols[0]
    for c in key_cols[1:]:
        base, sep, idx = c.rpartition(BYTE_KEY_SEP)
        if c in local.columns or not (sep and idx.isdigit() and base in g.members):
            return None
    if len(key_cols) > 1 + ops.FIX_TIES_MAX_COLS:
        return None
    used = max(off + _inner(local.columns[k])[0] for k, (off, _, _) in g.members.items())
    base, sep, idx = first.rpartition(BYTE_KEY_SEP)
    shift, be = 0, True
    if first not in local.columns and sep and idx == "0" and base in g.members and g.members[base][0] == 0 \
            and local.columns[base].shape[1] >= 8:
        shift = 8
    elif first not in local.columns:
        return None
    copy = (used - shift + 15) // 16 * 16
    if copy + shift > g.pitch and shift:
        return None
    return {"group": g, "used": used, "shift": shift, "be": be, "copy": min(copy, g.pitch), "base": base if shift else None}


def _p2p_partition_sort_records(local: DeviceBlock, sort_key: SortKey, n: int, dev, group, keycol: torch.Tensor, splitters: np.ndarray,
                                plan) -> DeviceBlock:
    """Partition-first distributed sort of a block held as records, with the transfer and the last gather overlapped:
      1. first key column (dense) -> splitters -> per-row (owner, slot) targets, as in `_p2p_partition_sort`;
      2. the key column is exchanged and the received keys are sorted on a second stream;
      3. the records follow in K chunks of rows (`b200x_scatter_records_by_target`: without padding, and without
         their first 8 bytes when those are the key prefix that went ahead in step 2), each chunk closed by a
         device-side barrier of all ranks;
      4. as soon as the keys are sorted, every chunk that has landed is gathered into its final rows on a third
         stream (`b200x_gather_records_prefixed` restricted to that chunk's slots, the key prefix put back from the
         sorted keys) while the next chunk is still crossing NVLink: only the last chunk's gather is exposed;
      5. positions whose first key equals a neighbour's are left out until everything has landed: `fix_ties` orders
         them from the key bytes in the received rows, then one small pass gathers them."""
    from . import _lib
    from .block import _TieVerdict

    world, rank = dist.get_world_size(group), dist.get_rank(group)
    key_cols = sort_key.get_columns()
    first, desc = key_cols[0], sort_key.get_descending()
    g, used, shift, copy = plan["group"], plan["used"], plan["shift"], plan["copy"]
    pitch = g.pitch
    in_group = set(g.members)
    extra = [c for c in local.column_names if c not in in_group and c != first]   # columns outside the record: sent whole, gathered last
    sp = torch.from_numpy(splitters.view(np.int64)).to(dev)
    target, counts, tile_table = ops.partition_targets(keycol, desc, sp, want_tile_table=True)
    tiles = tile_table.shape[0]
    # chunks: measured at 8 GPUs x 125 M gensort records: 1 -> 30.6 ms, 2 -> 30.4, 4 -> 31.8, 8 -> 33.6 per sort (the
    # gathers that run next to the transfer slow it by about what they hide; every chunk ends in a barrier of all ranks)
    K = int(os.environ.get("B200X_EXCHANGE_CHUNKS", "2")) if n >= int(os.environ.get("B200X_EXCHANGE_CHUNK_MIN_ROWS", 1 << 22)) else 1
    K = max(1, min(K, tiles))
    cut_tiles = [c * tiles // K for c in range(K)]
    rows_at = [t * ops.PT_TILE_ROWS for t in cut_tiles] + [n]
    sel = torch.tensor(cut_tiles, dtype=torch.int64, device=dev)
    table = torch.cat([tile_table.index_select(0, sel).to(torch.int64), counts.reshape(1, 16)], 0)
    cols = {first: (keycol, None), "\x00rec": (g.rec, None)}
    for c in extra:
        cols[c] = (_contig(local.columns[c]), None)
    ex = P2PExchange(cols, n, None, dev, group, target=target, chunk_table=table, recv_row_bytes={"\x00rec": copy})
    main = torch.cuda.current_stream(dev)
    ex.fence()
    ex.send([first])
    ex.fence()
    if ex.n_recv == 0:
        ex.send(extra)
        for c in range(K):
            ex.send_records("\x00rec", copy, shift, rows_at[c], rows_at[c + 1] - rows_at[c])
            ex.fence()
        return local.slice(0, 0)
    n_recv = ex.n_recv
    recv_keys = ex.view([first])[first]
    side, gstream = _side_stream(dev), _side_stream(dev, 1)
    landed, sorted_ev = torch.cuda.Event(), torch.cuda.Event()
    landed.record(main)
    out = torch.empty((n_recv, pitch), dtype=torch.uint8, device=dev)
    # the transfer runs on two blocks per SM (it is link-bound) so that the sort and the gathers fit next to it
    _lib.call("b200x_set_tuning", 1, int(os.environ.get("B200X_PAYLOAD_BLOCKS_PER_SM", "2")))
    _lib.call("b200x_set_tuning", 4, int(os.environ.get("B200X_SCATTER_BULK", "1")))
    chunk_ev = []
    try:
        ex.send(extra)
        # (wider later chunks, once the received-key sort is over, made the step SLOWER at 8 GPUs: 32.6 ms with 8
        # blocks per SM, 34.3 with one block per tile, against 29.0 with two blocks per SM throughout)
        late_bps = int(os.environ.get("B200X_PAYLOAD_BLOCKS_PER_SM_LATE", os.environ.get("B200X_PAYLOAD_BLOCKS_PER_SM", "2")))
        for c in range(K):
            if c > 0:
                _lib.call("b200x_set_tuning", 1, late_bps)
            ex.send_records("\x00rec", copy, shift, rows_at[c], rows_at[c + 1] - rows_at[c])
            ex.fence()
            ev = torch.cuda.Event()
            ev.record(main)
            chunk_ev.append(ev)
    finally:
        _lib.call("b200x_set_tuning", 1, 0)
    with torch.cuda.stream(side):
        side.wait_event(landed)
        perm2, codes2, skeys2 = ops.argsort(recv_keys, desc, want_keys=True, want_codes=recv_keys.dtype.is_floating_point)
        sorted_ev.record(side)
    probe = skeys2 if skeys2 is not None else codes2.view(torch.uint64)
    recv_rec = ex.view(["\x00rec"])["\x00rec"]
    later = len(key_cols) > 1
    with torch.cuda.stream(gstream):
        gstream.wait_event(sorted_ev)
        for t in (perm2, probe, out):
            t.record_stream(gstream)
        lists = [None]
        if K > 1:       # output positions grouped by the chunk their row arrives in (known as soon as the keys are sorted)
            starts, sizes = ex.chunk_layout()
            lists = ops.split_positions_by_chunk(perm2, starts, world, K, sizes)
        # while chunks are still crossing NVLink the gathers leave block slots to the transfer kernel
        gather_bps = int(os.environ.get("B200X_GATHER_BLOCKS_PER_SM", "6"))
        for c in range(K):
            gstream.wait_event(chunk_ev[c])
            _lib.call("b200x_set_tuning", 3, gather_bps if c + 1 < K else 16)
            try:
                ops.gather_records_prefixed(recv_rec, perm2, out, copy, used, probe if shift else None, plan["be"], shift,
                                            positions=lists[c])
            finally:
                _lib.call("b200x_set_tuning", 3, 16)
        verdict = None
        if later:
            pieces = []
            for c in key_cols[1:]:
                base, _, idx = c.rpartition(BYTE_KEY_SEP)
                off, _, shape = g.members[base]
                lo = off + 8 * int(idx) - shift                                   # where that chunk lies in a received row
                pieces.append(ops.KeyPiece(recv_rec, lo, min(8, int(shape[0]) - 8 * int(idx))))
            verdict = _TieVerdict(ops.fix_ties(probe, perm2, pieces, desc))
            ops.gather_records_prefixed(recv_rec, perm2, out, copy, used, probe, plan["be"], shift, tie_mode=2)
        extra_out = {}
        if extra:
            recv_extra = ex.view(extra)
            extra_out = dict(zip(extra, ops.gather_columns([recv_extra[c] for c in extra], perm2)))
        done = torch.cuda.Event()
        done.record(gstream)
    main.wait_event(done)
    perm2.record_stream(main)
    probe.record_stream(main)
    if verdict is not None and not verdict.ok():
        # long runs of equal first keys: order the received rows by all key columns, gather everything again
        chunks = {first: recv_keys}
        for c, piece in zip(key_cols[1:], pieces):
            chunks[c] = ops.pack_key_bytes(piece.col, piece.offset, piece.width)
        perm2 = None
        for i, c in enumerate(reversed(key_cols)):
            perm2, _, sk = ops.argsort(chunks[c], desc, perm_in=perm2, want_keys=(i == len(key_cols) - 1))
        probe = sk
        ops.gather_records_prefixed(recv_rec, perm2, out, copy, used, probe if shift else None, plan["be"], shift)
        if extra:
            extra_out = dict(zip(extra, ops.gather_columns([ex.view(extra)[c] for c in extra], perm2)))
    out_cols = dict(g.views(out))
    out_cols.update(extra_out)
    if first in local.columns and first not in in_group:
        out_cols[first] = probe if skeys2 is not None else ops.gather(recv_keys, perm2)
    aux = {first: probe} if (first not in local.columns and first.endswith(f"{BYTE_KEY_SEP}0") and probe is not None) else {}
    return DeviceBlock({c: out_cols[c] for c in local.column_names}, num_rows=n_recv, device=dev, aux=aux)


def _p2p_sort_exchange(local: DeviceBlock, sort_key: SortKey, perm, skeys, n: int, send_counts, dev, group) -> DeviceBlock:
    """Exchange in two groups: key columns first; while the payload columns are in flight over
    NVLink, the received keys are already being sorted on a second stream."""
    key_cols = sort_key.get_columns()
    first = key_cols[0]
    cols = {}
    for name in key_cols + [c for c in local.column_names if c not in key_cols]:
        col = local.key_column(name) if name in key_cols else _contig(local.columns[name])
        cols[name] = (skeys, None) if (skeys is not None and name == first) else (col, perm)
    rest = [c for c in cols if c not in key_cols]
    ex = P2PExchange(cols, n, send_counts, dev, group)
    main = torch.cuda.current_stream(dev)
    ex.fence()
    ex.send(key_cols)
    ex.fence()
    if ex.n_recv == 0:
        ex.send(rest)
        ex.fence()
        return DeviceBlock({k: v.clone() for k, v in ex.view(list(cols)).items()}, num_rows=0, device=dev)
    keys_block = DeviceBlock(ex.view(key_cols), num_rows=ex.n_recv, device=dev)
    side = _side_stream(dev)
    landed, sorted_ev = torch.cuda.Event(), torch.cuda.Event()
    landed.record(main)
    ex.send(rest)  # (enqueued before the sort, which may synchronise the host with its own stream)
    ex.fence()
    with torch.cuda.stream(side):
        side.wait_event(landed)
        perm2, _, skeys2 = DeviceBlockAccessor(keys_block)._argsort(sort_key, False)
        sorted_ev.record(side)
    main.wait_event(sorted_ev)
    perm2.record_stream(main)
    if skeys2 is not None:
        skeys2.record_stream(main)
    recv = ex.view(list(cols))
    names = [c for c in local.column_names if not (skeys2 is not None and c == first)]
    outs = ops.gather_columns([recv[c] for c in names], perm2)
    out_cols = dict(zip(names, outs))
    if skeys2 is not None:
        out_cols[first] = skeys2
    return DeviceBlock({c: out_cols[c] for c in local.column_names}, num_rows=ex.n_recv, device=dev)


def adopt_schema(local: DeviceBlock, dev: torch.device, group=None) -> Optional[DeviceBlock]:
    """A rank whose shard has no rows and no columns (every block empty, e.g. after a filter) still has to take part
    in the collectives of an exchange: it adopts the column names, dtypes and row shapes of the lowest rank that has
    columns and goes on with zero rows. Collective; returns None when no rank has any column."""
    world, rank = dist.get_world_size(group), dist.get_rank(group)
    flag = torch.tensor([1 if local.columns else 0], dtype=torch.int64, device=dev)
    flags = [torch.empty_like(flag) for _ in range(world)]
    dist.all_gather(flags, flag, group=group)
    have = [int(f.item()) for f in flags]
    if all(have):
        return local
    if not any(have):
        return None
    src = have.index(1)
    box = [[(k, str(v.dtype).replace("torch.", ""), tuple(v.shape[1:])) for k, v in local.columns.items()] if rank == src else None]
    dist.broadcast_object_list(box, src=dist.get_global_rank(group, src) if group is not None else src, group=group)
    if local.columns:
        return local
    return DeviceBlock({k: torch.empty((0,) + tuple(shape), dtype=getattr(torch, dt), device=dev) for k, dt, shape in box[0]},
                       num_rows=0, device=dev)


def _partial_like(partial: Optional[DeviceBlock], dev: torch.device, group=None) -> Optional[DeviceBlock]:
    """The partial-aggregate block of a rank without rows: zero rows in the schema the other ranks' partials have."""
    return adopt_schema(partial if partial is not None else DeviceBlock.empty(dev), dev, group)


def dist_random_shuffle(local: DeviceBlock, seed: Optional[int], group=None) -> DeviceBlock:
    """random_shuffle across the ranks. Over peer memory: ONE global seeded bijection, every row stored straight at its
    place in the shuffled dataset (`b200x_permute_scatter_to_peers`) - the concatenation of the ranks' results equals
    the single-GPU shuffle of the concatenated input with the same seed. Without peer memory: ShuffleTaskSpec.map /
    reduce with one mapper and one reducer per rank over NCCL all-to-all-v."""
    from .shuffle_task_spec import ShuffleTaskSpec

    world, rank = dist.get_world_size(group), dist.get_rank(group)
    dev = local.device
    if dev is not None and dev.type == "cuda" and _transport() != "nccl" and p2p_available(dev, group):
        return _p2p_random_shuffle(local, seed, dev, group)
    parts = ShuffleTaskSpec.map(rank, local, world, 0, None, True, seed)[:-1]
    send = DeviceBlock.concat(parts) if any(p.num_rows for p in parts) else local.slice(0, 0)
    received = exchange_block(send if send.columns else local.slice(0, 0), [p.num_rows for p in parts], group)
    out, _ = ShuffleTaskSpec.reduce(True, seed, received)
    return out


def _p2p_random_shuffle(local: DeviceBlock, seed: Optional[int], dev, group) -> DeviceBlock:
    world, rank = dist.get_world_size(group), dist.get_rank(group)
    n = local.num_rows
    has_groups = bool(record_groups(local)[0])
    draw = int.from_bytes(os.urandom(7), "little") if seed is None else int(seed) & ((1 << 62) - 1)
    mine = torch.tensor([n, draw, int(has_groups)], dtype=torch.int64, device=dev)
    parts = [torch.empty_like(mine) for _ in range(world)]
    dist.all_gather(parts, mine, group=group)
    info = torch.stack(parts).cpu().numpy()
    counts = [int(x) for x in info[:, 0]]
    total, first = sum(counts), sum(counts[:rank])
    use_seed = int(info[0, 1])                      # seed=None: rank 0's draw, the same on every rank
    if total == 0:
        return local.slice(0, 0)
    if total > 0xFFFFFFFF:
        raise ValueError(f"random_shuffle handles at most 2^32-1 rows in all, got {total}")
    # a rank without rows cannot know how the others lay out their records: then everybody moves dense columns
    if info[:, 2].any() and not (all(counts) and info[:, 2].all()):
        local = _densify(local)
    per = (total + world - 1) // world
    got = max(0, min(total, (rank + 1) * per) - rank * per)
    groups, loose = record_groups(local)
    units = [(f"\x00rec{i}", g.rec) for i, g in enumerate(groups)] + [(c, _contig(local.columns[c])) for c in loose]
    bufs = []
    for slot, (name, t) in enumerate(units):
        row_bytes = t.element_size()
        for d in t.shape[1:]:
            row_bytes *= int(d)
        bufs.append((_symm_buffer(slot, per * row_bytes, dev, group), row_bytes))
    handle = bufs[0][0][1]
    handle.barrier(channel=0)                       # everyone is done with the buffers' previous contents
    for (name, t), ((buf, h), row_bytes) in zip(units, bufs):
        ops.permute_scatter_to_peers(t, first, total, use_seed, per, [int(p) for p in h.buffer_ptrs], row_bytes)
    handle.barrier(channel=0)                       # every row has landed
    cols: Dict[str, torch.Tensor] = {}
    for (name, t), ((buf, _), row_bytes) in zip(units, bufs):
        flat = buf[: got * row_bytes].clone()       # the receive buffer belongs to the next exchange
        if name.startswith("\x00rec"):
            cols.update(groups[int(name[4:])].views(flat.reshape(got, row_bytes)))
        else:
            cols[name] = (flat.view(t.dtype) if t.dtype != torch.uint8 else flat).reshape((got,) + tuple(t.shape[1:]))
    return DeviceBlock({c: cols[c] for c in local.column_names}, num_rows=got, device=dev)


def _dist_aggregate_multi(local: DeviceBlock, keys: List[str], aggs, group, samples_per_rank: int) -> DeviceBlock:
    """groupby over several key columns across ranks: local partials (sorted by all key columns), exchanged by the range
    of the FIRST key column - equal first keys share an owner, so equal key tuples do - then merged by all columns."""
    from .aggregate import _argsort_keys, aggregate_sorted, merge_combined_blocks

    world = dist.get_world_size(group)
    dev = local.device
    n = local.num_rows
    partial = None
    if n > 0:
        perm, _ = _argsort_keys(local, keys, want_codes=False)
        partial = aggregate_sorted(local, keys, aggs, perm, None, finalize=False, n=n)
    partial = _partial_like(partial, dev, group)
    if partial is None:
        return DeviceBlock.empty(dev)
    pcodes = ops.encode_keys(_contig(partial.columns[keys[0]]), False)
    all_codes, _ = _gather_samples(_sample_codes(pcodes, samples_per_rank), samples_per_rank, world, dev, group)
    splitters = choose_splitters(all_codes, world)
    if partial.num_rows and splitters.size:
        cuts = ops.search_sorted_codes(pcodes, torch.from_numpy(splitters.view(np.int64)).to(dev), upper=False).cpu().tolist()
    else:
        cuts = [0] * (world - 1)
    received = exchange_block(partial, split_counts_from_cuts(cuts, partial.num_rows), group)
    if received.num_rows == 0:
        return received
    return merge_combined_blocks([received], keys, aggs, finalize=True)


def dist_aggregate(local: DeviceBlock, key: Optional[str], aggs, group=None, samples_per_rank: int = 4096) -> DeviceBlock:
    """Sort-based groupby across ranks: local partials -> range exchange of partial rows -> merge."""
    from .aggregate import aggregate_sorted, dense_aggregate, merge_combined_blocks
    from .aggregate_task_spec import SortAggregateTaskSpec

    world, rank = dist.get_world_size(group), dist.get_rank(group)
    dev = local.device
    local = SortAggregateTaskSpec._prune_unused_columns(local, key, aggs)
    if isinstance(key, list):
        return _dist_aggregate_multi(local, key, aggs, group, samples_per_rank)
    n = local.num_rows
    if key is None:
        partial = aggregate_sorted(local, None, aggs, None, None, finalize=False, n=n) if n else None
        partial = _partial_like(partial, dev, group)      # ranks without rows send nothing
        if partial is None:
            return DeviceBlock.empty(dev)
        # every rank sends its single partial row to rank 0
        counts = [partial.num_rows] + [0] * (world - 1)
        received = exchange_block(partial, counts, group)
        return merge_combined_blocks([received], None, aggs, finalize=True) if rank == 0 else DeviceBlock.empty(dev)
    partial = None
    if n > 0:
        partial = dense_aggregate(local, key, aggs, finalize=False)    # small-range integer keys: no sort needed
        if partial is None:
            perm, codes, _ = ops.argsort(_contig(local.columns[key]), False, want_codes=True)
            partial = aggregate_sorted(local, key, aggs, perm, codes, finalize=False, n=n)  # ascending by key
    partial = _partial_like(partial, dev, group)
    if partial is None:
        return DeviceBlock.empty(dev)
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
    if partial.num_rows == 0:
        cuts = [0] * (world - 1)
    send_counts = split_counts_from_cuts(cuts, partial.num_rows)
    received = exchange_block(partial, send_counts, group)
    if received.num_rows == 0:
        return received
    return merge_combined_blocks([received], key, aggs, finalize=True)
