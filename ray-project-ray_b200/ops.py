"""Thin tensor-level wrappers over the C ABI.

torch is used here only to own device buffers (`torch.empty`), to hand over
`data_ptr()` / the current CUDA stream, and to switch the current device. All
arithmetic happens in libb200x.so; there is no torch or CPU fallback.
"""
from __future__ import annotations

import ctypes
from typing import List, Optional, Sequence, Tuple

import torch

from . import _lib
from ._lib import (KEY_F32, KEY_F64, KEY_I8, KEY_I16, KEY_I32, KEY_I64, KEY_U8, KEY_U16, KEY_U32, KEY_U64, VAL_F64,
                   VAL_I64)

_KEY_TYPES = {
    torch.int64: KEY_I64, torch.float64: KEY_F64, torch.int32: KEY_I32, torch.float32: KEY_F32,
    torch.int16: KEY_I16, torch.int8: KEY_I8, torch.uint8: KEY_U8,
}
for _name, _kt in (("uint64", KEY_U64), ("uint32", KEY_U32), ("uint16", KEY_U16)):
    if hasattr(torch, _name):
        _KEY_TYPES[getattr(torch, _name)] = _kt

MAX_SORT_ROWS = (1 << 30) - 1


def key_type_of(t: torch.Tensor) -> int:
    try:
        return _KEY_TYPES[t.dtype]
    except KeyError:
        raise TypeError(f"unsupported key dtype {t.dtype}; supported: {sorted(map(str, _KEY_TYPES))}") from None


def _require_cuda(*tensors: Optional[torch.Tensor]) -> torch.device:
    dev = None
    for t in tensors:
        if t is None:
            continue
        if not t.is_cuda:
            raise RuntimeError("b200x ops need CUDA tensors: the device path has no CPU fallback")
        if not t.is_contiguous():
            raise ValueError("b200x ops need contiguous tensors")
        if dev is None:
            dev = t.device
        elif t.device != dev:
            raise ValueError(f"tensors on different devices: {dev} vs {t.device}")
    if dev is None:
        raise ValueError("no tensor given")
    return dev


def _stream(dev: torch.device) -> int:
    return torch.cuda.current_stream(dev).cuda_stream


def _ptr(t: Optional[torch.Tensor]) -> Optional[int]:
    return None if t is None else t.data_ptr()


def _temp(nbytes: int, dev: torch.device) -> torch.Tensor:
    return torch.empty(max(nbytes, 256), dtype=torch.uint8, device=dev)


def argsort(keys: torch.Tensor, descending: bool = False, perm_in: Optional[torch.Tensor] = None,
            want_codes: bool = False, want_keys: bool = False, end_bit: int = 0, _debug: Optional[dict] = None):
    """Stable argsort of a 1-D key column.

    Returns (perm, codes, sorted_keys): perm is an int32 tensor whose bits are the uint32 row
    indices; codes (int64 bits of the uint64 codes) only if want_codes; sorted_keys (the key
    column in sorted order, written by the last radix pass) only if want_keys and the key
    type is an integer type (float keys must be gathered: -0.0 / NaN bits are not in the code).
    """
    dev = _require_cuda(keys, perm_in)
    if keys.dim() != 1:
        raise ValueError("keys must be 1-D")
    n = keys.numel() if perm_in is None else perm_in.numel()
    if n > MAX_SORT_ROWS:
        raise ValueError(f"one sort call handles at most {MAX_SORT_ROWS} rows, got {n}")
    kt = key_type_of(keys)
    perm_out = torch.empty(n, dtype=torch.int32, device=dev)
    codes = torch.empty(n, dtype=torch.int64, device=dev) if want_codes else None
    skeys = torch.empty(n, dtype=keys.dtype, device=dev) if (want_keys and not keys.dtype.is_floating_point) else None
    with torch.cuda.device(dev):
        st = _stream(dev)
        need = _lib.query_temp("b200x_radix_argsort", _ptr(keys), kt, int(descending), _ptr(perm_in), _ptr(perm_out),
                               _ptr(codes), _ptr(skeys), n, 0, end_bit, stream=st)
        temp = _temp(need, dev)
        nb = ctypes.c_size_t(temp.numel())
        _lib.call("b200x_radix_argsort", _ptr(keys), kt, int(descending), _ptr(perm_in), _ptr(perm_out), _ptr(codes),
                  _ptr(skeys), n, 0, end_bit, temp.data_ptr(), ctypes.byref(nb), st)
        if _debug is not None and n:
            # tests only: which algorithm the device chose (SortCtl in csrc/radix_sort.cu, third 256-byte slot of temp)
            import struct

            raw = bytes(temp[512:512 + 56].cpu().numpy())
            cand, mode, shift_a, shift_b, max_run, tmode, shift_c = struct.unpack_from("<7I", raw, 24)
            _debug.update(candidate=cand, mode=("lsd", "hybrid")[mode], table=bool(tmode), max_run=max_run,
                          shiftA=shift_a, shiftB=shift_b, shiftC=shift_c)
    return perm_out, codes, skeys


def gather(col: torch.Tensor, idx: torch.Tensor) -> torch.Tensor:
    """out[i] = col[idx[i]] along dim 0 for a contiguous fixed-width column ([n] or [n, ...])."""
    dev = _require_cuda(col, idx)
    n = idx.numel()
    out = torch.empty((n,) + tuple(col.shape[1:]), dtype=col.dtype, device=dev)
    if n == 0:
        return out
    row_bytes = col.element_size()
    for s in col.shape[1:]:
        row_bytes *= s
    with torch.cuda.device(dev):
        _lib.call("b200x_gather_rows", col.data_ptr(), row_bytes, idx.data_ptr(), n, out.data_ptr(), _stream(dev))
    return out


def gather_records(rec: torch.Tensor, idx: torch.Tensor, row_bytes: Optional[int] = None) -> torch.Tensor:
    """out[i, :row_bytes] = rec[idx[i], :row_bytes] for a record buffer `rec` ([n, pitch] uint8, pitch and row_bytes
    multiples of 16); the result has the same pitch. Pitches of 16 .. 512 bytes that are powers of two take the
    lane-cooperative record gather (b200x_gather_records_prefixed without a prefix: 5.9 against 6.2 ms per 125 M
    128-byte records), the others b200x_gather_rows_pitched."""
    dev = _require_cuda(rec, idx)
    n, pitch = idx.numel(), rec.shape[1]
    row_bytes = pitch if row_bytes is None else row_bytes
    out = torch.empty((n, pitch), dtype=torch.uint8, device=dev)
    if n:
        with torch.cuda.device(dev):
            if row_bytes == pitch and 16 <= pitch <= 512 and pitch & (pitch - 1) == 0:
                _lib.call("b200x_gather_records_prefixed", rec.data_ptr(), pitch, pitch, idx.data_ptr(), n, None, 0, 0, pitch,
                          out.data_ptr(), pitch, 0, None, 0, _stream(dev))
            else:
                _lib.call("b200x_gather_rows_pitched", rec.data_ptr(), pitch, idx.data_ptr(), n, out.data_ptr(), pitch,
                          row_bytes, _stream(dev))
    return out


def gather_pitched_column(col: torch.Tensor, idx: torch.Tensor) -> torch.Tensor:
    """Rows `idx` of a [n, w] uint8 column whose rows lie at a pitch that is a multiple of 16 (a record member), as a
    [m, w] view of a small record buffer: only the 16-byte vectors that cover the column are read."""
    n, w = col.shape
    pitch = col.stride(0) if n > 1 else (w + 15) // 16 * 16
    a = col.data_ptr() % 16
    nb = (a + w + 15) // 16 * 16
    if pitch % 16 or nb > pitch:
        raise ValueError("not a record member column")
    m = idx.numel()
    out = torch.empty((m, nb), dtype=torch.uint8, device=col.device)
    if m:
        with torch.cuda.device(col.device):
            _lib.call("b200x_gather_rows_pitched", col.data_ptr() - a, pitch, idx.data_ptr(), m, out.data_ptr(), nb, nb,
                      _stream(col.device))
    return out[:, a:a + w]


def permute_records(rec: torch.Tensor, seed: int, inverse: bool = False, row_bytes: Optional[int] = None) -> torch.Tensor:
    """out[j, :row_bytes] = rec[pi_seed(j), :row_bytes] for a record buffer ([n, pitch] uint8)."""
    dev = _require_cuda(rec)
    n, pitch = rec.shape
    row_bytes = pitch if row_bytes is None else row_bytes
    out = torch.empty_like(rec)
    if n:
        with torch.cuda.device(dev):
            _lib.call("b200x_permute_rows_pitched", rec.data_ptr(), pitch, n, seed & 0xFFFFFFFFFFFFFFFF, int(inverse),
                      out.data_ptr(), pitch, row_bytes, _stream(dev))
    return out


def copy_rows_2d(src_ptr: int, src_pitch: int, dst_ptr: int, dst_pitch: int, row_bytes: int, n: int,
                 dev: torch.device) -> None:
    """dst[r, :row_bytes] = src[r, :row_bytes] between two row pitches (raw device addresses: the caller keeps the
    tensors alive)."""
    if n and row_bytes:
        with torch.cuda.device(dev):
            _lib.call("b200x_copy_rows_2d", src_ptr, src_pitch, dst_ptr, dst_pitch, row_bytes, n, _stream(dev))


class KeyPiece:
    """`width` big-endian bytes at byte `offset` of every row of a [n, w] uint8 column (rows dense or at a pitch): one
    piece of a byte-string key, compared where it lies instead of being packed into a column of its own."""

    __slots__ = ("col", "offset", "width")

    def __init__(self, col: torch.Tensor, offset: int, width: int):
        if col.dim() != 2 or col.dtype != torch.uint8 or not (1 <= width <= 8) or offset + width > col.shape[1]:
            raise ValueError("a key piece is 1..8 bytes inside a [n, w] uint8 column")
        self.col, self.offset, self.width = col, offset, width


def fix_ties(keys_sorted: torch.Tensor, perm: torch.Tensor, cols: Sequence, descending: bool) -> torch.Tensor:
    """Order every run of equal keys in `keys_sorted` (the first key column after a stable sort by it alone; int64
    codes count as uint64) by the remaining key columns `cols` (1-D tensors, dense or strided, or `KeyPiece`s), in
    place in `perm`. Returns int64[2] on the device: rows inside tie runs, longest run (runs above 64 rows are left
    untouched)."""
    dev = _require_cuda(keys_sorted, perm)
    k = len(cols)
    ptrs, types, strides, be = (ctypes.c_void_p * k)(), (ctypes.c_int * k)(), (ctypes.c_ulonglong * k)(), (ctypes.c_int * k)()
    for i, c in enumerate(cols):
        if isinstance(c, KeyPiece):
            t = c.col
            ptrs[i], types[i], be[i] = t.data_ptr() + c.offset, KEY_U8, c.width
            strides[i] = t.stride(0) if t.shape[0] > 1 else t.shape[1]
        else:
            if c.dim() != 1 or not c.is_cuda:
                raise ValueError("key columns are 1-D CUDA tensors")
            ptrs[i], types[i], be[i] = c.data_ptr(), key_type_of(c), 0
            strides[i] = (c.stride(0) if c.shape[0] > 1 else 1) * c.element_size()
    stats = torch.empty(2, dtype=torch.int64, device=dev)
    with torch.cuda.device(dev):
        _lib.call("b200x_fix_ties", keys_sorted.data_ptr(), key_type_of(keys_sorted), keys_sorted.numel(), perm.data_ptr(),
                  ptrs, types, strides, be, k, int(descending), stats.data_ptr(), _stream(dev))
    return stats


FIX_TIES_MAX_RUN = 64
FIX_TIES_MAX_COLS = 3


def gather_columns(cols: Sequence[torch.Tensor], idx: torch.Tensor) -> List[torch.Tensor]:
    """Gather several columns with one index vector; 8-byte 1-D columns share one launch."""
    outs: List[Optional[torch.Tensor]] = [None] * len(cols)
    n = idx.numel()
    eight = [i for i, c in enumerate(cols) if c.dim() == 1 and c.element_size() == 8]
    if len(eight) >= 2 and n > 0:
        dev = _require_cuda(idx, *[cols[i] for i in eight])
        for i in eight:
            outs[i] = torch.empty(n, dtype=cols[i].dtype, device=dev)
        srcs = (ctypes.c_void_p * len(eight))(*[cols[i].data_ptr() for i in eight])
        dsts = (ctypes.c_void_p * len(eight))(*[outs[i].data_ptr() for i in eight])
        with torch.cuda.device(dev):
            _lib.call("b200x_gather_columns8", srcs, dsts, len(eight), idx.data_ptr(), n, _stream(dev))
    for i, c in enumerate(cols):
        if outs[i] is None:
            outs[i] = gather(c, idx)
    return outs  # type: ignore[return-value]


def iota(n: int, dev: torch.device, start: int = 0) -> torch.Tensor:
    out = torch.empty(n, dtype=torch.int32, device=dev)
    if n:
        with torch.cuda.device(dev):
            _lib.call("b200x_iota_u32", out.data_ptr(), n, start, _stream(dev))
    return out


def encode_keys(keys: torch.Tensor, descending: bool = False) -> torch.Tensor:
    dev = _require_cuda(keys)
    out = torch.empty(keys.numel(), dtype=torch.int64, device=dev)
    with torch.cuda.device(dev):
        _lib.call("b200x_encode_keys", keys.data_ptr(), key_type_of(keys), int(descending), keys.numel(),
                  out.data_ptr(), _stream(dev))
    return out


def count_adjacent_equal(sorted_keys: torch.Tensor) -> torch.Tensor:
    """int64[1] on the device: how many rows of a sorted key column equal their predecessor."""
    dev = _require_cuda(sorted_keys)
    out = torch.empty(1, dtype=torch.int64, device=dev)
    with torch.cuda.device(dev):
        _lib.call("b200x_count_adjacent_equal", sorted_keys.data_ptr(), key_type_of(sorted_keys), sorted_keys.numel(),
                  out.data_ptr(), _stream(dev))
    return out


def _chunk_dtype(width: int) -> torch.dtype:
    return torch.uint16 if width <= 2 else (torch.uint32 if width <= 4 else torch.uint64)


def pack_key_bytes(col: torch.Tensor, offset: int, width: int) -> torch.Tensor:
    """Bytes [offset, offset + width) of every row of a [n, w] uint8 column as a big-endian unsigned integer
    column (uint16 / uint32 / uint64 by width, left-aligned). The column may lie at a row pitch (a record member)."""
    if not col.is_cuda:
        raise RuntimeError("b200x ops need CUDA tensors: the device path has no CPU fallback")
    if col.dim() != 2 or col.dtype != torch.uint8 or (col.shape[0] > 1 and col.stride(1) != 1):
        raise TypeError("byte-string keys are [n, width] uint8 columns")
    dev = col.device
    n, w = col.shape
    pitch = col.stride(0) if n > 1 else w
    if offset + width > w:
        raise ValueError("chunk outside the key column")
    out = torch.empty(n, dtype=_chunk_dtype(width), device=dev)
    if n:
        with torch.cuda.device(dev):
            _lib.call("b200x_pack_key_bytes", col.data_ptr(), pitch, offset, width, n, out.data_ptr(), out.element_size(),
                      _stream(dev))
    return out


def byte_key_chunks(col: torch.Tensor) -> List[torch.Tensor]:
    """A [n, w] uint8 key column as ceil(w / 8) unsigned integer columns, most significant first. Columns whose rows
    lie at a multiple of 16 bytes (record members) are packed in one pass with 16-byte loads."""
    n, w = col.shape
    pitch = col.stride(0) if n > 1 else w
    if col.is_cuda and n > 0 and pitch % 16 == 0 and w <= 32 and (n == 1 or col.stride(1) == 1):
        widths = [min(8, w - off) for off in range(0, w, 8)]
        outs = [torch.empty(n, dtype=_chunk_dtype(cw), device=col.device) for cw in widths]
        ptrs = (ctypes.c_void_p * len(outs))(*[o.data_ptr() for o in outs])
        nbytes = (ctypes.c_int * len(outs))(*[o.element_size() for o in outs])
        with torch.cuda.device(col.device):
            _lib.call("b200x_pack_key_chunks", col.data_ptr(), pitch, w, n, ptrs, nbytes, _stream(col.device))
        return outs
    return [pack_key_bytes(col, off, min(8, w - off)) for off in range(0, w, 8)]


def unpack_key_bytes(chunks: Sequence[torch.Tensor], width: int) -> torch.Tensor:
    """Inverse of byte_key_chunks: the [n, width] uint8 column whose 8-byte chunks are `chunks`."""
    dev = _require_cuda(*chunks)
    n = chunks[0].numel()
    out = torch.empty((n, width), dtype=torch.uint8, device=dev)
    if n:
        with torch.cuda.device(dev):
            for i, ch in enumerate(chunks):
                off = 8 * i
                _lib.call("b200x_unpack_key_bytes", ch.data_ptr(), ch.element_size(), min(8, width - off), out.data_ptr(),
                          width, off, n, _stream(dev))
    return out


def search_sorted_codes(codes: torch.Tensor, bound_codes: torch.Tensor, upper: bool) -> torch.Tensor:
    dev = _require_cuda(codes, bound_codes)
    out = torch.empty(bound_codes.numel(), dtype=torch.int64, device=dev)
    with torch.cuda.device(dev):
        _lib.call("b200x_search_sorted_codes", codes.data_ptr(), codes.numel(), bound_codes.data_ptr(),
                  bound_codes.numel(), int(upper), out.data_ptr(), _stream(dev))
    return out


def search_sorted_keys(keys_sorted: torch.Tensor, descending: bool, bound_codes: torch.Tensor, upper: bool) -> torch.Tensor:
    """Lower/upper bound of boundary codes in a raw key column sorted in the given direction."""
    dev = _require_cuda(keys_sorted, bound_codes)
    out = torch.empty(bound_codes.numel(), dtype=torch.int64, device=dev)
    with torch.cuda.device(dev):
        _lib.call("b200x_search_sorted_keys", keys_sorted.data_ptr(), key_type_of(keys_sorted), int(descending),
                  keys_sorted.numel(), bound_codes.data_ptr(), bound_codes.numel(), int(upper), out.data_ptr(),
                  _stream(dev))
    return out


def even_cuts(keys_sorted: torch.Tensor, descending: bool, parts: int) -> torch.Tensor:
    """parts-1 tie-respecting cut positions of a sorted key column (int64 tensor on the device)."""
    dev = _require_cuda(keys_sorted)
    out = torch.empty(max(parts - 1, 0), dtype=torch.int64, device=dev)
    if parts > 1 and keys_sorted.numel() > 0:
        with torch.cuda.device(dev):
            _lib.call("b200x_even_cuts", keys_sorted.data_ptr(), key_type_of(keys_sorted), int(descending),
                      keys_sorted.numel(), parts, out.data_ptr(), _stream(dev))
    return out


PT_TILE_ROWS = 2048   # rows per tile of b200x_partition_targets (csrc/exchange.cu PT_TILE)


def partition_targets(keys: torch.Tensor, descending: bool, splitter_codes: torch.Tensor, want_tile_table: bool = False):
    """(target int32[n] = owner << 28 | slot, counts int64[16]) of a stable split of the rows by key range.
    want_tile_table: also the int32 [tiles, 16] table of slots handed out before every tile of PT_TILE_ROWS rows
    (row t = rows per owner among the first t * PT_TILE_ROWS rows) - the library's scratch after the call."""
    dev = _require_cuda(keys, splitter_codes if splitter_codes.numel() else None)
    n = keys.numel()
    target = torch.empty(n, dtype=torch.int32, device=dev)
    counts = torch.zeros(16, dtype=torch.int64, device=dev)
    with torch.cuda.device(dev):
        st = _stream(dev)
        args = (keys.data_ptr(), key_type_of(keys), int(descending), n,
                splitter_codes.data_ptr() if splitter_codes.numel() else None, splitter_codes.numel(),
                target.data_ptr(), counts.data_ptr())
        need = _lib.query_temp("b200x_partition_targets", *args, stream=st)
        temp = _temp(need, dev)
        nb = ctypes.c_size_t(temp.numel())
        _lib.call("b200x_partition_targets", *args, temp.data_ptr(), ctypes.byref(nb), st)
    if want_tile_table:
        tiles = (n + PT_TILE_ROWS - 1) // PT_TILE_ROWS
        return target, counts, temp[: tiles * 64].view(torch.int32).reshape(tiles, 16)
    return target, counts


def scatter_rows_by_target(col: torch.Tensor, target: torch.Tensor, bases: Sequence[int], row_offsets: Sequence[int]) -> None:
    """Row r of `col` -> row row_offsets[owner] + slot of the buffer at device address bases[owner]."""
    dev = _require_cuda(col, target)
    n = col.shape[0]
    row_bytes = col.element_size()
    for s in col.shape[1:]:
        row_bytes *= s
    w = len(bases)
    b = (ctypes.c_void_p * w)(*[int(x) for x in bases])
    ro = (ctypes.c_ulonglong * w)(*[int(x) for x in row_offsets])
    with torch.cuda.device(dev):
        _lib.call("b200x_scatter_rows_by_target", col.data_ptr(), row_bytes, target.data_ptr(), n, b, ro, w, _stream(dev))


def scatter_records_by_target(rec: torch.Tensor, target: torch.Tensor, bases: Sequence[int], row_offsets: Sequence[int],
                              dst_pitch: int, copy_bytes: Optional[int] = None, src_offset: int = 0, row0: int = 0,
                              nrows: Optional[int] = None) -> None:
    """Bytes [src_offset, src_offset + copy_bytes) of rows [row0, row0 + nrows) of the record buffer `rec` ([n, pitch]
    uint8) -> the start of row row_offsets[owner] + slot of the buffer at bases[owner] (rows dst_pitch apart)."""
    dev = _require_cuda(rec, target)
    n, pitch = rec.shape
    nrows = n - row0 if nrows is None else nrows
    copy_bytes = pitch if copy_bytes is None else copy_bytes
    w = len(bases)
    b = (ctypes.c_void_p * w)(*[int(x) for x in bases])
    ro = (ctypes.c_ulonglong * w)(*[int(x) for x in row_offsets])
    with torch.cuda.device(dev):
        _lib.call("b200x_scatter_records_by_target", rec.data_ptr() + row0 * pitch + src_offset, pitch, copy_bytes,
                  target.data_ptr() + 4 * row0, nrows, b, ro, dst_pitch, w, _stream(dev))


def gather_records_prefixed(recv: torch.Tensor, idx: torch.Tensor, out: torch.Tensor, copy_bytes: int, used: int,
                            sorted_keys: Optional[torch.Tensor] = None, prefix_be: bool = True, shift: int = 0,
                            tie_mode: int = 0, positions: Optional[torch.Tensor] = None) -> None:
    """out[i] = [key prefix from sorted_keys[i] (shift == 8) | recv[idx[i], :copy_bytes]], zero from byte `used` on, for
    the output positions in `positions` (all if None) / selected by `tie_mode` (b200x_gather_records_prefixed)."""
    dev = _require_cuda(recv, idx, out, sorted_keys, positions)
    with torch.cuda.device(dev):
        _lib.call("b200x_gather_records_prefixed", recv.data_ptr(), recv.shape[1], copy_bytes, idx.data_ptr(), idx.numel(),
                  _ptr(sorted_keys), int(prefix_be), shift, used, out.data_ptr(), out.shape[1], tie_mode, _ptr(positions),
                  0 if positions is None else positions.numel(), _stream(dev))


def split_positions_by_chunk(idx: torch.Tensor, chunk_starts: Sequence[int], num_sources: int, num_chunks: int,
                             chunk_sizes: Sequence[int]) -> List[torch.Tensor]:
    """Output positions 0..n-1 grouped by the chunk their row (receive slot idx[i]) arrived in: one int32 tensor per
    chunk (views of one buffer; chunk_sizes are known from the count tables)."""
    dev = _require_cuda(idx)
    n = idx.numel()
    offs, run = [], 0
    for sz in chunk_sizes:
        offs.append(run)
        run += int(sz)
    if run != n:
        raise ValueError("chunk sizes must add up to the number of rows")
    positions = torch.empty(max(n, 1), dtype=torch.int32, device=dev)
    starts = (ctypes.c_uint32 * len(chunk_starts))(*[int(x) for x in chunk_starts])
    offsets = (ctypes.c_uint32 * num_chunks)(*offs)
    with torch.cuda.device(dev):
        st = _stream(dev)
        args = (idx.data_ptr(), n, starts, num_sources, num_chunks, offsets, positions.data_ptr())
        need = _lib.query_temp("b200x_split_positions_by_chunk", *args, stream=st)
        temp = _temp(need, dev)
        nb = ctypes.c_size_t(temp.numel())
        _lib.call("b200x_split_positions_by_chunk", *args, temp.data_ptr(), ctypes.byref(nb), st)
    return [positions[o:o + int(sz)] for o, sz in zip(offs, chunk_sizes)]


def permute_rows(col: torch.Tensor, seed: int, inverse: bool = False) -> torch.Tensor:
    """out[j] = col[pi_seed(j)] — seeded permutation fused with the row gather."""
    dev = _require_cuda(col)
    out = torch.empty_like(col)
    n = col.shape[0]
    if n == 0:
        return out
    row_bytes = col.element_size()
    for s in col.shape[1:]:
        row_bytes *= s
    with torch.cuda.device(dev):
        _lib.call("b200x_permute_rows", col.data_ptr(), row_bytes, n, seed & 0xFFFFFFFFFFFFFFFF, int(inverse),
                  out.data_ptr(), _stream(dev))
    return out


def perm_indices(n: int, seed: int, dev: torch.device, inverse: bool = False, start: int = 0,
                 count: Optional[int] = None) -> torch.Tensor:
    count = n - start if count is None else count
    out = torch.empty(count, dtype=torch.int32, device=dev)
    if count:
        with torch.cuda.device(dev):
            _lib.call("b200x_perm_indices", n, seed & 0xFFFFFFFFFFFFFFFF, int(inverse), start, count, out.data_ptr(),
                      _stream(dev))
    return out


def perm_indices_host(n: int, seed: int, inverse: bool = False, start: int = 0, count: Optional[int] = None):
    """The same permutation evaluated by the library on the host (small n: slice-order shuffles)."""
    import numpy as np

    count = n - start if count is None else count
    out = np.empty(count, dtype=np.uint32)
    if count:
        _lib.call("b200x_perm_indices_host", n, seed & 0xFFFFFFFFFFFFFFFF, int(inverse), start, count,
                  out.ctypes.data_as(ctypes.c_void_p))
        _lib.launch_count -= 1  # host-only entry point
    return out


def group_boundaries(codes: torch.Tensor) -> Tuple[torch.Tensor, torch.Tensor]:
    """(group_starts int32[n] (first num_groups valid), num_groups int32[1]) of an ascending code array."""
    dev = _require_cuda(codes)
    n = codes.numel()
    starts = torch.empty(max(n, 1), dtype=torch.int32, device=dev)
    ng = torch.zeros(1, dtype=torch.int32, device=dev)
    with torch.cuda.device(dev):
        st = _stream(dev)
        need = _lib.query_temp("b200x_group_boundaries", codes.data_ptr(), n, starts.data_ptr(), ng.data_ptr(),
                               stream=st)
        temp = _temp(need, dev)
        nb = ctypes.c_size_t(temp.numel())
        _lib.call("b200x_group_boundaries", codes.data_ptr(), n, starts.data_ptr(), ng.data_ptr(), temp.data_ptr(),
                  ctypes.byref(nb), st)
    return starts, ng


def group_boundaries_multi(codes_list: Sequence[torch.Tensor]) -> Tuple[torch.Tensor, torch.Tensor]:
    """Group boundaries for a multi-column key: a group starts where any of the (sorted-order) code columns changes."""
    if len(codes_list) == 1:
        return group_boundaries(codes_list[0])
    dev = _require_cuda(*codes_list)
    n = codes_list[0].numel()
    starts = torch.empty(max(n, 1), dtype=torch.int32, device=dev)
    ng = torch.zeros(1, dtype=torch.int32, device=dev)
    ptrs = (ctypes.c_void_p * len(codes_list))(*[c.data_ptr() for c in codes_list])
    with torch.cuda.device(dev):
        st = _stream(dev)
        need = _lib.query_temp("b200x_group_boundaries_multi", ptrs, len(codes_list), n, starts.data_ptr(), ng.data_ptr(),
                               stream=st)
        temp = _temp(need, dev)
        nb = ctypes.c_size_t(temp.numel())
        _lib.call("b200x_group_boundaries_multi", ptrs, len(codes_list), n, starts.data_ptr(), ng.data_ptr(),
                  temp.data_ptr(), ctypes.byref(nb), st)
    return starts, ng


def absmax_groups(mn: torch.Tensor, mx: torch.Tensor) -> torch.Tensor:
    dev = _require_cuda(mn, mx)
    out = torch.empty_like(mn)
    if mn.numel():
        with torch.cuda.device(dev):
            _lib.call("b200x_absmax_groups", mn.data_ptr(), mx.data_ptr(), value_type_of(mn), mn.numel(), out.data_ptr(),
                      _stream(dev))
    return out


def value_type_of(t: torch.Tensor) -> int:
    if t.dtype == torch.int64:
        return VAL_I64
    if t.dtype == torch.float64:
        return VAL_F64
    raise TypeError(f"aggregation needs an int64 or float64 column, got {t.dtype}")


def segmented_reduce(starts: torch.Tensor, ng: torch.Tensor, num_groups: int, perm: Optional[torch.Tensor],
                     values: torch.Tensor, n: int, want_sum=True, want_min=False, want_max=False,
                     center: Optional[torch.Tensor] = None, weights: Optional[torch.Tensor] = None,
                     addend: Optional[torch.Tensor] = None):
    """Per-group (sum, min, max) of values[perm] over sorted runs; entries not requested are None."""
    dev = _require_cuda(starts, ng, perm, values, center, weights, addend)
    vt = value_type_of(values)
    out_dtype = torch.float64 if center is not None else values.dtype
    mk = lambda want: torch.empty(num_groups, dtype=out_dtype, device=dev) if want else None  # noqa: E731
    s, mn, mx = mk(want_sum), mk(want_min), mk(want_max)
    with torch.cuda.device(dev):
        st = _stream(dev)
        need = _lib.query_temp("b200x_segmented_reduce", starts.data_ptr(), ng.data_ptr(), _ptr(perm),
                               values.data_ptr(), vt, n, _ptr(center), _ptr(weights), _ptr(addend), _ptr(s), _ptr(mn),
                               _ptr(mx), stream=st)
        temp = _temp(need, dev)
        nb = ctypes.c_size_t(temp.numel())
        _lib.call("b200x_segmented_reduce", starts.data_ptr(), ng.data_ptr(), _ptr(perm), values.data_ptr(), vt, n,
                  _ptr(center), _ptr(weights), _ptr(addend), _ptr(s), _ptr(mn), _ptr(mx), temp.data_ptr(),
                  ctypes.byref(nb), st)
    return s, mn, mx


def group_counts(starts: torch.Tensor, ng: torch.Tensor, num_groups: int, n: int) -> torch.Tensor:
    dev = _require_cuda(starts, ng)
    out = torch.empty(num_groups, dtype=torch.int64, device=dev)
    if num_groups:
        with torch.cuda.device(dev):
            _lib.call("b200x_group_counts", starts.data_ptr(), ng.data_ptr(), num_groups, n, out.data_ptr(),
                      _stream(dev))
    return out


def group_quantile(starts: torch.Tensor, ng: torch.Tensor, num_groups: int, perm: Optional[torch.Tensor],
                   values: torch.Tensor, n: int, q: float) -> torch.Tensor:
    """Per-group q-quantile (float64); `perm` visits rows ordered by (group key, value)."""
    dev = _require_cuda(starts, ng, values)
    out = torch.empty(num_groups, dtype=torch.float64, device=dev)
    if num_groups and n:
        with torch.cuda.device(dev):
            _lib.call("b200x_group_quantile", starts.data_ptr(), ng.data_ptr(), num_groups,
                      perm.data_ptr() if perm is not None else None, values.data_ptr(), value_type_of(values), n,
                      float(q), out.data_ptr(), _stream(dev))
    return out


def finalize_groups(sums: torch.Tensor, counts: torch.Tensor, mode: int, ddof: int = 1) -> torch.Tensor:
    dev = _require_cuda(sums, counts)
    out = torch.empty(sums.numel(), dtype=torch.float64, device=dev)
    if sums.numel():
        with torch.cuda.device(dev):
            _lib.call("b200x_finalize_groups", sums.data_ptr(), value_type_of(sums), counts.data_ptr(), sums.numel(),
                      mode, ddof, out.data_ptr(), _stream(dev))
    return out


def key_code_range(keys: torch.Tensor) -> Tuple[int, int]:
    """(smallest, largest) order-preserving code of a key column, as Python ints (one device -> host read)."""
    dev = _require_cuda(keys)
    out = torch.empty(2, dtype=torch.int64, device=dev)
    with torch.cuda.device(dev):
        _lib.call("b200x_key_code_range", keys.data_ptr(), key_type_of(keys), keys.numel(), out.data_ptr(), _stream(dev))
    lo, hi = out.cpu().tolist()
    return lo & 0xFFFFFFFFFFFFFFFF, hi & 0xFFFFFFFFFFFFFFFF


def dense_group_accumulate(keys: torch.Tensor, base_code: int, rng: int, values: Optional[torch.Tensor], count: Optional[torch.Tensor],
                           want_sum: bool, want_min: bool, want_max: bool):
    """Accumulate into int64 tables of `rng` entries indexed by code(key) - base_code. `count`: the count table to add
    to, or None to leave counts out. Returns (sum, min, max) tables (None where not wanted); all freshly initialised."""
    dev = _require_cuda(keys, values, count)
    mk = lambda want: torch.empty(rng, dtype=torch.int64, device=dev) if want else None  # noqa: E731
    s, mn, mx = mk(want_sum), mk(want_min), mk(want_max)
    with torch.cuda.device(dev):
        _lib.call("b200x_dense_group_accumulate", keys.data_ptr(), key_type_of(keys), keys.numel(), base_code, rng,
                  _ptr(values), 1, _ptr(count), _ptr(s), _ptr(mn), _ptr(mx), _stream(dev))
    return s, mn, mx


def dense_group_compact(base_code: int, rng: int, key_dtype: torch.dtype, count: torch.Tensor, tabs: Sequence[torch.Tensor],
                        max_groups: int):
    """(keys, counts, compacted tables, number of groups) of the non-empty table entries in key order."""
    dev = _require_cuda(count, *tabs)
    kt = _KEY_TYPES[key_dtype]
    keys_out = torch.empty(max_groups, dtype=key_dtype, device=dev)
    count_out = torch.empty(max_groups, dtype=torch.int64, device=dev)
    outs = [torch.empty(max_groups, dtype=torch.int64, device=dev) for _ in tabs]
    ng = torch.zeros(1, dtype=torch.int32, device=dev)
    tin = (ctypes.c_void_p * max(len(tabs), 1))(*[t.data_ptr() for t in tabs])
    tout = (ctypes.c_void_p * max(len(tabs), 1))(*[t.data_ptr() for t in outs])
    with torch.cuda.device(dev):
        st = _stream(dev)
        args = (base_code, rng, kt, count.data_ptr(), len(tabs), tin, keys_out.data_ptr(), count_out.data_ptr(), tout,
                ng.data_ptr())
        need = _lib.query_temp("b200x_dense_group_compact", *args, stream=st)
        temp = _temp(need, dev)
        nb = ctypes.c_size_t(temp.numel())
        _lib.call("b200x_dense_group_compact", *args, temp.data_ptr(), ctypes.byref(nb), st)
    g = int(ng.item())
    return keys_out[:g], count_out[:g], [o[:g] for o in outs], g


def permute_scatter_to_peers(col: torch.Tensor, first_global: int, total_rows: int, seed: int, rows_per_rank: int,
                             bases: Sequence[int], dst_pitch: int) -> None:
    """Rows of `col` ([n, ...] dense, or a [n, pitch] record buffer) = rows [first_global, first_global + n) of a dataset of
    total_rows rows: each is stored at its place in the shuffled dataset, on whichever rank that is."""
    dev = _require_cuda(col)
    n = col.shape[0]
    row_bytes = col.element_size()
    for d in col.shape[1:]:
        row_bytes *= int(d)
    w = len(bases)
    b = (ctypes.c_void_p * w)(*[int(x) for x in bases])
    with torch.cuda.device(dev):
        _lib.call("b200x_permute_scatter_to_peers", col.data_ptr(), row_bytes, row_bytes, n, first_global, total_rows,
                  seed & 0xFFFFFFFFFFFFFFFF, rows_per_rank, b, dst_pitch, w, _stream(dev))


def mask_values(values: torch.Tensor, valid: torch.Tensor, fill_mode: int) -> torch.Tensor:
    """values with null rows (valid == 0) replaced by the identity of a sum (0), min (1) or max (2) reduction."""
    dev = _require_cuda(values, valid)
    out = torch.empty_like(values)
    with torch.cuda.device(dev):
        _lib.call("b200x_mask_values", values.data_ptr(), value_type_of(values), valid.data_ptr(), values.numel(), fill_mode,
                  out.data_ptr(), None, _stream(dev))
    return out


def valid_as_int64(valid: torch.Tensor) -> torch.Tensor:
    """The validity bytes of a column as an int64 0 / 1 column (summed per group = number of non-null rows)."""
    dev = _require_cuda(valid)
    out = torch.empty(valid.numel(), dtype=torch.int64, device=dev)
    with torch.cuda.device(dev):
        _lib.call("b200x_mask_values", None, VAL_I64, valid.data_ptr(), valid.numel(), 0, None, out.data_ptr(), _stream(dev))
    return out


def group_validity(nvalid: torch.Tensor, rows: torch.Tensor, ignore_nulls: bool) -> torch.Tensor:
    """uint8 per group: does the aggregation have a result under the reference's null policy?"""
    dev = _require_cuda(nvalid, rows)
    out = torch.empty(nvalid.numel(), dtype=torch.uint8, device=dev)
    with torch.cuda.device(dev):
        _lib.call("b200x_group_validity", nvalid.data_ptr(), rows.data_ptr(), nvalid.numel(), int(ignore_nulls), out.data_ptr(),
                  _stream(dev))
    return out
