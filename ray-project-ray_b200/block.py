"""Device blocks: the columnar HBM block format and its BlockAccessor.

Mirrors the data-plane types of the reference (python/ray/data/block.py:54, :123-215,
:219-435) for a block whose columns live in HBM:

  Block            -> DeviceBlock: ordered {column name: torch CUDA tensor}, every column
                      `[n]` (any fixed-width dtype) or `[n, w]` (a packed fixed-width payload,
                      row-major so that a row moves as one contiguous w-byte record).
  BlockAccessor    -> DeviceBlockAccessor: the hot-path methods of ArrowBlockAccessor
                      (python/ray/data/_internal/arrow_block.py:159-680) implemented by calling
                      the C ABI; nothing is computed with torch ops or on the CPU.
  BlockMetadata / BlockExecStats -> same fields.

Record layout: several columns of a block may be views into ONE record buffer `[n, pitch]` (pitch a
multiple of 16 bytes; `DeviceBlock.to_records`): column c of row r lies at `r * pitch + offset_c`. The
columns keep their names, dtypes and shapes (they are strided torch views), but every row-moving operation
(sort, take, random_shuffle, the multi-GPU exchange) moves the record - one pass of aligned 16-byte accesses
per row whatever the column widths, instead of one pass per column (a gensort row = 10-byte key + 90-byte
payload in a 128-byte record moves as one DRAM line; as two dense columns it moves in 2-byte pieces).

A byte-string key column that lives inside a record may be accompanied by its **key prefix column**: the
key's first 8 bytes as a dense big-endian uint64 column (`DeviceBlock.aux["<col>::be0"]`), i.e. the key kept
columnar in directly radix-sortable form next to the row records (reading a 10-byte key out of 128-byte records
costs a pass over every record: the L2 fetches whole 128-byte lines). It is built at ingest
(`to_records(key_prefix=[...])`), comes out of every sort for free (the sort writes the sorted keys), and is simply
dropped by operations that do not maintain it - `key_column` packs it again on demand.

Null values: a nullable VALUE column is its values (0 where null) plus a hidden validity column `<col>::valid`
(uint8, 1 = present) that travels through every operation like any other column and never shows in the schema;
aggregations apply the reference's null policy to it (aggregate.py), egress turns it back into Arrow nulls / NaN.
Nulls in sort or groupby KEY columns and variable-width columns are outside the device format (SURVEY.md §8:
fixed-width int/float keys and payloads).
"""
from __future__ import annotations

import os
import time
from dataclasses import dataclass
from typing import Any, Dict, Iterator, List, Optional, Sequence, Tuple, Union

import numpy as np
import torch

from . import ops

try:  # resource is POSIX-only, as in the reference (block.py:21-24)
    import resource
except ImportError:  # pragma: no cover
    resource = None

KeyType = Any
AggType = Any


def default_device() -> torch.device:
    """The CUDA device new blocks are placed on: LOCAL_RANK's GPU under torchrun, else cuda:0."""
    if not torch.cuda.is_available():
        raise RuntimeError("the device block format needs a CUDA device; there is no CPU fallback")
    return torch.device("cuda", int(os.environ.get("LOCAL_RANK", "0")) % torch.cuda.device_count())


class BlockExecStats:
    """Execution stats of one map / merge / reduce step (reference: block.py:123-154)."""

    def __init__(self):
        self.start_time_s: Optional[float] = None
        self.end_time_s: Optional[float] = None
        self.wall_time_s: Optional[float] = None
        self.udf_time_s: Optional[float] = 0
        self.cpu_time_s: Optional[float] = None
        self.node_id = f"gpu{os.environ.get('LOCAL_RANK', '0')}"
        self.max_rss_bytes: int = 0
        self.task_idx: Optional[int] = None

    @staticmethod
    def builder() -> "_BlockExecStatsBuilder":
        return _BlockExecStatsBuilder()

    def __repr__(self):
        return repr({"wall_time_s": self.wall_time_s, "cpu_time_s": self.cpu_time_s,
                     "udf_time_s": self.udf_time_s, "node_id": self.node_id})


class _BlockExecStatsBuilder:
    def __init__(self):
        self.start_time = time.perf_counter()
        self.start_cpu = time.process_time()

    def build(self) -> BlockExecStats:
        stats = BlockExecStats()
        stats.start_time_s = self.start_time
        stats.end_time_s = time.perf_counter()
        stats.wall_time_s = stats.end_time_s - self.start_time
        stats.cpu_time_s = time.process_time() - self.start_cpu
        if resource is not None:
            stats.max_rss_bytes = int(resource.getrusage(resource.RUSAGE_SELF).ru_maxrss * 1e3)
        return stats


@dataclass
class BlockMetadata:
    """Reference: block.py:194-215."""

    num_rows: Optional[int]
    size_bytes: Optional[int]
    schema: Optional[Any]
    input_files: Optional[List[str]]
    exec_stats: Optional[BlockExecStats]

    def __post_init__(self):
        if self.input_files is None:
            self.input_files = []
        if self.size_bytes is not None:
            assert isinstance(self.size_bytes, int)


class Schema:
    """Column names and dtypes of a device block (stands in for pyarrow.Schema)."""

    def __init__(self, names: Sequence[str], types: Sequence[Tuple[torch.dtype, Tuple[int, ...]]]):
        self.names = list(names)
        self.types = list(types)

    def __eq__(self, other):
        return isinstance(other, Schema) and self.names == other.names and self.types == other.types

    def __repr__(self):
        cols = ", ".join(f"{n}: {str(t[0]).replace('torch.', '')}{list(t[1]) if t[1] else ''}"
                         for n, t in zip(self.names, self.types))
        return f"Schema({cols})"


_NP_TO_TORCH = {
    np.dtype(np.int64): torch.int64, np.dtype(np.float64): torch.float64, np.dtype(np.int32): torch.int32,
    np.dtype(np.float32): torch.float32, np.dtype(np.int16): torch.int16, np.dtype(np.int8): torch.int8,
    np.dtype(np.uint8): torch.uint8, np.dtype(np.bool_): torch.uint8,
}
_NP_UNSIGNED_VIEW = {np.dtype(np.uint64): (np.int64, "uint64"), np.dtype(np.uint32): (np.int32, "uint32"),
                     np.dtype(np.uint16): (np.int16, "uint16")}


def _from_numpy(col: np.ndarray) -> torch.Tensor:
    """torch view of a host array that is only read (Arrow buffers are read-only: no copy just to silence torch)."""
    if col.flags.writeable:
        return torch.from_numpy(col)
    import warnings

    with warnings.catch_warnings():
        warnings.simplefilter("ignore", UserWarning)
        return torch.from_numpy(col)


def _to_device(col: np.ndarray, device: torch.device) -> torch.Tensor:
    col = np.ascontiguousarray(col)
    if col.dtype in _NP_UNSIGNED_VIEW:
        signed, name = _NP_UNSIGNED_VIEW[col.dtype]
        return _from_numpy(col.view(signed)).to(device, non_blocking=True).view(getattr(torch, name))
    if col.dtype == np.bool_:
        col = col.view(np.uint8)
    if col.dtype not in _NP_TO_TORCH:
        raise TypeError(f"column dtype {col.dtype} is not a fixed-width type the device block format holds")
    return _from_numpy(col).to(device, non_blocking=True)


def _to_host(t: torch.Tensor) -> np.ndarray:
    if t.dtype in (getattr(torch, "uint64", None), getattr(torch, "uint32", None), getattr(torch, "uint16", None)):
        signed = {8: torch.int64, 4: torch.int32, 2: torch.int16}[t.element_size()]
        unsigned = {8: np.uint64, 4: np.uint32, 2: np.uint16}[t.element_size()]
        return t.view(signed).cpu().numpy().view(unsigned)
    return t.cpu().numpy()


# ------------------------------------------------------------------ record layout helpers
BYTE_KEY_SEP = "::be"   # `<column>::be<i>` = the i-th big-endian 8-byte chunk of a byte-string key column
VALID_SEP = "::valid"   # `<column>::valid` = hidden validity column (uint8, 1 = present) of a nullable value column


def valid_name(col: str) -> str:
    return f"{col}{VALID_SEP}"


def is_hidden(name: str) -> bool:
    return name.endswith(VALID_SEP)
RECORD_ALIGN = 16


def _inner(t: torch.Tensor) -> Tuple[int, bool]:
    """(bytes of one row, rows are dense inside) of a column tensor."""
    es, nbytes, expect, dense = t.element_size(), t.element_size(), 1, True
    for size, stride in zip(reversed(t.shape[1:]), reversed(t.stride()[1:])):
        if size != 1 and stride != expect:
            dense = False
        expect *= size
        nbytes *= size
    return nbytes, dense


def _row_pitch(t: torch.Tensor) -> Optional[int]:
    """Row pitch in bytes if `t` is a column whose rows are dense inside and lie further apart than their
    width (a record member or any other row-strided view), else None."""
    if t.dim() < 1 or t.shape[0] == 0:
        return None
    inner, dense = _inner(t)
    pitch = t.stride(0) * t.element_size()
    return pitch if dense and pitch > inner else None


def _contiguous_strides(shape: Sequence[int]) -> Tuple[int, ...]:
    out, run = [], 1
    for size in reversed(shape):
        out.append(run)
        run *= size
    return tuple(reversed(out))


def member_view(rec: torch.Tensor, offset: int, dtype: torch.dtype, inner_shape: Sequence[int]) -> torch.Tensor:
    """Column view `[m, *inner_shape]` of dtype `dtype` at byte `offset` of every record of `rec` ([m, pitch] uint8)."""
    m, pitch = rec.shape
    flat = rec.reshape(-1)
    es = torch.empty(0, dtype=dtype).element_size()
    typed = flat if dtype == torch.uint8 else flat.view(dtype)
    inner_shape = tuple(int(x) for x in inner_shape)
    return typed.as_strided((m,) + inner_shape, (pitch // es,) + _contiguous_strides(inner_shape),
                            rec.storage_offset() // es + offset // es)


@dataclass
class RecordGroup:
    """Columns of one block that are views into the same record buffer."""

    rec: torch.Tensor                                                   # [n, pitch] uint8 view of the records
    pitch: int
    members: Dict[str, Tuple[int, torch.dtype, Tuple[int, ...]]]        # name -> (byte offset, dtype, inner shape)

    def views(self, rec: torch.Tensor) -> Dict[str, torch.Tensor]:
        return {k: member_view(rec, off, dt, shape) for k, (off, dt, shape) in self.members.items()}


def record_groups(block: "DeviceBlock", names: Optional[Sequence[str]] = None) -> Tuple[List[RecordGroup], List[str]]:
    """Split the columns `names` of a block into record groups (moved record-wise) and loose columns.
    A group is used when its member columns fill at least half of the record (otherwise moving the dense
    columns one by one touches fewer bytes)."""
    names = list(block.columns) if names is None else list(names)
    found: Dict[Tuple[int, int, int], RecordGroup] = {}
    of: Dict[str, Tuple[int, int, int]] = {}
    n = block.num_rows
    for name in names:
        t = block.columns[name]
        pitch = _row_pitch(t) if t.is_cuda else None
        if pitch is None or pitch % RECORD_ALIGN or pitch > (1 << 20):
            continue
        inner, _ = _inner(t)
        storage = t.untyped_storage()
        off = t.data_ptr() - storage.data_ptr()
        col_off = off % pitch
        rec0 = off - col_off
        if col_off + inner > pitch or col_off % t.element_size() or (storage.data_ptr() + rec0) % RECORD_ALIGN \
                or rec0 + n * pitch > storage.nbytes():
            continue
        key = (storage.data_ptr(), pitch, rec0)
        if key not in found:
            rec = torch.empty(0, dtype=torch.uint8, device=t.device).set_(storage, rec0, (n, pitch), (pitch, 1))
            found[key] = RecordGroup(rec, pitch, {})
        found[key].members[name] = (col_off, t.dtype, tuple(t.shape[1:]))
        of[name] = key
    groups = []
    for key, g in found.items():
        filled = sum(_inner(block.columns[k])[0] for k in g.members)
        if 2 * filled >= g.pitch:
            groups.append(g)
        else:
            for k in g.members:
                del of[k]
    return groups, [k for k in names if k not in of]


def auto_pitch(used: int) -> int:
    """Record pitch for `used` bytes of columns: 16 / 32 / 64, else a multiple of 128 (whole DRAM lines)."""
    for p in (16, 32, 64):
        if used <= p:
            return p
    return (used + 127) // 128 * 128


class DeviceBlock:
    """A columnar block resident in HBM. Immutable by convention; slices are views."""

    __slots__ = ("columns", "_n", "device", "aux")

    def __init__(self, columns: Dict[str, torch.Tensor], num_rows: Optional[int] = None,
                 device: Optional[torch.device] = None, aux: Optional[Dict[str, torch.Tensor]] = None):
        self.columns: Dict[str, torch.Tensor] = dict(columns)
        # key prefix columns (module docstring): "<col>::be0" -> dense uint64 [n]; never part of the schema
        self.aux: Dict[str, torch.Tensor] = dict(aux) if aux else {}
        n = num_rows
        dev = device
        for name, t in self.columns.items():
            if not isinstance(t, torch.Tensor):
                raise TypeError(f"column {name!r} is not a tensor")
            if n is None:
                n = t.shape[0]
            elif t.shape[0] != n:
                raise ValueError(f"column {name!r} has {t.shape[0]} rows, expected {n}")
            if dev is None:
                dev = t.device
        self._n = int(n or 0)
        self.device = dev

    # -- construction -------------------------------------------------------------------
    @staticmethod
    def from_numpy(cols: Dict[str, np.ndarray], device: Optional[torch.device] = None) -> "DeviceBlock":
        device = device or default_device()
        return DeviceBlock({k: _to_device(np.asarray(v), device) for k, v in cols.items()}, device=device)

    @staticmethod
    def from_arrow(table, device: Optional[torch.device] = None) -> "DeviceBlock":
        import pyarrow as pa

        device = device or default_device()
        cols: Dict[str, torch.Tensor] = {}
        hidden: Dict[str, torch.Tensor] = {}
        for name in table.column_names:
            col = table[name]
            if isinstance(col, pa.ChunkedArray):
                col = col.combine_chunks()
            if col.null_count:
                if not (pa.types.is_integer(col.type) or pa.types.is_floating(col.type)):
                    raise ValueError(f"column {name!r} ({col.type}) has nulls; only numeric columns may be nullable in a device block")
                import pyarrow.compute as pac

                hidden[valid_name(name)] = _to_device(pac.is_valid(col).to_numpy(zero_copy_only=False).astype(np.uint8), device)
                col = pac.fill_null(col, 0)
            if pa.types.is_fixed_size_binary(col.type):
                w = col.type.byte_width
                buf = np.frombuffer(col.buffers()[1], dtype=np.uint8, count=len(col) * w, offset=col.offset * w)
                arr = buf.reshape(len(col), w)
            elif pa.types.is_null(col.type):
                raise ValueError(f"column {name!r} has null type")
            else:
                arr = col.to_numpy(zero_copy_only=False)
            cols[name] = _to_device(arr, device)
        cols.update(hidden)
        return DeviceBlock(cols, num_rows=table.num_rows, device=device)

    @staticmethod
    def from_pandas(df, device: Optional[torch.device] = None) -> "DeviceBlock":
        """pandas marks missing values with NaN / None and the reference turns them into Arrow nulls on the way in
        (pa.Table.from_pandas): numeric columns with missing values become nullable columns here too (NaN as a VALUE
        is what `from_numpy` ingests). Missing values in non-numeric columns are refused."""
        cols: Dict[str, np.ndarray] = {}
        hidden: Dict[str, np.ndarray] = {}
        for c in df.columns:
            ser = df[c]
            if ser.isna().any():
                import pandas as pd

                if not pd.api.types.is_numeric_dtype(ser.dtype):
                    raise ValueError(f"column {c!r} ({ser.dtype}) has missing values; only numeric columns may be nullable "
                                     f"in a device block")
                hidden[valid_name(str(c))] = (~ser.isna()).to_numpy().astype(np.uint8)
                ser = ser.fillna(0)
                if str(ser.dtype) in ("Int64", "Int32", "Int16", "Int8", "Float64", "Float32"):   # pandas' nullable dtypes
                    ser = ser.astype(str(ser.dtype).lower())
            cols[str(c)] = ser.to_numpy()
        cols.update(hidden)
        return DeviceBlock.from_numpy(cols, device)

    @staticmethod
    def empty(device: Optional[torch.device] = None) -> "DeviceBlock":
        return DeviceBlock({}, num_rows=0, device=device)

    def to_records(self, columns: Optional[Sequence[str]] = None, pitch: Optional[int] = None,
                   key_prefix: Optional[Sequence[str]] = None) -> "DeviceBlock":
        """The same block with `columns` (default: all) stored side by side in one record per row (module
        docstring, "Record layout"). Offsets follow the column order, each aligned to its element size; `pitch`
        defaults to `auto_pitch`. Padding bytes are zero. `key_prefix`: byte-string columns ([n, w] uint8) that get
        a key prefix column (module docstring)."""
        names = list(self.columns) if columns is None else list(columns)
        if not names or self._n == 0:
            return self
        layout, used = {}, 0
        for k in names:
            t = self.columns[k]
            es = t.element_size()
            used = (used + es - 1) // es * es
            inner, _ = _inner(t)
            layout[k] = (used, t.dtype, tuple(t.shape[1:]), inner)
            used += inner
        pitch = auto_pitch(used) if pitch is None else int(pitch)
        if pitch % RECORD_ALIGN or pitch < used:
            raise ValueError(f"record pitch {pitch} must be a multiple of {RECORD_ALIGN} and hold {used} bytes of columns")
        rec = torch.zeros((self._n, pitch), dtype=torch.uint8, device=self.device)
        cols = dict(self.columns)
        # Columns that already lie side by side in one row buffer (a [n, 100] tensor of gensort records cut into a key
        # and a payload view) and stay side by side in the record are copied as ONE run of bytes per row: wider
        # accesses (4 instead of 2 bytes for gensort) and one pass instead of one per column.
        runs: List[List[Any]] = []      # [source address of row 0, source pitch, record offset, bytes, tensors kept alive]
        for k, (off, dt, shape, inner) in layout.items():
            t = self.columns[k]
            src_pitch = _row_pitch(t)
            if src_pitch is None:
                t = t.contiguous()
                src_pitch = inner
            last = runs[-1] if runs else None
            if last is not None and last[1] == src_pitch and t.data_ptr() == last[0] + last[3] and off == last[2] + last[3] \
                    and last[3] + inner <= src_pitch:
                last[3] += inner
                last[4].append(t)
            else:
                runs.append([t.data_ptr(), src_pitch, off, inner, [t]])
            cols[k] = member_view(rec, off, dt, shape)
        for src_ptr, src_pitch, off, nbytes, _keep in runs:
            ops.copy_rows_2d(src_ptr, src_pitch, rec.data_ptr() + off, pitch, nbytes, self._n, self.device)
        aux = {}
        for k in key_prefix or []:
            aux[f"{k}{BYTE_KEY_SEP}0"] = ops.byte_key_chunks(cols[k])[0] if cols[k].shape[1] > 8 else \
                ops.pack_key_bytes(cols[k], 0, cols[k].shape[1])
        return DeviceBlock(cols, num_rows=self._n, device=self.device, aux=aux)

    def key_column(self, name: str) -> torch.Tensor:
        """The dense 1-D key column `name`. `<col>::be<i>` names the i-th big-endian 8-byte chunk of the
        byte-string column `col` (planner._expand_byte_keys): computed here when the block does not carry it."""
        if name in self.columns:
            return _contig(self.columns[name])
        if name in self.aux:
            return self.aux[name]
        base, sep, idx = name.rpartition(BYTE_KEY_SEP)
        if sep and base in self.columns and idx.isdigit():
            col = self.columns[base]
            return ops.byte_key_chunks(col if _row_pitch(col) is not None else _contig(col))[int(idx)]
        raise KeyError(name)

    def key_columns(self, names: Sequence[str]) -> List[torch.Tensor]:
        """`key_column` for several names at once: the chunks of one byte-string column are packed in a single pass."""
        out: Dict[str, torch.Tensor] = {}
        for name in names:
            if name in out:
                continue
            if name in self.aux:
                out[name] = self.aux[name]
                continue
            base, sep, idx = name.rpartition(BYTE_KEY_SEP)
            if name not in self.columns and sep and base in self.columns and idx.isdigit():
                col = self.columns[base]
                chunks = ops.byte_key_chunks(col if _row_pitch(col) is not None else _contig(col))
                for i, ch in enumerate(chunks):
                    out[f"{base}{BYTE_KEY_SEP}{i}"] = ch
            else:
                out[name] = self.key_column(name)
        return [out[name] for name in names]

    def rows_as_bytes(self, columns: Optional[Sequence[str]] = None) -> torch.Tensor:
        """The bytes of `columns` (default: all visible ones, in order) of every row as one dense `[n, width]` uint8
        tensor - the row-oriented form of the block (a gensort file is exactly this: 100-byte records). Columns that
        lie side by side in a record are copied as one run of bytes per row."""
        names = self.visible_names() if columns is None else list(columns)
        widths = [_inner(self.columns[k])[0] for k in names]
        out = torch.empty((self._n, sum(widths)), dtype=torch.uint8, device=self.device)
        runs: List[List[Any]] = []      # [source address, source pitch, output offset, bytes, tensors kept alive]
        off = 0
        for k, w in zip(names, widths):
            t = self.columns[k]
            src_pitch = _row_pitch(t)
            if src_pitch is None:
                t = t.contiguous()
                src_pitch = w
            last = runs[-1] if runs else None
            if last is not None and last[1] == src_pitch and t.data_ptr() == last[0] + last[3] and last[3] + w <= src_pitch:
                last[3] += w
                last[4].append(t)
            else:
                runs.append([t.data_ptr(), src_pitch, off, w, [t]])
            off += w
        for src_ptr, src_pitch, o, nbytes, _keep in runs:
            ops.copy_rows_2d(src_ptr, src_pitch, out.data_ptr() + o, out.shape[1], nbytes, self._n, self.device)
        return out

    def key_piece(self, name: str):
        """A later key column as `fix_ties` takes it: a chunk of a byte-string column is read where it lies
        (`ops.KeyPiece`), anything else is the dense column."""
        base, sep, idx = name.rpartition(BYTE_KEY_SEP)
        if name not in self.columns and name not in self.aux and sep and base in self.columns and idx.isdigit():
            col = self.columns[base]
            off = 8 * int(idx)
            return ops.KeyPiece(col, off, min(8, col.shape[1] - off))
        return self.key_column(name)

    # -- introspection ------------------------------------------------------------------
    @property
    def num_rows(self) -> int:
        return self._n

    @property
    def column_names(self) -> List[str]:
        return list(self.columns)

    def __len__(self) -> int:
        return self._n

    def __getitem__(self, name: str) -> torch.Tensor:
        return self.columns[name]

    def __repr__(self):
        return f"DeviceBlock(rows={self._n}, {self.schema()})"

    def visible_names(self) -> List[str]:
        """The columns a user sees: everything but the hidden validity columns."""
        return [k for k in self.columns if not is_hidden(k)]

    def validity(self, col: str) -> Optional[torch.Tensor]:
        """uint8 [n] (1 = present) if `col` is nullable, else None."""
        return self.columns.get(valid_name(col))

    def schema(self) -> Optional[Schema]:
        if not self.columns:
            return None
        names = self.visible_names()
        return Schema(names, [(self.columns[k].dtype, tuple(self.columns[k].shape[1:])) for k in names])

    def size_bytes(self) -> int:
        return int(sum(t.numel() * t.element_size() for t in self.columns.values()))

    # -- zero-copy structure ops --------------------------------------------------------
    def slice(self, start: int, end: int) -> "DeviceBlock":
        start = max(0, min(start, self._n))
        end = max(start, min(end, self._n))
        return DeviceBlock({k: v[start:end] for k, v in self.columns.items()}, num_rows=end - start,
                           device=self.device, aux={k: v[start:end] for k, v in self.aux.items()})

    def select(self, names: Sequence[str]) -> "DeviceBlock":
        names = list(names) + [valid_name(k) for k in names if valid_name(k) in self.columns and valid_name(k) not in names]
        return DeviceBlock({k: self.columns[k] for k in names}, num_rows=self._n, device=self.device,
                           aux={k: v for k, v in self.aux.items() if k.rpartition(BYTE_KEY_SEP)[0] in names})

    @staticmethod
    def concat(blocks: Sequence["DeviceBlock"]) -> "DeviceBlock":
        """Concatenate same-schema blocks in order (empty / schema-less blocks are dropped).

        Adjacent views of one buffer (the normal case for blocks cut from one dataset buffer)
        are re-joined without a copy."""
        blocks = [b for b in blocks if b.num_rows > 0]
        if not blocks:
            return DeviceBlock.empty()
        if len(blocks) == 1:
            return blocks[0]
        names = blocks[0].column_names
        for b in blocks[1:]:    # a column may be nullable in some blocks only: the others are all-present
            names += [k for k in b.column_names if is_hidden(k) and k not in names]
        out: Dict[str, torch.Tensor] = {}
        for name in names:
            parts = [b.columns[name] if name in b.columns else torch.ones(b.num_rows, dtype=torch.uint8, device=b.device)
                     for b in blocks]
            out[name] = _join_views(parts)
        aux = {k: _join_views([b.aux[k] for b in blocks]) for k in blocks[0].aux if all(k in b.aux for b in blocks)}
        return DeviceBlock(out, device=blocks[0].device, aux=aux)

    # -- egress (D2H happens only here) -------------------------------------------------
    def to_numpy(self, columns: Optional[Sequence[str]] = None) -> Dict[str, np.ndarray]:
        """Host copies of the (visible) columns. Null entries of a nullable column come back as NaN - the column as
        float64 if it holds integers, which is what pyarrow's to_numpy gives a nullable integer column."""
        names = list(columns) if columns is not None else self.visible_names()
        out = {}
        for k in names:
            v = _to_host(_contig(self.columns[k]))
            m = self.columns.get(valid_name(k))
            if m is not None and v.ndim == 1:
                ok = _to_host(_contig(m)).astype(bool)
                if not ok.all():
                    v = v.astype(np.float64) if not np.issubdtype(v.dtype, np.floating) else v.copy()
                    v[~ok] = np.nan
            out[k] = v
        return out

    def to_arrow(self):
        import pyarrow as pa

        arrays, names = [], []
        for k in self.visible_names():
            v = _to_host(_contig(self.columns[k]))
            m = self.columns.get(valid_name(k))
            if v.ndim == 2:
                arrays.append(pa.FixedSizeBinaryArray.from_buffers(pa.binary(v.shape[1] * v.dtype.itemsize), len(v),
                                                                   [None, pa.py_buffer(np.ascontiguousarray(v))]))
            elif m is not None:
                arrays.append(pa.array(v, mask=~_to_host(_contig(m)).astype(bool)))
            else:
                arrays.append(pa.array(v))
            names.append(k)
        return pa.table(arrays, names=names) if names else pa.table({})

    def to_pandas(self):
        import pandas as pd

        return pd.DataFrame({k: (list(v) if v.ndim == 2 else v) for k, v in self.to_numpy().items()})


def _join_views(parts: List[torch.Tensor]) -> torch.Tensor:
    first = parts[0]
    row = first.stride(0) * first.element_size() if first.dim() >= 1 and first.shape[0] > 0 else 0
    # dense columns, or columns at a common row pitch (record members): adjacent views re-join without a copy
    contiguous = all(p.dtype == first.dtype and p.shape[1:] == first.shape[1:] and p.stride() == first.stride()
                     and _inner(p)[1] for p in parts)
    if contiguous and row:
        ptr = first.data_ptr()
        same = True
        for p in parts:
            if p.data_ptr() != ptr or p.untyped_storage().data_ptr() != first.untyped_storage().data_ptr():
                same = False
                break
            ptr += p.shape[0] * row
        if same:
            total = sum(p.shape[0] for p in parts)
            return first.as_strided((total,) + tuple(first.shape[1:]), first.stride(), first.storage_offset())
    return torch.cat([p for p in parts], dim=0)


Block = DeviceBlock


def _as_device_block(block: Any) -> DeviceBlock:
    if isinstance(block, DeviceBlock):
        return block
    try:
        import pyarrow as pa

        if isinstance(block, pa.Table):
            return DeviceBlock.from_arrow(block)
    except ImportError:  # pragma: no cover
        pass
    try:
        import pandas as pd

        if isinstance(block, pd.DataFrame):
            return DeviceBlock.from_pandas(block)
    except ImportError:  # pragma: no cover
        pass
    if isinstance(block, dict):
        return DeviceBlock.from_numpy(block)
    raise TypeError(f"Not a block type: {block!r} ({type(block)})")


class BlockAccessor:
    """`BlockAccessor.for_block` (reference: block.py:380-407). Arrow tables, DataFrames and
    dicts of numpy arrays are moved to the device first; everything runs on DeviceBlockAccessor."""

    @staticmethod
    def for_block(block: Any) -> "DeviceBlockAccessor":
        return DeviceBlockAccessor(_as_device_block(block))


# multi-column sorts try "is the first column alone enough?" from this many rows on, on a sample of this size
_DISTINCT_FIRST_MIN_ROWS = 1 << 16
_DISTINCT_SAMPLE = 1 << 14


class _TieVerdict:
    """Outcome of `ops.fix_ties`, read back without stalling the device: the two counters are copied to pinned host
    memory behind the kernel; `ok()` waits for that copy only."""

    def __init__(self, stats: torch.Tensor):
        self._host = torch.empty(2, dtype=torch.int64).pin_memory()
        self._host.copy_(stats, non_blocking=True)
        self._event = torch.cuda.Event()
        self._event.record(torch.cuda.current_stream(stats.device))
        self._keep = stats

    def ok(self) -> bool:
        self._event.synchronize()
        return int(self._host[1]) <= ops.FIX_TIES_MAX_RUN


class DeviceBlockAccessor(BlockAccessor):
    def __init__(self, block: DeviceBlock):
        self._block = block

    # -- plain accessors (block.py:227-378) ---------------------------------------------
    def num_rows(self) -> int:
        return self._block.num_rows

    def size_bytes(self) -> int:
        return self._block.size_bytes()

    def schema(self):
        return self._block.schema()

    def column_names(self) -> List[str]:
        return self._block.visible_names()

    def to_block(self) -> DeviceBlock:
        return self._block

    def to_arrow(self):
        return self._block.to_arrow()

    def to_pandas(self):
        return self._block.to_pandas()

    def to_numpy(self, columns: Optional[Union[str, List[str]]] = None):
        if isinstance(columns, str):
            return self._block.to_numpy([columns])[columns]
        return self._block.to_numpy(columns)

    def iter_rows(self, public_row_format: bool = True) -> Iterator[Dict[str, Any]]:
        host = self._block.to_numpy()
        nullable = {k for k in host if self._block.validity(k) is not None}

        def item(k, v, i):
            if v.ndim != 1:
                return v[i]
            x = v[i].item()
            return None if (k in nullable and x != x) else x      # a null comes back from to_numpy as NaN

        for i in range(self._block.num_rows):
            yield {k: item(k, v, i) for k, v in host.items()}

    def slice(self, start: int, end: int, copy: bool = False) -> DeviceBlock:
        out = self._block.slice(start, end)
        if copy:
            out = DeviceBlock({k: v.clone() for k, v in out.columns.items()}, num_rows=out.num_rows, device=out.device)
        return out

    def select(self, columns: List[str]) -> DeviceBlock:
        if not all(isinstance(c, str) for c in columns):
            raise ValueError(f"Columns must be a list of column name strings, but got: {columns}.")
        return self._block.select(columns)

    def get_metadata(self, input_files: Optional[List[str]] = None,
                     exec_stats: Optional[BlockExecStats] = None) -> BlockMetadata:
        return BlockMetadata(num_rows=self.num_rows(), size_bytes=self.size_bytes(), schema=self.schema(),
                             input_files=input_files, exec_stats=exec_stats)

    # -- gathers ------------------------------------------------------------------------
    def _take_columns(self, indices: torch.Tensor, names: Sequence[str]) -> Dict[str, torch.Tensor]:
        """Columns `names` at rows `indices`: record groups move as whole records (one pass per group), the
        other columns one by one (8-byte columns share a launch)."""
        b = self._block
        groups, loose = record_groups(b, names)
        cols: Dict[str, torch.Tensor] = {}
        for g in groups:
            cols.update(g.views(ops.gather_records(g.rec, indices)))
        outs = ops.gather_columns([_contig(b.columns[k]) for k in loose], indices)
        cols.update(zip(loose, outs))
        return cols

    def take(self, indices: torch.Tensor) -> DeviceBlock:
        """Rows at `indices` (int32 tensor holding uint32 row numbers). arrow_block.py:346-355."""
        b = self._block
        cols = self._take_columns(indices, b.column_names)
        return DeviceBlock({k: cols[k] for k in b.column_names}, num_rows=indices.numel(), device=b.device)

    def random_shuffle(self, random_seed: Optional[int]) -> DeviceBlock:
        """Seeded permutation of the rows. arrow_block.py:249-255 (permutation + take, fused here)."""
        b = self._block
        seed = 0 if random_seed is None else int(random_seed)
        if random_seed is None:
            seed = int.from_bytes(os.urandom(8), "little")
        groups, loose = record_groups(b)
        cols: Dict[str, torch.Tensor] = {}
        for g in groups:
            cols.update(g.views(ops.permute_records(g.rec, seed)))
        for k in loose:
            cols[k] = ops.permute_rows(_contig(b.columns[k]), seed)
        return DeviceBlock({k: cols[k] for k in b.column_names}, num_rows=b.num_rows, device=b.device)

    def sample(self, n_samples: int, sort_key: "SortKey") -> DeviceBlock:
        """`n_samples` rows of the key columns. table_block.py:252-262, arrow_block.py:365-368.

        The reference draws with the unseeded `random.sample`; here rows are taken at evenly spaced
        positions, which is deterministic (splitters are load-balance hints, not part of parity)."""
        b = self._block
        cols = sort_key.get_columns()
        if b.num_rows == 0 or not b.columns:
            return DeviceBlock.empty(b.device)
        n_samples = min(n_samples, b.num_rows)
        pos = ((np.arange(n_samples, dtype=np.float64) + 0.5) * (b.num_rows / max(n_samples, 1))).astype(np.int64)
        idx = torch.from_numpy(pos.astype(np.int32)).to(b.device)
        return DeviceBlock({c: ops.gather(b.key_column(c), idx) for c in cols}, num_rows=n_samples, device=b.device)

    # -- sort (arrow_block.py:449-464, :554-567) ------------------------------------------
    def _argsort(self, sort_key: "SortKey", want_codes: bool):
        """Stable multi-column argsort -> (perm, codes, sorted first key column)."""
        perm, codes, skeys, verdict = self._argsort_deferred(sort_key, want_codes)
        if verdict is not None and not verdict.ok():
            perm, codes, skeys = self._argsort_chain(sort_key, want_codes)
        return perm, codes, skeys

    def _argsort_chain(self, sort_key: "SortKey", want_codes: bool):
        """One stable sort per key column, from the last to the first."""
        b = self._block
        cols, desc = sort_key.get_columns(), sort_key.get_descending()
        perm = codes = skeys = None
        for i, col in enumerate(reversed(b.key_columns(cols))):
            lastcall = i == len(cols) - 1
            perm, codes, skeys = ops.argsort(col, desc, perm_in=perm, want_codes=want_codes and lastcall, want_keys=lastcall)
        return perm, codes, skeys

    def _argsort_deferred(self, sort_key: "SortKey", want_codes: bool):
        """(perm, codes, sorted first key column, verdict): `verdict` is None, or a `_TieVerdict` whose `ok()` must be
        asked before the result is trusted (False -> call `_argsort_chain`). Work that only depends on `perm` can be
        enqueued first: asking blocks the host until the sort has run, not the device."""
        b = self._block
        cols, desc = sort_key.get_columns(), sort_key.get_descending()
        nullable = [c for c in cols if valid_name(c) in b.columns]
        if nullable:
            if len(cols) > 1:
                raise NotImplementedError(f"multi-column sort keys with a nullable column ({nullable[0]!r}) are not supported "
                                          f"on device blocks")
            # Arrow puts nulls last in BOTH directions, after NaN, in input order (transform_pyarrow.py:17-21):
            # a stable sort by the values, then a stable sort by "is present" descending on top of it
            perm, _, _ = ops.argsort(b.key_column(cols[0]), desc)
            perm, _, _ = ops.argsort(_contig(b.columns[valid_name(cols[0])]), True, perm_in=perm)
            return perm, None, None, None
        if 1 < len(cols) <= 1 + ops.FIX_TIES_MAX_COLS and b.num_rows >= _DISTINCT_FIRST_MIN_ROWS:
            first = self._argsort_first_column_then_ties(cols, desc, want_codes)
            if first is not None:
                return first
        return self._argsort_chain(sort_key, want_codes) + (None,)

    def _argsort_first_column_then_ties(self, cols: List[str], desc: bool, want_codes: bool):
        """A multi-column sort whose first column has (almost) no duplicates is the sort by that column alone,
        plus a repair of the few runs of equal first keys (gensort's 10-byte key: 8 random bytes already separate
        nearly all records; timestamp + id; ...). If a sample of the first column shows no duplicate: one sort by
        it, then `fix_ties` orders every run of equal first keys by the remaining columns in place (byte-string
        chunks are compared where they lie in the rows). A run longer than the repair handles (the column was
        duplicate-heavy after all) is reported through the verdict. None = not applicable: run the chain."""
        b = self._block
        n, first = b.num_rows, cols[0]
        m = min(_DISTINCT_SAMPLE, n)
        pos = torch.from_numpy(((np.arange(m, dtype=np.float64) + 0.5) * (n / m)).astype(np.int64).astype(np.int32)).to(b.device)
        base, sep, idx = first.rpartition(BYTE_KEY_SEP)
        col = None
        if first not in b.columns and first not in b.aux and sep and base in b.columns and idx.isdigit() \
                and (_row_pitch(b.columns[base]) or 1) % RECORD_ALIGN == 0:
            # the sample is read straight from the records, before the pass that packs the whole column
            kcol = b.columns[base]
            off = 8 * int(idx)
            if kcol.shape[1] - off < 4:
                return None
            sample_col = ops.pack_key_bytes(ops.gather_pitched_column(kcol, pos), off, min(8, kcol.shape[1] - off))
        else:
            col = b.key_column(first)
            if col.dim() != 1 or col.element_size() < 4:
                return None
            sample_col = ops.gather(col, pos)
        sample = np.sort(ops.encode_keys(sample_col, False).cpu().numpy())   # (np.unique takes 20x as long)
        if bool((sample[1:] == sample[:-1]).any()):
            return None
        if col is None:
            col = b.key_column(first)
        perm, codes, skeys = ops.argsort(col, desc, want_codes=want_codes or col.dtype.is_floating_point, want_keys=True)
        probe = skeys if skeys is not None else codes.view(torch.uint64)
        stats = ops.fix_ties(probe, perm, [b.key_piece(c) for c in cols[1:]], desc)
        return perm, (codes if want_codes else None), skeys, _TieVerdict(stats)

    def sort(self, sort_key: "SortKey", want_codes: bool = False, on_sorted_key=None):
        """Whole block in sorted order (transform_pyarrow.sort: sort_indices + take).

        `on_sorted_key(column)`, if given, is called with the sorted first key column as soon as it exists
        and before the other columns are gathered: work that only needs the keys (cutting the result into
        blocks, including its device -> host read) then overlaps the payload gather."""
        b = self._block
        perm, codes, skeys, verdict = self._argsort_deferred(sort_key, want_codes)
        first_key = sort_key.get_columns()[0]
        for attempt in (0, 1):
            if skeys is None and on_sorted_key is not None:  # float keys are gathered, not decoded by the sort
                skeys = ops.gather(b.key_column(first_key), perm)
            if on_sorted_key is not None:
                on_sorted_key(skeys)
            # the sorted first key column comes out of the sort itself, unless it travels inside a record anyway
            decoded = skeys is not None and first_key in b.columns and _row_pitch(b.columns[first_key]) is None
            names = [k for k in b.column_names if not (decoded and k == first_key)]
            cols = self._take_columns(perm, names)
            if verdict is None or verdict.ok():
                break
            perm, codes, skeys = self._argsort_chain(sort_key, want_codes)   # rare: long runs of equal first keys
            verdict = None
        if decoded:
            cols[first_key] = skeys
        aux = {}
        if skeys is not None and first_key not in b.columns and first_key.endswith(f"{BYTE_KEY_SEP}0"):
            aux[first_key] = skeys       # the key prefix column of the sorted block, for free
        out = DeviceBlock({k: cols[k] for k in b.column_names}, num_rows=b.num_rows, device=b.device, aux=aux)
        return out, codes

    def sort_and_partition(self, boundaries: List[Any], sort_key: "SortKey") -> List[DeviceBlock]:
        b = self._block
        if b.num_rows == 0:
            return [DeviceBlock.empty(b.device) for _ in range(len(boundaries) + 1)]
        table, _ = self.sort(sort_key)
        if len(boundaries) == 0:
            return [table]
        if any(b.validity(c) is not None for c in sort_key.get_columns()):
            raise NotImplementedError("range partitioning by a nullable sort key is not supported on device blocks "
                                      "(the default fused scheduler sorts such a dataset in one pass)")
        return find_partitions(table, boundaries, sort_key)

    @staticmethod
    def merge_sorted_blocks(blocks: List[DeviceBlock], sort_key: "SortKey") -> Tuple[DeviceBlock, BlockMetadata]:
        stats = BlockExecStats.builder()
        blocks = [_as_device_block(b) for b in blocks]
        blocks = [b for b in blocks if b.num_rows > 0]
        if not blocks:
            ret = DeviceBlock.empty()
        else:
            ret, _ = DeviceBlockAccessor(DeviceBlock.concat(blocks)).sort(sort_key)
        return ret, DeviceBlockAccessor(ret).get_metadata(None, exec_stats=stats.build())

    # -- per-column reductions over the whole block (arrow_block.py:370-447) --------------
    def count(self, on: str) -> Optional[int]:
        if not isinstance(on, str):
            raise ValueError(f"on must be a string when aggregating on device blocks, but got:{type(on)}.")
        return None if self.num_rows() == 0 else self.num_rows()

    def _whole_block_reduce(self, on: str, want: str):
        from .aggregate import _global_reduce

        if not isinstance(on, str):
            raise ValueError(f"on must be a string when aggregating on device blocks, but got:{type(on)}.")
        if self.num_rows() == 0:
            return None
        return _global_reduce(self._block.columns[on], want)

    def sum(self, on: str, ignore_nulls: bool = True):
        return self._whole_block_reduce(on, "sum")

    def min(self, on: str, ignore_nulls: bool = True):
        return self._whole_block_reduce(on, "min")

    def max(self, on: str, ignore_nulls: bool = True):
        return self._whole_block_reduce(on, "max")

    def mean(self, on: str, ignore_nulls: bool = True):
        s = self._whole_block_reduce(on, "sum")
        return None if s is None else s / self.num_rows()

    def sum_of_squared_diffs_from_mean(self, on: str, ignore_nulls: bool = True, mean: Optional[float] = None):
        from .aggregate import _global_reduce

        if self.num_rows() == 0:
            return None
        if mean is None:
            mean = self.mean(on, ignore_nulls)
        return _global_reduce(self._block.columns[on], "m2", center=float(mean))

    # -- groupby (arrow_block.py:466-680) -------------------------------------------------
    def combine(self, key: Union[str, List[str], None], aggs: Sequence["AggregateFn"]) -> DeviceBlock:
        from .aggregate import combine_block

        if key is not None and not isinstance(key, (str, list)):
            raise ValueError(f"key must be a string, list of strings or None when aggregating on device blocks, "
                             f"but got: {type(key)}.")
        return combine_block(self._block, key, aggs, presorted=True)

    @staticmethod
    def aggregate_combined_blocks(blocks: List[DeviceBlock], key: Union[str, List[str], None],
                                  aggs: Sequence["AggregateFn"], finalize: bool) -> Tuple[DeviceBlock, BlockMetadata]:
        from .aggregate import merge_combined_blocks

        stats = BlockExecStats.builder()
        ret = merge_combined_blocks([_as_device_block(b) for b in blocks], key, aggs, finalize)
        return ret, DeviceBlockAccessor(ret).get_metadata(None, exec_stats=stats.build())

    @staticmethod
    def builder():
        from .builder import DeviceBlockBuilder

        return DeviceBlockBuilder()


def _contig(t: torch.Tensor) -> torch.Tensor:
    """Dense copy of a column unless it is dense already (record members are copied out by the 2-D copy kernel)."""
    if t.is_contiguous():
        return t
    pitch = _row_pitch(t) if t.is_cuda else None
    if pitch is None:
        return t.contiguous()
    out = torch.empty(t.shape, dtype=t.dtype, device=t.device)
    ops.copy_rows_2d(t.data_ptr(), pitch, out.data_ptr(), _inner(t)[0], _inner(t)[0], t.shape[0], t.device)
    return out


def boundary_codes(boundaries: Sequence[Any], key_col: torch.Tensor, descending: bool) -> torch.Tensor:
    """Encode scalar boundary values with the key column's own code (util.py:719-758 compares
    boundary values with column values; an int column against a float boundary compares as
    `x < b`  <=>  `x < ceil(b)`)."""
    dt = key_col.dtype
    vals = []
    for b in boundaries:
        if isinstance(b, (tuple, list)):
            b = b[0]
        if hasattr(b, "item"):
            b = b.item()
        vals.append(b)
    if dt.is_floating_point:
        host = np.asarray(vals, dtype=np.float64 if dt == torch.float64 else np.float32)
    else:
        info = np.iinfo(_to_host(torch.empty(0, dtype=dt)).dtype)
        ints = [int(min(max(int(np.ceil(v)), info.min), info.max)) if isinstance(v, float) else
                int(min(max(int(v), info.min), info.max)) for v in vals]
        host = np.asarray(ints, dtype=_to_host(torch.empty(0, dtype=dt)).dtype)
    dev_vals = _to_device(host, key_col.device)
    return ops.encode_keys(dev_vals, descending)


def _boundary_above_range(b: Any, dt: torch.dtype) -> bool:
    """True for a boundary above every value an integer column of dtype `dt` can hold (no row is >= it)."""
    if dt.is_floating_point:
        return False
    if isinstance(b, (tuple, list)):
        b = b[0]
    if hasattr(b, "item"):
        b = b.item()
    if isinstance(b, float) and np.isnan(b):
        raise ValueError("a NaN boundary cannot cut an integer key column")
    info = np.iinfo(_to_host(torch.empty(0, dtype=dt)).dtype)
    return (int(np.ceil(b)) if isinstance(b, float) else int(b)) > info.max


def find_partition_indices(sorted_keys: torch.Tensor, boundaries: Sequence[Any], descending: bool) -> List[int]:
    """find_partition_index for every boundary (util.py:719-758) on the SORTED first key column.
    Ascending: #rows < b (lower bound). Descending: #rows >= b (upper bound of the complemented code).
    A boundary above the column's integer range is not clamped to a representable key (rows equal to the dtype's
    maximum would then fall on the wrong side): ascending every row is < b, descending no row is >= b."""
    if len(boundaries) == 0:
        return []
    bc = boundary_codes(boundaries, sorted_keys, descending)
    idx = ops.search_sorted_keys(_contig(sorted_keys), descending, bc, upper=descending)
    n = sorted_keys.shape[0]
    return [(0 if descending else n) if _boundary_above_range(b, sorted_keys.dtype) else int(x)
            for b, x in zip(boundaries, idx.cpu().tolist())]


def find_partitions(table: DeviceBlock, boundaries: Sequence[Any], sort_key: "SortKey") -> List[DeviceBlock]:
    """Cut a sorted block at the boundaries (util.py:761-779). Single key column boundaries;
    multi-column boundaries partition on the first key column's value only (a coarser, still
    order-respecting cut: every row of partition i sorts before every row of partition i+1)."""
    first = sort_key.get_columns()[0]
    bounds = find_partition_indices(table.key_column(first), boundaries, sort_key.get_descending())
    parts, last = [], 0
    for idx in bounds:
        idx = max(idx, last)
        parts.append(table.slice(last, idx))
        last = idx
    parts.append(table.slice(last, table.num_rows))
    return parts
