"""GPU tests of the record layout (several columns of a row side by side in one record at a 16-byte
aligned pitch, block.py "Record layout") and of the tie repair of multi-column sorts. Everything is
compared bit for bit with the CPU oracle on the same inputs; the kernels are called through the C ABI."""

import numpy as np
import pytest
import torch

import ray_project_ray_b200.data as rd
from oracle import exchange_oracle as orc
from ray_project_ray_b200 import _lib, ops
from ray_project_ray_b200.block import (BlockAccessor, DeviceBlock, _row_pitch, auto_pitch, record_groups)
from ray_project_ray_b200.context import DataContext
from ray_project_ray_b200.sort_task_spec import SortKey

pytestmark = pytest.mark.gpu


@pytest.fixture(autouse=True)
def _restore_context():
    ctx = DataContext.get_current()
    saved = (ctx.use_push_based_shuffle, ctx.use_fused_device_exchange)
    yield
    ctx.use_push_based_shuffle, ctx.use_fused_device_exchange = saved


def _gensort(n, seed, dup_hi=0, dup_full=0):
    rng = np.random.default_rng(seed)
    key = rng.integers(0, 256, (n, 10)).astype(np.uint8)
    if dup_hi:      # rows that share their first 8 key bytes with another row: the last two bytes decide
        src = rng.integers(0, n, dup_hi)
        dst = rng.integers(0, n, dup_hi)
        key[dst, :8] = key[src, :8]
    if dup_full:    # exact duplicates: input order decides
        src = rng.integers(0, n, dup_full)
        dst = rng.integers(0, n, dup_full)
        key[dst] = key[src]
    payload = rng.integers(0, 256, (n, 90)).astype(np.uint8)
    payload[:, :8] = np.arange(n, dtype=np.int64).view(np.uint8).reshape(n, 8)
    return key, payload


def test_to_records_layout_and_round_trip(cuda_device):
    key, payload = _gensort(5000, 1)
    v = np.arange(5000, dtype=np.int64) * 3
    b = DeviceBlock.from_numpy({"key": key, "v": v, "payload": payload}, cuda_device)
    r = b.to_records()
    assert auto_pitch(10 + 6 + 8 + 90) == 128
    groups, loose = record_groups(r)
    assert len(groups) == 1 and not loose and groups[0].pitch == 128
    assert groups[0].members["key"][0] == 0 and groups[0].members["v"][0] == 16 and groups[0].members["payload"][0] == 24
    assert all(_row_pitch(r.columns[k]) == 128 for k in r.columns)
    assert r.schema() == b.schema()
    got = r.to_numpy()
    np.testing.assert_array_equal(got["key"], key)
    np.testing.assert_array_equal(got["v"], v)
    np.testing.assert_array_equal(got["payload"], payload)
    # padding is zero, slices stay record views and re-join without a copy
    rec = groups[0].rec.cpu().numpy()
    assert not rec[:, 10:16].any() and not rec[:, 114:].any()
    parts = [r.slice(0, 1234), r.slice(1234, 1234), r.slice(1234, 5000)]
    joined = DeviceBlock.concat(parts)
    assert joined.columns["payload"].data_ptr() == r.columns["payload"].data_ptr()
    assert len(record_groups(joined)[0]) == 1
    np.testing.assert_array_equal(joined.to_numpy()["payload"], payload)
    # a chosen pitch, a subset of the columns
    r2 = b.to_records(["key", "payload"], pitch=112)
    g2, loose2 = record_groups(r2)
    assert g2[0].pitch == 112 and loose2 == ["v"]
    np.testing.assert_array_equal(r2.to_numpy()["payload"], payload)
    with pytest.raises(ValueError):
        b.to_records(pitch=100)


@pytest.mark.parametrize("desc", [False, True])
@pytest.mark.parametrize("n,dup_hi,dup_full", [(200_000, 0, 0), (200_000, 3000, 500), (70_000, 70_000, 5000), (3000, 10, 10)])
def test_sort_gensort_records(n, dup_hi, dup_full, desc, cuda_device):
    """Dataset.sort by the 10-byte key of gensort records kept as 128-byte records: pyarrow's order incl. ties."""
    key, payload = _gensort(n, 7 + n, dup_hi, dup_full)
    ds = rd.from_numpy_columns({"key": key, "payload": payload}, override_num_blocks=4, records=True)
    assert all(_row_pitch(b.columns["payload"]) == 128 for b in ds.get_internal_block_refs())
    out = ds.sort("key", descending=desc)
    assert out.columns() == ["key", "payload"] and out.num_blocks() == 4
    want = orc.stable_argsort_bytes(key, desc)
    got = out.to_numpy()
    np.testing.assert_array_equal(got["key"], key[want])
    np.testing.assert_array_equal(got["payload"], payload[want])
    # the output blocks are records again
    assert all(len(record_groups(b)[0]) == 1 for b in out.get_internal_block_refs() if b.num_rows)


@pytest.mark.parametrize("sched", ["pull", "push"])
def test_sort_records_task_schedulers(sched, cuda_device):
    ctx = DataContext.get_current()
    ctx.use_fused_device_exchange = False
    ctx.use_push_based_shuffle = sched == "push"
    key, payload = _gensort(30_000, 3, 2000, 100)
    ds = rd.from_numpy_columns({"key": key, "payload": payload}, override_num_blocks=5, records=True)
    got = ds.sort("key").to_numpy()
    want = orc.stable_argsort_bytes(key, False)
    np.testing.assert_array_equal(got["key"], key[want])
    np.testing.assert_array_equal(got["payload"], payload[want])


def test_sort_int_key_inside_record(cuda_device):
    rng = np.random.default_rng(5)
    n = 150_000
    ident = rng.integers(-1000, 1000, n, dtype=np.int64)
    x = rng.random(n)
    payload = rng.integers(0, 256, (n, 64)).astype(np.uint8)
    ds = rd.from_numpy_columns({"id": ident, "x": x, "payload": payload}, override_num_blocks=3, records=True)
    for desc in (False, True):
        got = ds.sort("id", descending=desc).to_numpy()
        order = orc.stable_argsort(ident, desc)
        np.testing.assert_array_equal(got["id"], ident[order])
        np.testing.assert_array_equal(got["x"].view(np.int64), x[order].view(np.int64))
        np.testing.assert_array_equal(got["payload"], payload[order])
    # shuffle and groupby see the same columns
    sh = ds.random_shuffle(seed=11).to_numpy()
    p = orc.feistel_permutation(n, 11)
    np.testing.assert_array_equal(sh["id"], ident[p])
    np.testing.assert_array_equal(sh["payload"], payload[p])
    agg = ds.groupby("id").count().to_numpy()
    ids, cnt = np.unique(ident, return_counts=True)
    np.testing.assert_array_equal(agg["id"], ids)
    np.testing.assert_array_equal(agg["count()"], cnt)


def test_take_and_select_on_records(cuda_device):
    key, payload = _gensort(10_000, 9)
    b = DeviceBlock.from_numpy({"key": key, "payload": payload}, cuda_device).to_records()
    idx = np.random.default_rng(0).integers(0, 10_000, 777).astype(np.int32)
    t = BlockAccessor.for_block(b).take(torch.from_numpy(idx).to(cuda_device)).to_numpy()
    np.testing.assert_array_equal(t["key"], key[idx])
    np.testing.assert_array_equal(t["payload"], payload[idx])
    only_key = BlockAccessor.for_block(b.select(["key"])).take(torch.from_numpy(idx).to(cuda_device))
    assert not record_groups(only_key)[0]          # 10 of 128 bytes: moved as a dense column instead
    np.testing.assert_array_equal(only_key.to_numpy()["key"], key[idx])


def test_pitched_kernels_through_the_abi(cuda_device):
    """b200x_gather_rows_pitched / b200x_permute_rows_pitched / b200x_copy_rows_2d against numpy."""
    rng = np.random.default_rng(2)
    n, sp, dp, w = 4097, 128, 160, 112
    src = rng.integers(0, 256, (n, sp)).astype(np.uint8)
    idx = rng.integers(0, n, 3001).astype(np.int32)
    d_src = torch.from_numpy(src).to(cuda_device)
    d_idx = torch.from_numpy(idx).to(cuda_device)
    dst = torch.zeros((idx.size, dp), dtype=torch.uint8, device=cuda_device)
    st = torch.cuda.current_stream().cuda_stream
    _lib.call("b200x_gather_rows_pitched", d_src.data_ptr(), sp, d_idx.data_ptr(), idx.size, dst.data_ptr(), dp, w, st)
    got = dst.cpu().numpy()
    np.testing.assert_array_equal(got[:, :w], src[idx, :w])
    assert not got[:, w:].any()
    out = torch.zeros((n, dp), dtype=torch.uint8, device=cuda_device)
    _lib.call("b200x_permute_rows_pitched", d_src.data_ptr(), sp, n, 99, 0, out.data_ptr(), dp, w, st)
    p = orc.feistel_permutation(n, 99)
    np.testing.assert_array_equal(out.cpu().numpy()[:, :w], src[p, :w])
    for row_bytes, s_off in ((90, 10), (8, 16), (3, 1), (112, 0)):
        dense = torch.zeros((n, row_bytes), dtype=torch.uint8, device=cuda_device)
        _lib.call("b200x_copy_rows_2d", d_src.data_ptr() + s_off, sp, dense.data_ptr(), row_bytes, row_bytes, n, st)
        np.testing.assert_array_equal(dense.cpu().numpy(), src[:, s_off:s_off + row_bytes])
    lib = _lib.load()
    assert lib.b200x_gather_rows_pitched(d_src.data_ptr(), sp, d_idx.data_ptr(), 10, dst.data_ptr(), dp, 100, st) != 0
    assert lib.b200x_gather_rows_pitched(d_src.data_ptr(), sp, d_idx.data_ptr(), 10, dst.data_ptr(), dp, 256, st) != 0
    assert b"16" in lib.b200x_last_error_string() or b"pitch" in lib.b200x_last_error_string()


@pytest.mark.parametrize("desc", [False, True])
def test_fix_ties_against_lexsort(desc, cuda_device):
    """Sort by the first column + b200x_fix_ties == stable sort by all columns (runs up to 64 rows)."""
    rng = np.random.default_rng(12)
    n = 300_000
    a = rng.integers(0, n // 3, n, dtype=np.int64)            # runs of ~3 rows
    a[1000:1050] = 17                                          # one run of >= 50 rows
    b = rng.integers(0, 4, n).astype(np.uint16)
    c = rng.integers(-2, 2, n).astype(np.int32)
    da, db, dc = (torch.from_numpy(x.view(np.int16) if x.dtype == np.uint16 else x).to(cuda_device) for x in (a, b, c))
    db = db.view(torch.uint16)
    perm, _, skeys = ops.argsort(da, desc, want_keys=True)
    stats = ops.fix_ties(skeys, perm, [db, dc], desc).cpu().numpy()
    sgn = -1 if desc else 1
    want = np.lexsort([sgn * c.astype(np.int64), sgn * b.astype(np.int64), sgn * a])
    np.testing.assert_array_equal(perm.cpu().numpy().view(np.uint32), want.astype(np.uint32))
    assert stats[0] > 0 and 50 <= stats[1] <= 64
    # a run above the limit is reported and left alone
    a2 = a.copy()
    a2[:100] = 5
    perm2, _, sk2 = ops.argsort(torch.from_numpy(a2).to(cuda_device), desc, want_keys=True)
    assert ops.fix_ties(sk2, perm2, [db], desc).cpu().numpy()[1] > ops.FIX_TIES_MAX_RUN
    # through the accessor: the chain takes over for such columns
    blk = DeviceBlock.from_numpy({"a": a2, "b": b.view(np.int16), "c": c}, cuda_device)
    perm3, _, _ = BlockAccessor.for_block(blk)._argsort(SortKey(["a", "b", "c"], descending=[desc] * 3), False)
    want3 = np.lexsort([sgn * c.astype(np.int64), sgn * b.view(np.int16).astype(np.int64), sgn * a2])
    np.testing.assert_array_equal(perm3.cpu().numpy().view(np.uint32), want3.astype(np.uint32))


@pytest.mark.parametrize("width,offset,pitch", [(10, 0, 128), (10, 3, 48), (3, 13, 16), (17, 15, 64), (32, 9, 48), (8, 8, 16)])
def test_pack_key_chunks_any_alignment(width, offset, pitch, cuda_device):
    """b200x_pack_key_chunks (all chunks of a key held inside a record, one pass) == b200x_pack_key_bytes per chunk
    == numpy's big-endian view."""
    rng = np.random.default_rng(width * 100 + offset)
    n = 10_007
    rec = rng.integers(0, 256, (n, pitch)).astype(np.uint8)
    d = torch.from_numpy(rec).to(cuda_device)
    col = d[:, offset:offset + width]
    fused = ops.byte_key_chunks(col)
    assert len(fused) == (width + 7) // 8
    for i, ch in enumerate(fused):
        cw = min(8, width - 8 * i)
        single = ops.pack_key_bytes(col, 8 * i, cw)
        assert ch.dtype == single.dtype
        got = ch.view({2: torch.int16, 4: torch.int32, 8: torch.int64}[ch.element_size()]).cpu().numpy()
        np.testing.assert_array_equal(got, single.view(got.dtype if False else {2: torch.int16, 4: torch.int32, 8: torch.int64}[ch.element_size()]).cpu().numpy())
        padded = np.zeros((n, ch.element_size()), dtype=np.uint8)
        padded[:, :cw] = rec[:, offset + 8 * i: offset + 8 * i + cw]
        want = padded.view({2: ">u2", 4: ">u4", 8: ">u8"}[ch.element_size()]).ravel()
        np.testing.assert_array_equal(got.view({2: np.uint16, 4: np.uint32, 8: np.uint64}[ch.element_size()]), want.astype(want.dtype.newbyteorder("=")))


def test_key_prefix_column(cuda_device):
    """The key prefix column (block.py): built by to_records, used by the sort instead of a pass over the records,
    comes out of the sort, survives slice / concat, never shows in the schema; results do not depend on it."""
    key, payload = _gensort(120_000, 31, 1500, 300)
    plain = rd.from_numpy_columns({"key": key, "payload": payload}, override_num_blocks=3, records=True)
    withp = rd.from_numpy_columns({"key": key, "payload": payload}, override_num_blocks=3, records=True, key_prefix=["key"])
    b0 = withp.get_internal_block_refs()[0]
    assert list(b0.aux) == ["key::be0"] and b0.aux["key::be0"].dtype == torch.uint64
    assert b0.column_names == ["key", "payload"] and withp.schema() == plain.schema()
    want_hi = key[:b0.num_rows, :8].copy().view(">u8").astype(np.uint64).ravel()
    np.testing.assert_array_equal(b0.aux["key::be0"].view(torch.int64).cpu().numpy().view(np.uint64), want_hi)
    joined = DeviceBlock.concat(withp.get_internal_block_refs())
    assert joined.aux["key::be0"].numel() == 120_000 and joined.slice(5, 50).aux["key::be0"].numel() == 45
    for desc in (False, True):
        a = plain.sort("key", descending=desc)
        b = withp.sort("key", descending=desc)
        want = orc.stable_argsort_bytes(key, desc)
        for out in (a, b):
            got = out.to_numpy()
            np.testing.assert_array_equal(got["key"], key[want])
            np.testing.assert_array_equal(got["payload"], payload[want])
            for blk in out.get_internal_block_refs():       # the sorted blocks carry the prefix of THEIR keys
                if blk.num_rows:
                    hi = blk.to_numpy()["key"][:, :8].copy().view(">u8").astype(np.uint64).ravel()
                    np.testing.assert_array_equal(blk.aux["key::be0"].view(torch.int64).cpu().numpy().view(np.uint64), hi)
        again = b.sort("key", descending=not desc).to_numpy()    # sorting a sorted dataset uses the column it carries
        np.testing.assert_array_equal(again["payload"], payload[orc.stable_argsort_bytes(key, not desc)])
    sh = withp.random_shuffle(seed=4)
    assert not sh.get_internal_block_refs()[0].aux                # not maintained by shuffle: packed again on demand
    np.testing.assert_array_equal(sh.sort("key").to_numpy()["key"], key[orc.stable_argsort_bytes(key, False)])


@pytest.mark.parametrize("n,world", [(5000, 2), (300_001, 8)])
def test_record_exchange_kernels_with_local_peers(n, world, cuda_device):
    """b200x_scatter_records_by_target + b200x_gather_records_prefixed on one GPU, the 'peers' being local buffers:
    gensort records leave without their first 8 bytes (the key prefix, exchanged as a column) and without padding,
    in chunks of rows; the receiver puts the prefix back from the sorted keys while gathering into sorted order."""
    key, payload = _gensort(n, n, dup_hi=n // 20, dup_full=n // 50)
    blk = DeviceBlock.from_numpy({"key": key, "payload": payload}, cuda_device).to_records(key_prefix=["key"])
    rec = record_groups(blk)[0][0].rec
    hi = blk.aux["key::be0"]
    hi_np = key[:, :8].copy().view(">u8").astype(np.uint64).ravel()
    split = np.sort(np.random.default_rng(1).integers(0, 2 ** 63, world - 1).astype(np.uint64) * np.uint64(2))
    owner = np.searchsorted(split, hi_np, side="right")
    target, counts, table = ops.partition_targets(hi, False, torch.from_numpy(split.view(np.int64)).to(cuda_device),
                                                  want_tile_table=True)
    tab = table.cpu().numpy()
    for t in (0, 1, tab.shape[0] // 2, tab.shape[0] - 1):      # slots handed out before tile t = rows per owner before it
        np.testing.assert_array_equal(tab[t][:world], np.bincount(owner[: t * ops.PT_TILE_ROWS], minlength=world))
    cnt = np.bincount(owner, minlength=world)
    np.testing.assert_array_equal(counts.cpu().numpy()[:world], cnt)
    bufs = [torch.full((int(cnt[d]) + 3, 128), 7, dtype=torch.uint8, device=cuda_device) for d in range(world)]
    kbufs = [torch.zeros(int(cnt[d]) + 3, dtype=torch.int64, device=cuda_device) for d in range(world)]
    ops.scatter_rows_by_target(hi, target, [b.data_ptr() for b in kbufs], [3] * world)
    cuts = [0, n // 3 // 2048 * 2048, n // 2, n]           # three chunks of rows, sent one after the other
    for lo, hi_row in zip(cuts[:-1], cuts[1:]):
        ops.scatter_records_by_target(rec, target, [b.data_ptr() for b in bufs], [3] * world, 128, copy_bytes=96,
                                      src_offset=8, row0=lo, nrows=hi_row - lo)
    full = np.zeros((n, 128), dtype=np.uint8)
    full[:, :10], full[:, 10:100] = key, payload
    # dense slots (96 bytes apart, as the exchange receives them): written by one bulk copy per owner and tile, or by
    # 16-byte stores - the same bytes either way, nothing before the first or after the last slot
    for bulk in (1, 0):
        _lib.call("b200x_set_tuning", 4, bulk)
        dense = [torch.full((int(cnt[d]) + 5, 96), 7, dtype=torch.uint8, device=cuda_device) for d in range(world)]
        try:
            ops.scatter_records_by_target(rec, target, [b.data_ptr() for b in dense], [2] * world, 96, copy_bytes=96, src_offset=8)
        finally:
            _lib.call("b200x_set_tuning", 4, 1)
        for d in range(world):
            got = dense[d].cpu().numpy()
            np.testing.assert_array_equal(got[2:2 + int(cnt[d])], full[owner == d][:, 8:104])
            assert (got[:2] == 7).all() and (got[2 + int(cnt[d]):] == 7).all()
    for d in range(world):
        got = bufs[d].cpu().numpy()
        mine = full[owner == d]
        np.testing.assert_array_equal(got[3:, :96], mine[:, 8:104])
        assert (got[:, 96:] == 7).all() and (got[:3] == 7).all()       # nothing outside the 96 bytes of a slot
        # receiver d: sort the exchanged prefixes, repair ties from the key bytes in the rows, gather + re-insert
        m = int(cnt[d])
        if m < 2:
            continue
        recv, rkeys = bufs[d][3:], kbufs[d][3:].view(torch.uint64)
        perm, _, skeys = ops.argsort(rkeys, False, want_keys=True)
        out = torch.full((m, 128), 9, dtype=torch.uint8, device=cuda_device)
        # the rows "arrive" in 3 chunks from 2 sources: positions are grouped by chunk, gathered chunk by chunk
        a, b = m // 3, m // 2
        starts = [0, a // 2, a, b, b + (m - b) // 4, b + (m - b) // 2]           # source 0: slots [0, b), source 1: [b, m)
        sizes = [(a // 2 - 0) + (starts[4] - b), (a - a // 2) + (starts[5] - starts[4]), (b - a) + (m - starts[5])]
        lists = ops.split_positions_by_chunk(perm, starts, 2, 3, sizes)
        seen = np.sort(np.concatenate([x.cpu().numpy() for x in lists]))
        np.testing.assert_array_equal(seen, np.arange(m))                        # every position in exactly one list
        pn = perm.cpu().numpy().view(np.uint32)
        for c, lst in enumerate(lists):
            assert (np.diff(lst.cpu().numpy().astype(np.int64)) > 0).all()       # every list in ascending order
            slots = pn[lst.cpu().numpy()]
            in_c = ((slots >= starts[c]) & (slots < (starts[c + 1] if c < 2 else b))) | \
                   ((slots >= starts[3 + c]) & (slots < (starts[4 + c] if c < 2 else m)))
            assert in_c.all()
            ops.gather_records_prefixed(recv, perm, out, 96, 100, skeys, True, 8, positions=lst)
        ops.fix_ties(skeys, perm, [ops.KeyPiece(recv[:, 0:2], 0, 2)], False)      # key bytes 8..9 lie at bytes 0..1 of a row
        ops.gather_records_prefixed(recv, perm, out, 96, 100, skeys, True, 8, tie_mode=2)
        # the same in one pass, ties left out first (tie_mode 1) and put in afterwards
        out2 = torch.full((m, 128), 9, dtype=torch.uint8, device=cuda_device)
        ops.gather_records_prefixed(recv, perm, out2, 96, 100, skeys, True, 8, tie_mode=1)
        ops.gather_records_prefixed(recv, perm, out2, 96, 100, skeys, True, 8, tie_mode=2)
        np.testing.assert_array_equal(out2.cpu().numpy(), out.cpu().numpy())
        want = mine[orc.stable_argsort_bytes(mine[:, :10], False)]
        np.testing.assert_array_equal(out.cpu().numpy(), want)
    # shift 0 (whole rows, no prefix), all positions, record pitch that is not a power of two
    src = torch.from_numpy(np.random.default_rng(3).integers(0, 256, (1000, 48)).astype(np.uint8)).to(cuda_device)
    idx = torch.from_numpy(np.random.default_rng(4).permutation(1000).astype(np.int32)).to(cuda_device)
    out = torch.full((1000, 48), 9, dtype=torch.uint8, device=cuda_device)
    ops.gather_records_prefixed(src, idx, out, 48, 40)
    want = src.cpu().numpy()[idx.cpu().numpy()].copy()
    want[:, 40:] = 0
    np.testing.assert_array_equal(out.cpu().numpy(), want)


def test_row_oriented_ingest_and_egress(cuda_device):
    """Records that arrive row-oriented (one [n, 100] buffer, the gensort file format) become a record block with ONE
    copy per row run, and leave the same way (`rows_as_bytes`)."""
    key, payload = _gensort(20_000, 44, 300, 50)
    rows = np.concatenate([key, payload], axis=1)
    d = torch.from_numpy(rows).to(cuda_device)
    ds = rd.from_torch_columns({"key": d[:, :10], "payload": d[:, 10:]}, override_num_blocks=3, records=True, key_prefix=["key"])
    blk = DeviceBlock.concat(ds.get_internal_block_refs())
    np.testing.assert_array_equal(blk.rows_as_bytes().cpu().numpy(), rows)
    np.testing.assert_array_equal(blk.rows_as_bytes(["payload"]).cpu().numpy(), payload)
    out = DeviceBlock.concat(ds.sort("key").get_internal_block_refs())
    want = rows[orc.stable_argsort_bytes(key, False)]
    np.testing.assert_array_equal(out.rows_as_bytes().cpu().numpy(), want)
    # dense columns go through the same call
    dense = DeviceBlock.from_numpy({"key": key, "payload": payload}, cuda_device)
    np.testing.assert_array_equal(dense.rows_as_bytes().cpu().numpy(), rows)
