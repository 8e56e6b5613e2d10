"""world_size-2 gloo tests (CPU) of the host side of the multi-GPU exchange: splitter choice,
count exchange and the all-to-all-v of rows with ragged splits, in source-rank order."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from ray_project_ray_b200 import distributed as D


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, q):
    try:
        os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
        dist.init_process_group("gloo", rank=rank, world_size=world)
        assert D.is_distributed()
        # every rank draws its own keys, sorts them (numpy stands in for the device sort), then runs the
        # same planning + exchange code the GPU path runs
        rng = np.random.default_rng(10 + rank)
        n = 5000 + 700 * rank
        keys = np.sort(rng.integers(0, 1000, n).astype(np.uint64))
        rowid = (np.arange(n, dtype=np.int64) + rank * 1_000_000)
        pay = rng.integers(0, 256, (n, 24), dtype=np.uint8)
        samples = keys[:: max(1, n // 64)][:64]
        padded = torch.full((65,), -1, dtype=torch.int64)
        padded[0] = len(samples)
        padded[1:1 + len(samples)] = torch.from_numpy(samples.view(np.int64))
        gathered = [torch.empty_like(padded) for _ in range(world)]
        dist.all_gather(gathered, padded)
        host = torch.stack(gathered).numpy()
        all_codes = np.concatenate([r[1:1 + int(r[0])].view(np.uint64) for r in host])
        splitters = D.choose_splitters(all_codes, world)
        cuts = np.searchsorted(keys, splitters, side="left").tolist()
        send_counts = D.split_counts_from_cuts(cuts, n)
        C = D.all_gather_counts(send_counts)
        assert C.shape == (world, world) and C[rank].tolist() == send_counts
        recv_counts = C[:, rank].tolist()
        got_keys = D.exchange_rows(torch.from_numpy(keys.view(np.int64)), send_counts, recv_counts)
        got_rows = D.exchange_rows(torch.from_numpy(rowid), send_counts, recv_counts)
        got_pay = D.exchange_rows(torch.from_numpy(pay), send_counts, recv_counts)
        q.put((rank, splitters.tolist(), send_counts, got_keys.numpy().view(np.uint64).tolist(), got_rows.numpy().tolist(),
               got_pay.numpy().sum(axis=1).tolist(), keys.tolist(), rowid.tolist(), pay.sum(axis=1).tolist()))
        dist.barrier()
        dist.destroy_process_group()
    except Exception as e:  # surface failures to the parent
        q.put(("error", rank, repr(e)))
        raise


def test_two_rank_range_exchange_on_gloo():
    world = 2
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, q)) for r in range(world)]
    for p in procs:
        p.start()
    results = {}
    for _ in range(world):
        item = q.get(timeout=120)
        assert item[0] != "error", item
        results[item[0]] = item[1:]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    # identical splitters on every rank
    assert results[0][0] == results[1][0] and len(results[0][0]) == world - 1
    split = results[0][0][0]
    all_keys = np.concatenate([np.asarray(results[r][5], dtype=np.uint64) for r in range(world)])
    all_rows = np.concatenate([np.asarray(results[r][6]) for r in range(world)])
    all_pay = np.concatenate([np.asarray(results[r][7]) for r in range(world)])
    for r in range(world):
        _, send_counts, k, rows, paysum, *_ = results[r]
        k = np.asarray(k, dtype=np.uint64)
        # rank r owns exactly one key range; equal keys never straddle ranks
        assert ((k < split) if r == 0 else (k >= split)).all()
        want = (all_keys < split) if r == 0 else (all_keys >= split)
        # received rows are in source-rank order, and within a source in that source's (sorted) order
        np.testing.assert_array_equal(np.asarray(rows), all_rows[want])
        np.testing.assert_array_equal(k, all_keys[want])
        np.testing.assert_array_equal(np.asarray(paysum), all_pay[want])
    assert sum(len(results[r][2]) for r in range(world)) == len(all_keys)


def test_choose_splitters_and_cuts():
    codes = np.arange(1000, dtype=np.uint64)[::-1].copy()
    s = D.choose_splitters(codes, 4)
    assert s.tolist() == [250, 500, 749] or s.tolist() == [250, 499, 749] or len(s) == 3
    assert (np.diff(s.astype(np.int64)) > 0).all()
    assert D.choose_splitters(np.zeros(0, np.uint64), 4).size == 0
    assert D.choose_splitters(codes, 1).size == 0
    assert D.split_counts_from_cuts([10, 10, 70], 100) == [10, 0, 60, 30]
    assert D.split_counts_from_cuts([], 5) == [5]
    assert D.split_counts_from_cuts([7, 3], 10) == [7, 0, 3]  # non-monotone cuts are clamped


def _schema_worker(rank, world, port, q):
    try:
        os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
        dist.init_process_group("gloo", rank=rank, world_size=world)
        from ray_project_ray_b200.block import DeviceBlock

        cpu = torch.device("cpu")
        full = DeviceBlock({"k": torch.arange(5, dtype=torch.int64), "p": torch.zeros((5, 3), dtype=torch.uint8)}, device=cpu)
        # rank 1 has no rows and no columns: it adopts rank 0's schema and takes part with zero rows
        mine = full if rank == 0 else DeviceBlock.empty(cpu)
        got = D.adopt_schema(mine, cpu)
        assert got.column_names == ["k", "p"] and got.columns["p"].shape[1:] == (3,) and got.columns["k"].dtype == torch.int64
        assert got.num_rows == (5 if rank == 0 else 0)
        # nobody has columns -> None on every rank; everybody has columns -> unchanged, no broadcast
        assert D.adopt_schema(DeviceBlock.empty(cpu), cpu) is None
        assert D.adopt_schema(full, cpu) is full
        # the zero-row block goes through the all-to-all-v like any other
        out = D.exchange_block(got, [got.num_rows, 0])
        assert out.num_rows == (5 if rank == 0 else 0) and out.column_names == ["k", "p"]
        q.put(("ok", rank))
        dist.barrier()
        dist.destroy_process_group()
    except Exception as e:
        q.put(("error", rank, repr(e)))
        raise


def test_rank_without_rows_adopts_the_schema_on_gloo():
    world, port = 2, _free_port()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_schema_worker, args=(r, world, port, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = [q.get(timeout=120) for _ in range(world)]
    for p in procs:
        p.join(timeout=60)
    assert all(r[0] == "ok" for r in res), res


def test_rank_splitters_from_user_boundaries():
    """Which boundary separates which ranks (planner/sort.py:47-54 across ranks): R partitions dealt to W ranks in
    order; equal-to-boundary rows go right when ascending, left when descending (util.py:745-758)."""
    codes = np.array([10, 20, 30, 40, 50], dtype=np.uint64)            # 6 partitions
    sp, edges = D.rank_splitters_from_boundaries(codes, 3, False)
    assert edges == [0, 2, 4, 6] and sp.tolist() == [20, 40]
    sp, edges = D.rank_splitters_from_boundaries(codes, 3, True)       # descending: boundary + 1 (codes are complemented)
    assert sp.tolist() == [21, 41]
    sp, edges = D.rank_splitters_from_boundaries(np.array([7], dtype=np.uint64), 4, False)   # 2 partitions, 4 ranks
    assert edges == [0, 0, 1, 1, 2] and sp.tolist() == [0, 7, 7]
    top = np.uint64(0xFFFFFFFFFFFFFFFF)
    sp, _ = D.rank_splitters_from_boundaries(np.array([top], dtype=np.uint64), 2, True)
    assert sp.tolist() == [int(top)]
    # ownership under the device rule: owner = number of splitters <= code
    sp, edges = D.rank_splitters_from_boundaries(codes, 3, False)
    owner = lambda c: int((sp <= np.uint64(c)).sum())                   # noqa: E731
    assert [owner(c) for c in (0, 19, 20, 39, 40, 99)] == [0, 0, 1, 1, 2, 2]


def test_chunk_layout_of_a_chunked_exchange():
    """Bookkeeping of the chunked record exchange: which receive slots every (source rank, chunk) fills. Checked
    against a brute-force simulation of the stable split."""
    rng = np.random.default_rng(5)
    world, K = 3, 4
    n = [1000, 0, 777]
    owners = [rng.integers(0, world, m) for m in n]
    cuts = [[(c * m) // K for c in range(K)] + [m] for m in n]                      # chunk c = rows [cuts[c], cuts[c+1])
    tables = np.zeros((world, K + 1, world), dtype=np.int64)
    for r in range(world):
        for c in range(K + 1):
            tables[r, c] = np.bincount(owners[r][:cuts[r][c]], minlength=world)
    for me in range(world):
        starts, sizes = D.chunk_layout_from_tables(tables, me)
        assert starts == sorted(starts) and sum(sizes) == sum(int((o == me).sum()) for o in owners)
        # simulate: the receive buffer is filled source by source, rows in row order; label every slot with its chunk
        label = []
        for r in range(world):
            for i in np.flatnonzero(owners[r] == me):
                label.append((r, int(np.searchsorted(cuts[r], i, side="right")) - 1))
        for r in range(world):
            for c in range(K):
                lo = starts[r * K + c]
                hi = starts[r * K + c + 1] if r * K + c + 1 < len(starts) else len(label)
                got = [x for x in label[lo:hi] if x[0] == r]                          # (empty chunks share their start)
                assert all(x == (r, c) for x in got) and len(got) == int(tables[r, c + 1, me] - tables[r, c, me])
        assert sizes == [sum(1 for x in label if x[1] == c) for c in range(K)]
