"""Multi-GPU parity check (run under torchrun, one rank per GPU):
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port 29501 scripts/dist_check.py
Each rank holds a shard; results of Dataset.sort / groupby().aggregate / random_shuffle across ranks are
gathered on rank 0 and compared with the CPU oracle run on the concatenated input."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
import torch.distributed as dist

import ray_project_ray_b200.data as rd
from oracle import exchange_oracle as orc
from ray_project_ray_b200.aggregate import Count, Mean, Std, Sum


def gather_np(arr, world, rank):
    objs = [None] * world
    dist.all_gather_object(objs, arr)
    return objs


def main():
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    local = int(os.environ.get("LOCAL_RANK", rank))
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    from ray_project_ray_b200.context import DataContext
    from ray_project_ray_b200 import distributed as D

    if torch.cuda.device_count() >= 2:
        # entry points run on the GPU that owns their buffers whatever device is current on the calling thread
        from ray_project_ray_b200 import _lib

        other = (local + 1) % torch.cuda.device_count()
        t = torch.zeros(1000, dtype=torch.int32, device=f"cuda:{local}")
        torch.cuda.synchronize()
        torch.cuda.set_device(other)
        _lib.call("b200x_iota_u32", t.data_ptr(), 1000, 5, None)
        assert torch.cuda.current_device() == other
        torch.cuda.set_device(local)
        torch.cuda.synchronize()
        assert t.cpu().tolist() == list(range(5, 1005))
        print(f"rank {rank}: call with another device current: OK", flush=True)
    n = 200_000 + 1000 * rank
    rng = np.random.default_rng(100 + rank)
    for transport, kind in (("nccl", "i64_ties"), ("nccl", "f64"), ("auto", "i64_ties"), ("auto", "f64"), ("p2p_sort_first", "i64_ties"), ("p2p_sort_first", "f64"), ("auto", "i64_ties")):
        DataContext.get_current().set_config("exchange_transport", transport)
        if kind == "i64_ties":
            key = rng.integers(0, 5000, n, dtype=np.int64)
        else:
            key = np.round(rng.standard_normal(n), 3)
            key[::50] = np.nan
            key[::70] = -0.0
        rowid = np.arange(n, dtype=np.int64) + rank * 10_000_000
        pay = rng.integers(0, 256, (n, 40), dtype=np.uint8)
        ds = rd.from_numpy_columns({"k": key, "row": rowid, "pay": pay}, override_num_blocks=4)
        for desc in (False, True):
            out = ds.sort("k", descending=desc)
            assert out.num_blocks() == 4
            got = out.to_numpy()
            parts = gather_np((got["row"], got["pay"].sum(axis=1)), world, rank)
            ins = gather_np((key, rowid, pay.sum(axis=1)), world, rank)
            if rank == 0:
                allk = np.concatenate([p[0] for p in ins])
                allr = np.concatenate([p[1] for p in ins])
                allp = np.concatenate([p[2] for p in ins])
                order = orc.stable_argsort(allk, desc)
                np.testing.assert_array_equal(np.concatenate([p[0] for p in parts]), allr[order])
                np.testing.assert_array_equal(np.concatenate([p[1] for p in parts]), allp[order])
                sizes = [len(p[0]) for p in parts]
                print(f"sort {kind} desc={desc} transport={transport} (p2p={D._p2p_ok['state']}): OK, rows per rank {sizes}", flush=True)
    # gensort-shaped records (configs[3]): 10-byte key as (hi uint64, lo uint16) + 90-byte payload
    m = 60_000
    hi = rng.integers(0, 2**63, m, dtype=np.int64).astype(np.uint64) * 2 + rng.integers(0, 2, m).astype(np.uint64)
    hi[: m // 3] = hi[m // 3: 2 * (m // 3)]
    lo = rng.integers(0, 2**16, m).astype(np.uint16)
    pay = rng.integers(0, 256, (m, 90), dtype=np.uint8)
    out = rd.from_numpy_columns({"hi": hi, "lo": lo, "payload": pay}, override_num_blocks=2).sort(["hi", "lo"]).to_numpy()
    parts = gather_np((out["hi"], out["lo"], out["payload"].sum(axis=1)), world, rank)
    ins = gather_np((hi, lo, pay.sum(axis=1)), world, rank)
    if rank == 0:
        ahi, alo, ap = (np.concatenate([p[i] for p in ins]) for i in range(3))
        order = orc.stable_argsort_multi([ahi, alo], False)
        np.testing.assert_array_equal(np.concatenate([p[0] for p in parts]), ahi[order])
        np.testing.assert_array_equal(np.concatenate([p[1] for p in parts]), alo[order])
        np.testing.assert_array_equal(np.concatenate([p[2] for p in parts]), ap[order])
        print("gensort-shaped 2-column key sort: OK", flush=True)
    # gensort records in the record layout (the bench workload): [n, 10] byte key + [n, 90] payload, 128-byte records,
    # with and without the key prefix column; ties on the first 8 bytes and exact duplicates inside and across ranks
    shared = np.random.default_rng(5).integers(0, 256, (500, 10)).astype(np.uint8)      # the same keys on every rank
    for m, prefix in ((300_000 + 777 * rank, ["key"]), (300_000 + 777 * rank, None), (90_000, ["key"]), (300_123, ["key"]), (0 if rank == 1 else 50_000, ["key"])):
        # the fourth case sends the records in 4 chunks with overlapped gathers, the last one leaves rank 1 without rows
        os.environ["B200X_EXCHANGE_CHUNK_MIN_ROWS"] = "1000" if m == 300_123 else str(1 << 22)
        bkey = rng.integers(0, 256, (m, 10)).astype(np.uint8)
        if m:
            dst, src = rng.integers(0, m, m // 50), rng.integers(0, m, m // 50)
            bkey[dst, :8] = bkey[src, :8]
            bkey[rng.integers(0, m, 3000)] = shared[rng.integers(0, 500, 3000)]
        if m == 90_000:
            bkey[:, :8] = bkey[rng.integers(0, 40, m), :8]      # duplicate-heavy first 8 bytes: the tie repair gives up
        bpay = rng.integers(0, 256, (m, 90), dtype=np.uint8)
        bpay[:, :8] = (np.arange(m, dtype=np.int64) + rank * 10_000_000).view(np.uint8).reshape(m, 8)
        dsr = rd.from_numpy_columns({"key": bkey, "payload": bpay}, override_num_blocks=3, records=True, key_prefix=prefix)
        for desc in (False, True):
            got = dsr.sort("key", descending=desc).to_numpy()
            parts = gather_np((got["key"], got["payload"]), world, rank)
            ins = gather_np((bkey, bpay), world, rank)
            if rank == 0:
                ak, ap = (np.concatenate([p[i] for p in ins]) for i in range(2))
                order = orc.stable_argsort_bytes(ak, desc)
                np.testing.assert_array_equal(np.concatenate([p[0] for p in parts]), ak[order])
                np.testing.assert_array_equal(np.concatenate([p[1] for p in parts]), ap[order])
                print(f"gensort records rows={m} prefix={bool(prefix)} desc={desc}: OK, rows per rank {[len(p[0]) for p in parts]}", flush=True)
    # user boundaries across ranks (planner/sort.py:47-54): the partitions are dealt to the ranks in order; the blocks of
    # all ranks, in rank order, are the reference's partitions (rows equal to a boundary included on the right side)
    bk = rng.integers(0, 1000, 50_000 + 100 * rank, dtype=np.int64)
    brow = np.arange(len(bk), dtype=np.int64) + rank * 10_000_000
    for bounds in ([100, 250, 251, 600, 999, 5000], [500], [10, 20, 30, 40, 50, 60, 70, 80, 90]):
        for desc in (False, True):
            res = rd.from_numpy_columns({"k": bk, "row": brow}, override_num_blocks=3).sort("k", descending=desc, boundaries=bounds)
            sizes = [b.num_rows for b in res.get_internal_block_refs()]
            got = res.to_numpy()["row"]
            parts, ins = gather_np((sizes, got), world, rank), gather_np((bk, brow), world, rank)
            if rank == 0:
                ak, ar = np.concatenate([x[0] for x in ins]), np.concatenate([x[1] for x in ins])
                order = orc.stable_argsort(ak, desc)
                np.testing.assert_array_equal(np.concatenate([p[1] for p in parts]), ar[order])
                sk = ak[order]
                bs = sorted(bounds, reverse=desc)
                cuts = [int((sk >= b).sum()) if desc else int((sk < b).sum()) for b in bs]
                want_sizes = np.diff([0] + cuts + [len(sk)]).tolist()
                assert [x for p in parts for x in p[0]] == want_sizes, (bounds, desc, [p[0] for p in parts], want_sizes)
                print(f"sort with user boundaries {bounds} desc={desc}: OK, blocks per rank {[len(p[0]) for p in parts]}", flush=True)
    # groupby by two key columns across ranks
    g1, g2 = rng.integers(0, 40, n, dtype=np.int64), rng.integers(0, 25, n, dtype=np.int64)
    gvv = rng.integers(-100, 100, n, dtype=np.int64)
    got = rd.from_numpy_columns({"a": g1, "b": g2, "v": gvv}, override_num_blocks=4).groupby(["a", "b"]).aggregate(Sum("v"), Count()).to_numpy()
    parts, ins = gather_np(got, world, rank), gather_np((g1, g2, gvv), world, rank)
    if rank == 0:
        A1, A2, AV = (np.concatenate([p[i] for p in ins]) for i in range(3))
        cat = {c: np.concatenate([p[c] for p in parts]) for c in parts[0]}
        combo = A1 * 1000 + A2
        keys, inv = np.unique(combo, return_inverse=True)
        np.testing.assert_array_equal(cat["a"] * 1000 + cat["b"], keys)
        np.testing.assert_array_equal(cat["sum(v)"], np.bincount(inv, weights=AV).astype(np.int64))
        np.testing.assert_array_equal(cat["count()"], np.bincount(inv))
        print("groupby by two key columns: OK, groups per rank", [len(p["a"]) for p in parts], flush=True)
    # groupby
    gk = rng.integers(0, 3000, n, dtype=np.int64)
    gv = rng.integers(-2**31, 2**31, n, dtype=np.int64)
    gf = rng.random(n)
    ds = rd.from_numpy_columns({"k": gk, "v": gv, "f": gf}, override_num_blocks=4)
    got = ds.groupby("k").aggregate(Sum("v"), Mean("f"), Count(), Std("f")).to_numpy()
    parts = gather_np(got, world, rank)
    ins = gather_np((gk, gv, gf), world, rank)
    if rank == 0:
        allk, allv, allf = (np.concatenate([p[i] for p in ins]) for i in range(3))
        want = orc.concat(orc.aggregate_dataset([{"k": allk, "v": allv, "f": allf}], "k",
                                                [("sum", "v"), ("mean", "f"), ("count", None), ("std", "f")], []))
        cat = {c: np.concatenate([p[c] for p in parts]) for c in parts[0]}
        np.testing.assert_array_equal(cat["k"], want["k"])
        np.testing.assert_array_equal(cat["sum(v)"], want["sum(v)"])
        np.testing.assert_array_equal(cat["count()"], want["count()"])
        np.testing.assert_allclose(cat["mean(f)"], want["mean(f)"], rtol=1e-6)
        np.testing.assert_allclose(cat["std(f)"], want["std(f)"], rtol=1e-6)
        print("groupby: OK, groups per rank", [len(p["k"]) for p in parts], flush=True)
    g = ds.aggregate(Sum("v"), Count())
    if rank == 0:
        assert g == {"sum(v)": int(np.concatenate([p[1] for p in ins]).sum()), "count()": sum(len(p[0]) for p in ins)}, g
        print("global aggregate: OK", flush=True)
    # random shuffle
    rowid = np.arange(n, dtype=np.int64) + rank * 10_000_000
    ds = rd.from_numpy_columns({"row": rowid}, override_num_blocks=4)
    a = ds.random_shuffle(seed=0).to_numpy()["row"]
    b = ds.random_shuffle(seed=0).to_numpy()["row"]
    c = ds.random_shuffle(seed=1).to_numpy()["row"]
    pa, pb, pc = (gather_np(x, world, rank) for x in (a, b, c))
    ins = gather_np(rowid, world, rank)
    if rank == 0:
        A, B, C = (np.concatenate(x) for x in (pa, pb, pc))
        assert np.array_equal(A, B) and (A != C).mean() > 0.9
        assert np.array_equal(np.sort(A), np.sort(np.concatenate(ins)))
        for x in pa:
            assert len(set(x // 10_000_000)) == world  # every rank's output draws from every rank
        print("random_shuffle: OK (deterministic per seed, exact multiset), rows per rank", [len(x) for x in pa], flush=True)
        if D._p2p_ok["state"]:     # over peer memory: exactly the single-GPU shuffle of the concatenated input
            allrows = np.concatenate(ins)
            np.testing.assert_array_equal(A, allrows[orc.feistel_permutation(len(allrows), 0)])
            print("random_shuffle: equals the one-GPU permutation of the concatenated input", flush=True)
    # the same with a payload, as records, with unequal shard sizes; and a rank without rows
    for m, recs in ((123_457 + 1000 * rank, True), (123_457 + 1000 * rank, False), (0 if rank == 1 else 77_000, True)):
        rid = np.arange(m, dtype=np.int64) + rank * 10_000_000
        pay = np.random.default_rng(rank).integers(0, 256, (m, 40), dtype=np.uint8)
        got = rd.from_numpy_columns({"row": rid, "pay": pay}, override_num_blocks=3, records=recs).random_shuffle(seed=9).to_numpy()
        outs, ins2 = gather_np((got["row"], got["pay"]), world, rank), gather_np((rid, pay), world, rank)
        if rank == 0:
            ar, ap = np.concatenate([x[0] for x in ins2]), np.concatenate([x[1] for x in ins2])
            p = orc.feistel_permutation(len(ar), 9)
            np.testing.assert_array_equal(np.concatenate([x[0] for x in outs]), ar[p])
            np.testing.assert_array_equal(np.concatenate([x[1] for x in outs]), ap[p])
            print(f"random_shuffle rows={m} records={recs}: OK, rows per rank {[len(x[0]) for x in outs]}", flush=True)
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
