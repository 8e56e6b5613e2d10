"""BASELINE.json sizes on one GPU, checked through size-independent properties (the oracle cannot
run at these sizes in seconds): sortedness, stable tie order, row integrity (every payload still
belongs to its key), multiset checksums, permutation round trips, conservation of counts and sums.
torch ops are used here only as the CHECKER."""
import numpy as np
import pytest
import torch

import ray_project_ray_b200.data as rd
from ray_project_ray_b200.aggregate import Count, Mean, Sum

pytestmark = pytest.mark.gpu


def _mix(x: torch.Tensor) -> torch.Tensor:
    """A cheap 64-bit mixer used to tie a payload to its key / row id."""
    x = x * -7046029254386353131
    return x ^ (x >> 29)


def test_config1_sort_desc_10m_rows_8_payload_columns(cuda_device):
    """configs[0]: range(10_000_000) ids (here permuted) + 8 int64 payload columns, sort('id', descending=True)."""
    n = 10_000_000
    g = torch.Generator(device=cuda_device).manual_seed(1)
    ids = torch.randperm(n, device=cuda_device, generator=g)
    cols = {"id": ids}
    for j in range(8):
        cols[f"p{j}"] = _mix(ids + j)
    out = rd.from_torch_columns(cols, override_num_blocks=8).sort("id", descending=True)
    assert out.num_blocks() == 8 and out.count() == n
    got = {k: torch.cat([b[k] for b in out.get_internal_block_refs()]) for k in cols}
    assert torch.equal(got["id"], torch.arange(n - 1, -1, -1, device=cuda_device))
    for j in range(8):
        assert torch.equal(got[f"p{j}"], _mix(got["id"] + j))


def test_config2_shape_sort_100m_rows_64_byte_payload(cuda_device):
    """The bench workload: 100 M rows x (int64 key + 64-byte payload)."""
    n = 100_000_000
    g = torch.Generator(device=cuda_device).manual_seed(2)
    keys = torch.randint(0, 2**63 - 1, (n,), dtype=torch.int64, device=cuda_device, generator=g)
    keys[: n // 100] = keys[n // 100: 2 * (n // 100)]          # 1 % exact duplicates: ties
    pay = torch.empty((n, 8), dtype=torch.int64, device=cuda_device)
    rowid = torch.arange(n, device=cuda_device)
    pay[:, 0] = _mix(keys)        # belongs to the key
    pay[:, 1] = rowid             # original position: tie order
    for j in range(2, 8):
        pay[:, j] = _mix(keys + j)
    ksum, kxor = int(keys.sum()), int(_mix(keys).sum())
    ds = rd.from_torch_columns({"id": keys, "payload": pay.view(torch.uint8).reshape(n, 64)}, override_num_blocks=8)
    del pay
    out = ds.sort("id")
    blocks = out.get_internal_block_refs()
    assert len(blocks) == 8 and sum(b.num_rows for b in blocks) == n
    k = torch.cat([b["id"] for b in blocks])
    p = torch.cat([b["payload"] for b in blocks]).view(torch.int64).reshape(n, 8)
    assert bool((k[1:] >= k[:-1]).all())                                   # sorted
    assert int(k.sum()) == ksum and int(_mix(k).sum()) == kxor               # same multiset of keys
    assert torch.equal(p[:, 0], _mix(k))                                     # every payload travelled with its key
    for j in range(2, 8):
        assert torch.equal(p[:, j], _mix(k + j))
    ties = k[1:] == k[:-1]
    assert int(ties.sum()) >= n // 100
    assert bool((p[1:, 1][ties] > p[:-1, 1][ties]).all())                    # equal keys keep input order (stable)
    # equal keys never straddle output blocks
    edges = np.cumsum([b.num_rows for b in blocks])[:-1]
    for e in edges:
        assert int(k[e]) != int(k[e - 1])


def test_config2_random_shuffle_100m_rows(cuda_device):
    """configs[1]: random_shuffle(seed=0) of 100 M rows x 64-byte payload as a per-seed multiset permutation."""
    n = 100_000_000
    ids = torch.arange(n, device=cuda_device)
    pay = torch.empty((n, 8), dtype=torch.int64, device=cuda_device)
    for j in range(8):
        pay[:, j] = _mix(ids + j)
    ds = rd.from_torch_columns({"id": ids, "payload": pay.view(torch.uint8).reshape(n, 64)}, override_num_blocks=8)
    del pay
    a = ds.random_shuffle(seed=0)
    ia = torch.cat([b["id"] for b in a.get_internal_block_refs()])
    pa = torch.cat([b["payload"] for b in a.get_internal_block_refs()]).view(torch.int64).reshape(n, 8)
    for j in range(8):
        assert torch.equal(pa[:, j], _mix(ia + j))                          # rows stayed intact
    del pa
    assert torch.equal(torch.sort(ia).values, ids)                          # exact permutation
    assert float((ia != ids).float().mean()) > 0.99
    b = ds.random_shuffle(seed=0)
    assert torch.equal(torch.cat([x["id"] for x in b.get_internal_block_refs()]), ia)   # deterministic per seed
    del b
    c = ds.random_shuffle(seed=1)
    ic = torch.cat([x["id"] for x in c.get_internal_block_refs()])
    assert float((ic != ia).float().mean()) > 0.99
    # every output block draws from every input block
    for blk in a.get_internal_block_refs():
        assert len(torch.unique(blk["id"] // (n // 8))) == 8


@pytest.mark.parametrize("skew", ["uniform", "zipf"])
def test_config3_groupby_200m_rows_1m_keys(cuda_device, skew):
    """configs[2]: groupby(int64 key, 1 M distinct).aggregate(Sum, Mean, Count) over 200 M rows."""
    n, K = 200_000_000, 1_000_000
    g = torch.Generator(device=cuda_device).manual_seed(4)
    if skew == "uniform":
        k = torch.randint(0, K, (n,), dtype=torch.int64, device=cuda_device, generator=g)
    else:  # Zipf(1.1) ranks by inverse CDF, ranks scattered over the key space
        w = torch.arange(1, K + 1, device=cuda_device, dtype=torch.float64).pow(-1.1)
        cdf = torch.cumsum(w, 0)
        cdf /= cdf[-1].clone()
        u = torch.rand(n, device=cuda_device, dtype=torch.float64, generator=g)
        ranks = torch.searchsorted(cdf, u).clamp_(max=K - 1)
        del u
        k = (ranks * 7919) % K      # 7919 is coprime with 10^6: a bijection of the key space
        del ranks
    v = torch.randint(-2**31, 2**31, (n,), dtype=torch.int64, device=cuda_device, generator=g)
    out = rd.from_torch_columns({"k": k, "v": v}, override_num_blocks=8).groupby("k").aggregate(Sum("v"), Mean("v"), Count())
    assert out.num_blocks() == 8
    blocks = out.get_internal_block_refs()
    keys = torch.cat([b["k"] for b in blocks])
    sums = torch.cat([b["sum(v)"] for b in blocks])
    means = torch.cat([b["mean(v)"] for b in blocks])
    counts = torch.cat([b["count()"] for b in blocks])
    assert out.columns() == ["k", "sum(v)", "mean(v)", "count()"]
    assert bool((keys[1:] > keys[:-1]).all())                               # ascending, one row per key
    want_counts = torch.bincount(k, minlength=K)
    present = want_counts > 0
    assert torch.equal(keys, torch.nonzero(present).flatten())
    assert torch.equal(counts, want_counts[present])                        # exact counts
    want_sums = torch.zeros(K, dtype=torch.int64, device=cuda_device).scatter_add_(0, k, v)
    assert torch.equal(sums, want_sums[present])                            # exact integer sums
    torch.testing.assert_close(means, sums.double() / counts.double(), rtol=1e-6, atol=0)


def test_config4_shape_gensort_records_30m(cuda_device):
    """configs[3] shape on one GPU: 100-byte records = 10-byte key + 90-byte payload, sorted by the byte-string
    key. Checked by the packed key (8 big-endian bytes, then 2), row integrity (the payload's first 8 bytes are
    the record number, from which the key can be recomputed) and the record-number checksum."""
    from ray_project_ray_b200 import ops

    n = 30_000_000
    g = torch.Generator(device=cuda_device).manual_seed(9)
    recno = torch.arange(n, dtype=torch.int64, device=cuda_device)
    key = torch.empty((n, 10), dtype=torch.uint8, device=cuda_device)
    key[:, :8] = _mix(recno).view(torch.uint8).reshape(n, 8)
    key[:, 8:] = _mix(recno + 12345).view(torch.uint8).reshape(n, 8)[:, :2]
    key[: n // 1000, :8] = key[0, :8]                        # 30 000 records tie on the first 8 bytes: the last 2 decide
    payload = torch.randint(0, 256, (n, 90), dtype=torch.uint8, device=cuda_device, generator=g)
    payload[:, :8] = recno.view(torch.uint8).reshape(n, 8)
    out = rd.from_torch_columns({"key": key, "payload": payload}, override_num_blocks=4).sort("key")
    assert out.columns() == ["key", "payload"] and out.count() == n and out.num_blocks() == 4
    skey = torch.cat([b["key"] for b in out.get_internal_block_refs()])
    spay = torch.cat([b["payload"] for b in out.get_internal_block_refs()])
    hi = ops.pack_key_bytes(skey, 0, 8).view(torch.int64) ^ (-0x8000000000000000)   # unsigned order as signed
    lo = ops.pack_key_bytes(skey, 8, 2).view(torch.int16).to(torch.int32) & 0xFFFF
    assert bool(((hi[1:] > hi[:-1]) | ((hi[1:] == hi[:-1]) & (lo[1:] >= lo[:-1]))).all())
    got_rec = spay[:, :8].contiguous().view(torch.int64).flatten()
    assert torch.equal(skey, key[got_rec])                                           # every payload still has its key
    assert int(got_rec.sum().item()) == n * (n - 1) // 2


def test_config5_shape_zipf_keys_50m(cuda_device):
    """configs[4] shape on one GPU: Zipf(1.2) int64 keys (ranks hashed over the key space): heavy duplicates send
    the sort down its LSD path. Sortedness, stability (equal keys keep their row order) and the multiset."""
    n, K = 50_000_000, 1 << 22
    g = torch.Generator(device=cuda_device).manual_seed(12)
    w = torch.arange(1, K + 1, device=cuda_device, dtype=torch.float64).pow(-1.2)
    cdf = torch.cumsum(w, 0)
    cdf /= cdf[-1].clone()
    ranks = torch.searchsorted(cdf, torch.rand(n, device=cuda_device, dtype=torch.float64, generator=g)).clamp_(max=K - 1)
    k = _mix(ranks)                                           # spread the ranks over the int64 space
    del ranks, cdf, w
    row = torch.arange(n, dtype=torch.int64, device=cuda_device)
    out = rd.from_torch_columns({"k": k, "row": row}, override_num_blocks=8).sort("k")
    sk = torch.cat([b["k"] for b in out.get_internal_block_refs()])
    sr = torch.cat([b["row"] for b in out.get_internal_block_refs()])
    assert bool((sk[1:] >= sk[:-1]).all())
    assert bool(((sk[1:] > sk[:-1]) | (sr[1:] > sr[:-1])).all())                     # ties in input order
    assert torch.equal(k[sr], sk)
    assert int(sr.sum().item()) == n * (n - 1) // 2
    blocks = out.get_internal_block_refs()                                           # equal keys never straddle blocks
    for a, b in zip(blocks[:-1], blocks[1:]):
        if a.num_rows and b.num_rows:
            assert int(a["k"][-1]) < int(b["k"][0])
