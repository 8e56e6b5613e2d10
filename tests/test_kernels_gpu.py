"""GPU parity tests of the individual C-ABI kernels against numpy restatements."""
import numpy as np
import pytest
import torch

from ray_project_ray_b200 import ops
from oracle import exchange_oracle as orc

pytestmark = pytest.mark.gpu


def _u32(t):
    return t.cpu().numpy().view(np.uint32)


def _keys(kind, n, rng):
    if kind == "i64_full":
        return rng.integers(np.iinfo(np.int64).min, np.iinfo(np.int64).max, n, dtype=np.int64)
    if kind == "i64_small":
        return rng.integers(-500, 500, n, dtype=np.int64)
    if kind == "i64_perm":
        return rng.permutation(n).astype(np.int64)
    if kind == "i64_const":
        return np.full(n, 7, dtype=np.int64)
    if kind == "f64":
        x = rng.standard_normal(n)
        if n >= 16:
            x[rng.integers(0, n, n // 8)] = np.nan
            x[rng.integers(0, n, n // 8)] = 0.0
            x[rng.integers(0, n, n // 8)] = -0.0
            x[rng.integers(0, n, 3)] = np.inf
            x[rng.integers(0, n, 3)] = -np.inf
            x[rng.integers(0, n, 3)] = np.float64(np.nan) * -1
        return x
    if kind == "f64_ties":
        return rng.integers(-20, 20, n).astype(np.float64) / 4
    if kind == "i32":
        return rng.integers(-2**31, 2**31 - 1, n, dtype=np.int64).astype(np.int32)
    if kind == "f32":
        x = rng.standard_normal(n).astype(np.float32)
        if n >= 16:
            x[rng.integers(0, n, n // 8)] = np.nan
            x[rng.integers(0, n, n // 8)] = -0.0
        return x
    if kind == "u8":
        return rng.integers(0, 256, n, dtype=np.int64).astype(np.uint8)
    if kind == "i16":
        return rng.integers(-2**15, 2**15, n, dtype=np.int64).astype(np.int16)
    raise KeyError(kind)


@pytest.mark.parametrize("kind", ["i64_full", "i64_small", "i64_perm", "i64_const", "f64", "f64_ties", "i32", "f32", "u8", "i16"])
@pytest.mark.parametrize("n", [0, 1, 2, 31, 33, 1000, 6143, 6144, 6145, 100_003, 1_500_000])
@pytest.mark.parametrize("descending", [False, True])
def test_argsort_matches_oracle(cuda_device, kind, n, descending):
    rng = np.random.default_rng(1234 + n)
    k = _keys(kind, n, rng)
    want = orc.stable_argsort(k, descending)
    t = torch.from_numpy(k).to(cuda_device)
    perm, codes, skeys = ops.argsort(t, descending, want_codes=True, want_keys=True)
    torch.cuda.synchronize()
    got = _u32(perm)
    assert got.dtype == np.uint32 and got.shape == (n,)
    np.testing.assert_array_equal(got, want.astype(np.uint32))
    want_codes = orc.encode_keys(k, descending)[want]
    np.testing.assert_array_equal(codes.cpu().numpy().view(np.uint64), want_codes)
    if k.dtype.kind in "iu":
        np.testing.assert_array_equal(skeys.cpu().numpy(), k[want])
    else:
        assert skeys is None


@pytest.fixture
def force_hybrid():
    """Let the sort consider its HYBRID mode at any size (default: >= 2^22 rows)."""
    from ray_project_ray_b200 import _lib

    _lib.call("b200x_set_tuning", 0, 1)
    yield
    _lib.call("b200x_set_tuning", 0, 1 << 22)


@pytest.mark.parametrize("kind", ["i64_full", "i64_small", "i64_perm", "f64", "f64_ties", "i32", "f32", "i64_dups", "i64_heavy",
                                  "i64_lowbits", "i64_window_ties", "i64_cluster", "i64_straddle"])
@pytest.mark.parametrize("n", [2, 33, 1000, 70_000, 1_500_000])
@pytest.mark.parametrize("descending", [False, True])
def test_argsort_hybrid_mode_matches_oracle(cuda_device, force_hybrid, kind, n, descending):
    """Same contract with HYBRID forced on: spread keys finish in shared memory, duplicate-heavy or
    skewed inputs must fall back (runs too long / many equal windows) without changing the result."""
    rng = np.random.default_rng(99 + n)
    if kind == "i64_dups":        # few distinct wide keys: long runs of equal keys inside a 16-bit bucket
        k = rng.integers(np.iinfo(np.int64).min, np.iinfo(np.int64).max, 50, dtype=np.int64)[rng.integers(0, 50, n)]
    elif kind == "i64_heavy":     # one heavy hitter + wide tail
        k = rng.integers(np.iinfo(np.int64).min, np.iinfo(np.int64).max, n, dtype=np.int64)
        k[rng.random(n) < 0.6] = 123456789
    elif kind == "i64_lowbits":   # top 40 bits from a small set, low 24 bits random: ties on the 24-bit window
        k = (rng.integers(0, 300, n, dtype=np.int64) << 40) | rng.integers(0, 1 << 24, n, dtype=np.int64)
    elif kind == "i64_cluster":   # narrow range far from zero (one rank's slice of a distributed sort)
        k = np.int64(5_000_000_000_000_000_123) + rng.integers(0, 1 << 45, n, dtype=np.int64)
    elif kind == "i64_straddle":  # range that crosses a high-bit boundary of the code space
        k = rng.integers(-(1 << 40), 1 << 40, n, dtype=np.int64)
    elif kind == "i64_window_ties":  # equal down to bit 8, differing only in the lowest byte
        k = (rng.integers(0, 1 << 30, max(n // 6, 1), dtype=np.int64)[rng.integers(0, max(n // 6, 1), n)] << 8) | rng.integers(0, 256, n, dtype=np.int64)
    else:
        k = _keys(kind, n, rng)
    want = orc.stable_argsort(k, descending)
    t = torch.from_numpy(k).to(cuda_device)
    perm, codes, skeys = ops.argsort(t, descending, want_codes=True, want_keys=True)
    np.testing.assert_array_equal(_u32(perm), want.astype(np.uint32))
    np.testing.assert_array_equal(codes.cpu().numpy().view(np.uint64), orc.encode_keys(k, descending)[want])
    if k.dtype.kind in "iu":
        np.testing.assert_array_equal(skeys.cpu().numpy(), k[want])


def test_argsort_hybrid_default_threshold_large(cuda_device):
    """Above the default threshold (2^22 rows) spread keys take HYBRID; check against numpy."""
    rng = np.random.default_rng(7)
    n = 6_000_000
    for k in (rng.integers(np.iinfo(np.int64).min, np.iinfo(np.int64).max, n, dtype=np.int64), rng.standard_normal(n),
              rng.permutation(n).astype(np.int64)):
        perm, _, _ = ops.argsort(torch.from_numpy(k).to(cuda_device), False)
        np.testing.assert_array_equal(_u32(perm), np.argsort(k, kind="stable").astype(np.uint32))


@pytest.mark.parametrize("descending", [False, True])
@pytest.mark.parametrize("kind", ["f64_normal", "f64_lognormal", "f64_mixture_tiny", "f32_normal", "i64_gauss_narrow",
                                  "i64_gauss_wide", "i64_two_clusters", "f64_with_specials"])
def test_argsort_table_binned_hybrid(cuda_device, force_hybrid, kind, descending):
    """Keys from smooth but far-from-uniform distributions: the uniform 65536-way cut overflows, the
    sampled bucket table takes over (HYBRID, table=True). Sparse coarse buckets whose low bits do not
    fit the rewritten key (doubles: tiny magnitudes) are re-read from the column by the finisher."""
    rng = np.random.default_rng(4321)
    n = 3_000_000
    if kind == "f64_normal":
        k = rng.standard_normal(n)
    elif kind == "f64_lognormal":
        k = np.exp(rng.standard_normal(n) * 3)
    elif kind == "f64_mixture_tiny":      # populated binades down to 1e-300: many single-bin (re-read) buckets
        k = rng.standard_normal(n) * 10.0 ** rng.integers(-300, 3, n)
    elif kind == "f32_normal":
        k = rng.standard_normal(n).astype(np.float32)
    elif kind == "i64_gauss_narrow":
        k = (rng.standard_normal(n) * 2**30).astype(np.int64) + 10**15
    elif kind == "i64_gauss_wide":
        k = (rng.standard_normal(n) * 2**59).astype(np.int64)
    elif kind == "i64_two_clusters":
        k = np.where(rng.random(n) < 0.5, rng.integers(0, 1 << 40, n), (1 << 62) + rng.integers(0, 1 << 50, n)).astype(np.int64)
    else:
        k = rng.standard_normal(n)
        k[rng.integers(0, n, 1000)] = np.nan
        k[rng.integers(0, n, 1000)] = -0.0
        k[rng.integers(0, n, 1000)] = 0.0
        k[rng.integers(0, n, 10)] = np.inf
        k[rng.integers(0, n, 10)] = -np.inf
    want = orc.stable_argsort(k, descending)
    dbg = {}
    t = torch.from_numpy(k).to(cuda_device)
    perm, codes, skeys = ops.argsort(t, descending, want_codes=True, want_keys=True, _debug=dbg)
    np.testing.assert_array_equal(_u32(perm), want.astype(np.uint32))
    np.testing.assert_array_equal(codes.cpu().numpy().view(np.uint64), orc.encode_keys(k, descending)[want])
    if k.dtype.kind in "iu":
        np.testing.assert_array_equal(skeys.cpu().numpy(), k[want])
    assert dbg["mode"] == "hybrid", dbg
    if kind in ("f64_normal", "f32_normal", "f64_with_specials"):
        assert dbg["table"], dbg  # (the other kinds may still fit the uniform cut at this size)
    # through a row indirection (second key of a multi-column sort)
    p0 = torch.from_numpy(rng.permutation(n).astype(np.uint32).view(np.int32)).to(cuda_device)
    perm2, _, _ = ops.argsort(t, descending, perm_in=p0)
    src = _u32(p0)
    np.testing.assert_array_equal(_u32(perm2), src[orc.stable_argsort(k[src], descending)])


def test_argsort_uint64_and_uint16(cuda_device):
    rng = np.random.default_rng(5)
    n = 200_000
    hi = rng.integers(0, 2**63, n, dtype=np.int64).astype(np.uint64) * 2 + rng.integers(0, 2, n).astype(np.uint64)
    hi[: n // 4] = hi[n // 4 : n // 2]  # ties on the major key
    lo = rng.integers(0, 2**16, n, dtype=np.int64).astype(np.uint16)
    want = orc.stable_argsort_multi([hi, lo], False)
    thi = torch.from_numpy(hi.view(np.int64)).to(cuda_device).view(torch.uint64)
    tlo = torch.from_numpy(lo.view(np.int16)).to(cuda_device).view(torch.uint16)
    p1, _, _ = ops.argsort(tlo, False)
    p2, _, _ = ops.argsort(thi, False, perm_in=p1)
    np.testing.assert_array_equal(_u32(p2), want.astype(np.uint32))
    # descending multi-column: both columns descending, ties keep input order
    want = orc.stable_argsort_multi([hi, lo], True)
    p1, _, _ = ops.argsort(tlo, True)
    p2, _, _ = ops.argsort(thi, True, perm_in=p1)
    np.testing.assert_array_equal(_u32(p2), want.astype(np.uint32))


@pytest.mark.parametrize("shape,dtype", [((1000,), np.int64), ((1000,), np.float64), ((5000, 8), np.int64),
                                         ((3000, 64), np.uint8), ((3000, 90), np.uint8), ((3000, 100), np.uint8),
                                         ((777, 3), np.int16), ((1000,), np.int32), ((1000,), np.uint8), ((0, 8), np.int64),
                                         ((100_003, 90), np.uint8), ((70_001, 25), np.uint8), ((40_000, 82), np.uint8),
                                         ((33_333, 31), np.uint8), ((5_001, 1000), np.uint8), ((9_999, 10), np.int16)])
def test_gather_rows(cuda_device, shape, dtype):
    rng = np.random.default_rng(7)
    src = rng.integers(0, 120, shape).astype(dtype)
    n = shape[0]
    idx = rng.integers(0, max(n, 1), 2 * n).astype(np.int32) if n else np.zeros(0, np.int32)
    out = ops.gather(torch.from_numpy(src).to(cuda_device), torch.from_numpy(idx).to(cuda_device))
    np.testing.assert_array_equal(out.cpu().numpy(), src[idx])


def test_gather_columns(cuda_device):
    rng = np.random.default_rng(8)
    n = 50_000
    cols = [rng.integers(0, 2**62, n, dtype=np.int64) for _ in range(9)] + [rng.standard_normal(n)]
    pay = rng.integers(0, 255, (n, 24)).astype(np.uint8)
    idx = rng.permutation(n).astype(np.int32)
    outs = ops.gather_columns([torch.from_numpy(c).to(cuda_device) for c in cols] + [torch.from_numpy(pay).to(cuda_device)],
                              torch.from_numpy(idx).to(cuda_device))
    for o, c in zip(outs, cols + [pay]):
        np.testing.assert_array_equal(o.cpu().numpy(), c[idx])


@pytest.mark.parametrize("n", [1, 2, 3, 17, 1000, 65_536, 65_537, 1_000_003])
@pytest.mark.parametrize("seed", [0, 1, 2**40 + 3])
def test_permutation_is_bijection_and_matches_oracle(cuda_device, n, seed):
    fwd = _u32(ops.perm_indices(n, seed, cuda_device))
    assert np.array_equal(np.sort(fwd), np.arange(n, dtype=np.uint32))
    np.testing.assert_array_equal(fwd, orc.feistel_permutation(n, seed))
    np.testing.assert_array_equal(fwd, ops.perm_indices_host(n, seed))
    inv = _u32(ops.perm_indices(n, seed, cuda_device, inverse=True))
    np.testing.assert_array_equal(inv[fwd], np.arange(n, dtype=np.uint32))
    if n > 1000:
        other = _u32(ops.perm_indices(n, seed + 1, cuda_device))
        assert (other != fwd).mean() > 0.9
        assert (fwd != np.arange(n)).mean() > 0.9


@pytest.mark.parametrize("shape,dtype", [((10_000,), np.int64), ((10_000, 64), np.uint8), ((4_097, 72), np.uint8),
                                         ((5_000, 100), np.uint8), ((3, 8), np.int64)])
def test_permute_rows(cuda_device, shape, dtype):
    rng = np.random.default_rng(9)
    src = rng.integers(0, 120, shape).astype(dtype)
    out = ops.permute_rows(torch.from_numpy(src).to(cuda_device), seed=11)
    p = orc.feistel_permutation(shape[0], 11)
    np.testing.assert_array_equal(out.cpu().numpy(), src[p])
    back = ops.permute_rows(out, seed=11, inverse=True)
    # out[j] = src[pi(j)]; back[j] = out[pi^-1(j)] = src[pi(pi^-1(j))] = src[j]
    np.testing.assert_array_equal(back.cpu().numpy(), src)


@pytest.mark.parametrize("n,ngroups", [(1, 1), (10, 3), (2048, 1), (2049, 2049), (100_000, 7), (100_000, 50_000), (300_000, 1000)])
@pytest.mark.parametrize("vkind", ["i64", "f64"])
def test_groupby_kernels(cuda_device, n, ngroups, vkind):
    rng = np.random.default_rng(n + ngroups)
    keys = np.sort(rng.integers(0, ngroups, n, dtype=np.int64))
    vals = rng.integers(-2**40, 2**40, n, dtype=np.int64) if vkind == "i64" else rng.standard_normal(n) * 1e3
    perm = rng.permutation(n).astype(np.int32)
    vals_unsorted = np.empty_like(vals)
    vals_unsorted[perm] = vals  # values[perm[i]] == vals[i]
    codes = torch.from_numpy(orc.encode_keys(keys, False).view(np.int64)).to(cuda_device)
    starts, ng = ops.group_boundaries(codes)
    G = int(ng.item())
    uk, ustart, ucount = np.unique(keys, return_index=True, return_counts=True)
    assert G == len(uk)
    np.testing.assert_array_equal(starts[:G].cpu().numpy(), ustart.astype(np.int32))
    s, mn, mx = ops.segmented_reduce(starts, ng, G, torch.from_numpy(perm).to(cuda_device),
                                     torch.from_numpy(vals_unsorted).to(cuda_device), n, True, True, True)
    cnt = ops.group_counts(starts, ng, G, n)
    np.testing.assert_array_equal(cnt.cpu().numpy(), ucount)
    ws = np.add.reduceat(vals, ustart)
    wmn = np.minimum.reduceat(vals, ustart)
    wmx = np.maximum.reduceat(vals, ustart)
    if vkind == "i64":
        np.testing.assert_array_equal(s.cpu().numpy(), ws)
    else:
        np.testing.assert_allclose(s.cpu().numpy(), ws, rtol=1e-9, atol=1e-6)
    np.testing.assert_array_equal(mn.cpu().numpy(), wmn)
    np.testing.assert_array_equal(mx.cpu().numpy(), wmx)
    mean = ops.finalize_groups(s, cnt, 0)
    np.testing.assert_allclose(mean.cpu().numpy(), ws / ucount, rtol=1e-9, atol=1e-9)
    # sum of squared differences from the group mean (Std)
    m2, _, _ = ops.segmented_reduce(starts, ng, G, torch.from_numpy(perm).to(cuda_device),
                                    torch.from_numpy(vals_unsorted).to(cuda_device), n, True, False, False, center=mean)
    want_m2 = np.add.reduceat((vals.astype(np.float64) - np.repeat(ws / ucount, ucount)) ** 2, ustart)
    np.testing.assert_allclose(m2.cpu().numpy(), want_m2, rtol=1e-6, atol=1e-3)
    std = ops.finalize_groups(m2, cnt, 1, 1)
    want_std = np.where(ucount < 2, 0.0, np.sqrt(want_m2 / np.maximum(ucount - 1, 1)))
    np.testing.assert_allclose(std.cpu().numpy(), want_std, rtol=1e-6, atol=1e-6)
    # second run is bit-identical (fixed reduction order)
    s2, _, _ = ops.segmented_reduce(starts, ng, G, torch.from_numpy(perm).to(cuda_device),
                                    torch.from_numpy(vals_unsorted).to(cuda_device), n, True, False, False)
    assert torch.equal(s, s2)


def test_search_sorted_codes(cuda_device):
    rng = np.random.default_rng(3)
    keys = np.sort(rng.integers(0, 1000, 20_000, dtype=np.int64))
    codes = orc.encode_keys(keys, False)
    bounds = np.array([-5, 0, 1, 500, 999, 1000, 5000], dtype=np.int64)
    bc = orc.encode_keys(bounds, False)
    tc = torch.from_numpy(codes.view(np.int64)).to(cuda_device)
    tb = torch.from_numpy(bc.view(np.int64)).to(cuda_device)
    lo = ops.search_sorted_codes(tc, tb, upper=False).cpu().numpy()
    hi = ops.search_sorted_codes(tc, tb, upper=True).cpu().numpy()
    np.testing.assert_array_equal(lo, np.searchsorted(keys, bounds, side="left"))
    np.testing.assert_array_equal(hi, np.searchsorted(keys, bounds, side="right"))
    np.testing.assert_array_equal(ops.encode_keys(torch.from_numpy(keys).to(cuda_device)).cpu().numpy().view(np.uint64), codes)


@pytest.mark.parametrize("n,world", [(1, 2), (1000, 2), (2048, 4), (2049, 8), (300_000, 8), (1_000_003, 16)])
@pytest.mark.parametrize("descending", [False, True])
def test_partition_targets_and_scatter(cuda_device, n, world, descending):
    """Stable W-way split by key range + scatter into per-owner buffers (the 'peers' are local buffers here)."""
    rng = np.random.default_rng(n + world)
    keys = rng.integers(-1000, 1000, n, dtype=np.int64)
    codes = orc.encode_keys(keys, descending)
    split = np.sort(orc.encode_keys(rng.integers(-900, 900, world - 1, dtype=np.int64), descending))
    owner = np.searchsorted(split, codes, side="right")          # number of splitters <= code
    t_keys = torch.from_numpy(keys).to(cuda_device)
    t_split = torch.from_numpy(split.view(np.int64)).to(cuda_device)
    target, counts = ops.partition_targets(t_keys, descending, t_split)
    tg = target.cpu().numpy().view(np.uint32)
    np.testing.assert_array_equal(tg >> 28, owner)
    want_counts = np.bincount(owner, minlength=16)
    np.testing.assert_array_equal(counts.cpu().numpy(), want_counts)
    for d in range(world):  # slots follow row order inside every owner
        np.testing.assert_array_equal(tg[owner == d] & ((1 << 28) - 1), np.arange(want_counts[d]))
    pay = rng.integers(0, 256, (n, 24), dtype=np.uint8)
    t_pay = torch.from_numpy(pay).to(cuda_device)
    bufs = [torch.zeros((int(want_counts[d]) + 3, 24), dtype=torch.uint8, device=cuda_device) for d in range(world)]
    ops.scatter_rows_by_target(t_pay, target, [b.data_ptr() for b in bufs], [3] * world)
    kbufs = [torch.zeros(int(want_counts[d]) + 3, dtype=torch.int64, device=cuda_device) for d in range(world)]
    ops.scatter_rows_by_target(t_keys, target, [b.data_ptr() for b in kbufs], [3] * world)
    for d in range(world):
        np.testing.assert_array_equal(bufs[d].cpu().numpy()[3:], pay[owner == d])
        np.testing.assert_array_equal(kbufs[d].cpu().numpy()[3:], keys[owner == d])
    # row widths that are not a multiple of 16 bytes (realigned in shared memory, aligned 16-byte stores):
    # exact contents, and nothing written before the first or after the last row of any owner
    for w, off in ((90, 3), (82, 1), (30, 0), (100, 5), (17, 2), (25, 7)):
        pw = rng.integers(1, 256, (n, w), dtype=np.uint8)
        t_pw = torch.from_numpy(pw).to(cuda_device)
        wb = [torch.zeros((int(want_counts[d]) + off + 2, w), dtype=torch.uint8, device=cuda_device) for d in range(world)]
        ops.scatter_rows_by_target(t_pw, target, [b.data_ptr() for b in wb], [off] * world)
        for d in range(world):
            got = wb[d].cpu().numpy()
            np.testing.assert_array_equal(got[off:off + int(want_counts[d])], pw[owner == d])
            assert not got[:off].any() and not got[off + int(want_counts[d]):].any()
