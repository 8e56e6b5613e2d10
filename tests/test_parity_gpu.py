"""GPU parity tests of the whole path (Dataset API -> planner -> scheduler -> C ABI kernels)
against the reference outputs in tests/golden/ and the CPU oracle. Sort order, block contents,
integer aggregates and counts are bit-exact; float sums / means / stds within 1e-6 relative
(the tolerance BASELINE.json's north_star states)."""
import numpy as np
import pyarrow as pa
import pyarrow.compute as pac
import pytest
import torch

import ray_project_ray_b200.data as rd
from oracle import exchange_oracle as orc
from ray_project_ray_b200.aggregate import Count, Max, Mean, Min, Std, Sum
from ray_project_ray_b200.aggregate_task_spec import SortAggregateTaskSpec
from ray_project_ray_b200.block import BlockAccessor, DeviceBlock
from ray_project_ray_b200.context import DataContext
from ray_project_ray_b200.interfaces import RefBundle, TaskContext
from ray_project_ray_b200.schedulers import (GpuExchangeScheduler, PullBasedShuffleTaskScheduler,
                                             PushBasedShuffleTaskScheduler)
from ray_project_ray_b200.shuffle_task_spec import ShuffleTaskSpec
from ray_project_ray_b200.sort_task_spec import SortKey, SortTaskSpec
from tests import _golden as G

pytestmark = pytest.mark.gpu
FLOAT_RTOL = 1e-6

SCHEDULERS = {"fused": GpuExchangeScheduler, "pull": PullBasedShuffleTaskScheduler, "push": PushBasedShuffleTaskScheduler}


@pytest.fixture(autouse=True)
def _restore_context():
    ctx = DataContext.get_current()
    saved = (ctx.use_push_based_shuffle, ctx.use_fused_device_exchange)
    yield
    ctx.use_push_based_shuffle, ctx.use_fused_device_exchange = saved


def _set_scheduler(name):
    ctx = DataContext.get_current()
    ctx.use_fused_device_exchange = name == "fused"
    ctx.use_push_based_shuffle = name == "push"


def _bundles(blocks):
    dev = [DeviceBlock.from_numpy(b) if b else DeviceBlock.empty() for b in blocks]
    return [RefBundle([(b, BlockAccessor.for_block(b).get_metadata(None, None))], owns_blocks=True) for b in dev]


def _run(sched, spec, blocks, R):
    out, stats = SCHEDULERS[sched](spec).execute(_bundles(blocks), R, TaskContext(task_idx=0, target_max_block_size=1 << 30))
    return [b.blocks[0][0] for b in out], stats


# ------------------------------------------------------------------------------- sort
@pytest.mark.parametrize("sched", ["fused", "pull", "push"])
@pytest.mark.parametrize("case", [c["tag"] for c in G.meta()["sort_cases"] if not c["push"]])
def test_sort_blocks_equal_reference(case, sched, cuda_device):
    """Same input blocks + the boundaries the reference sampled -> identical output blocks."""
    info = next(c for c in G.meta()["sort_cases"] if c["tag"] == case)
    a = G.arrays()
    blocks = G.sort_case_blocks(info["case"])
    desc = info["descending"]
    bounds = [(b.item(),) for b in a[case + "_bounds"]]
    if desc:
        bounds.reverse()  # generate_sort_fn hands descending boundaries to the spec (planner/sort.py:52-54)
    outs, stats = _run(sched, SortTaskSpec(bounds, SortKey("k", descending=desc)), blocks, len(bounds) + 1)
    want = G.split_by_sizes(a[case + "_rows"], a[case + "_sizes"])
    assert [b.num_rows for b in outs] == [len(w) for w in want]
    key = a[f"sort_{info['case']}_key"]
    for got, w in zip(outs, want):
        if len(w):
            host = got.to_numpy()
            np.testing.assert_array_equal(host["row"], w)
            np.testing.assert_array_equal(host["k"].view(np.int64), key[w].view(np.int64))  # bit-exact incl. -0.0
    assert sorted(stats.keys()) == (["map", "merge", "reduce"] if sched == "push" else ["map", "reduce"])


@pytest.mark.parametrize("sched", ["fused", "pull", "push"])
@pytest.mark.parametrize("desc", [False, True])
def test_dataset_sort_nan_keys_global_order(desc, sched, cuda_device):
    _set_scheduler(sched)
    a = G.arrays()
    key = a["sort_f64_nan_key"]
    ds = rd.from_numpy_columns({"k": key, "row": np.arange(len(key), dtype=np.int64)}, override_num_blocks=8)
    out = ds.sort("k", descending=desc)
    assert out.num_blocks() == 8
    host = out.to_numpy()
    np.testing.assert_array_equal(host["row"], a[f"sort_f64_nan_{'desc' if desc else 'asc'}_rows"])
    np.testing.assert_array_equal(host["k"].view(np.int64), key[host["row"]].view(np.int64))


@pytest.mark.parametrize("sched", ["fused", "pull"])
@pytest.mark.parametrize("desc", [False, True])
def test_dataset_sort_two_columns(desc, sched, cuda_device):
    _set_scheduler(sched)
    a = G.arrays()
    n = len(a["sort2_k1"])
    ds = rd.from_numpy_columns({"a": a["sort2_k1"], "b": a["sort2_k2"], "row": np.arange(n, dtype=np.int64)},
                               override_num_blocks=8)
    out = ds.sort(["a", "b"], descending=desc)
    np.testing.assert_array_equal(out.to_numpy()["row"], a[f"sort2_{'desc' if desc else 'asc'}_rows"])


@pytest.mark.parametrize("sched", ["fused", "pull", "push"])
def test_sort_with_specified_boundaries(sched, cuda_device):
    """python/ray/data/tests/test_sort.py:23-51."""
    _set_scheduler(sched)
    for case in G.meta()["specified_boundaries"]:
        ds = rd.range(1000, override_num_blocks=20).sort("id", case["descending"], case["boundaries"])
        blocks = [list(b["id"]) if b else [] for b in ds.iter_batches(batch_size=None)]
        assert blocks == case["blocks"]


@pytest.mark.parametrize("sched", ["fused", "pull", "push"])
def test_sort_simple_and_empty(sched, cuda_device):
    """python/ray/data/tests/test_sort.py:54-76."""
    _set_scheduler(sched)
    rng = np.random.default_rng(0)
    xs = rng.permutation(100).tolist()
    ds = rd.from_items(xs, override_num_blocks=4)
    assert [r["item"] for r in ds.sort("item").take(100)] == list(range(100))
    assert len([n for n in ds.sort("item")._block_num_rows() if n > 0]) == 4
    assert [r["item"] for r in ds.sort("item", descending=True).take(100)] == list(reversed(range(100)))
    empty = rd.from_items([])
    assert empty.sort("item").count() == 0 and empty.sort("item").take() == []
    assert rd.range(10).filter(lambda r: r["id"] > 10).sort("id").count() == 0


@pytest.mark.parametrize("sched", ["fused", "pull", "push"])
def test_sort_partition_same_key_to_same_block(sched, cuda_device):
    """python/ray/data/tests/test_sort.py:79-94."""
    _set_scheduler(sched)
    ds = rd.from_items([1] * 100).repartition(100).sort("item")
    sizes = ds._block_num_rows()
    assert len(sizes) == 100 and sum(1 for s in sizes if s > 0) == 1 and sum(sizes) == 100


def test_sort_empty_blocks_block_level(cuda_device):
    """python/ray/data/tests/test_sort.py:142-191."""
    acc = BlockAccessor.for_block(pa.Table.from_pydict({}))
    assert acc.sample(10, SortKey("A")).num_rows == 0
    parts = acc.sort_and_partition([1, 5, 10], SortKey("A"))
    assert len(parts) == 4 and all(p.num_rows == 0 for p in parts)
    assert BlockAccessor.for_block(pa.Table.from_pydict({})).merge_sorted_blocks([pa.Table.from_pydict({})], SortKey("A"))[0].num_rows == 0
    ds = rd.from_items([{"A": x % 3, "B": x} for x in range(3)], override_num_blocks=3).filter(lambda r: r["A"] == 0)
    assert list(ds.sort("A").iter_rows()) == [{"A": 0, "B": 0}]
    assert len(SortTaskSpec.sample_boundaries(rd.range(10).filter(lambda r: r["id"] > 10).get_internal_block_refs(),
                                              SortKey("id"), 3)) == 2


@pytest.mark.parametrize("sched", ["fused", "pull"])
def test_config1_shape_sort_desc_matches_pyarrow(sched, cuda_device):
    """BASELINE configs[0] at reduced size: range(n) ids permuted, 8 int64 payload columns,
    sort('id', descending=True); the whole table must equal take(sort_indices)."""
    _set_scheduler(sched)
    n = 200_000
    rng = np.random.default_rng(1)
    ids = rng.integers(0, n // 4, n, dtype=np.int64)  # ties: tie order is part of the contract
    cols = {"id": ids}
    for j in range(8):
        cols[f"p{j}"] = rng.integers(0, 2**62, n, dtype=np.int64)
    out = rd.from_numpy_columns(cols, override_num_blocks=8).sort("id", descending=True)
    assert out.num_blocks() == 8
    got = out.to_numpy()
    table = pa.table(cols)
    want = table.take(pac.sort_indices(table, [("id", "descending")]))
    for c in cols:
        np.testing.assert_array_equal(got[c], want[c].to_numpy())


def test_sort_packed_payload_and_narrow_keys(cuda_device):
    """gensort-shaped rows: (hi uint64, lo uint16) key + 90-byte packed payload (configs[3])."""
    n = 50_000
    rng = np.random.default_rng(2)
    hi = rng.integers(0, 2**63, n, dtype=np.int64).astype(np.uint64) * 2 + 1
    hi[: n // 3] = hi[n // 3: 2 * (n // 3)]
    lo = rng.integers(0, 2**16, n).astype(np.uint16)
    pay = rng.integers(0, 256, (n, 90)).astype(np.uint8)
    out = rd.from_numpy_columns({"hi": hi, "lo": lo, "payload": pay}, override_num_blocks=4).sort(["hi", "lo"]).to_numpy()
    order = orc.stable_argsort_multi([hi, lo], False)
    np.testing.assert_array_equal(out["hi"], hi[order])
    np.testing.assert_array_equal(out["lo"], lo[order])
    np.testing.assert_array_equal(out["payload"], pay[order])


# ---------------------------------------------------------------------------- groupby
AGGS = lambda: [Sum("v"), Mean("f"), Count(), Min("v"), Max("f"), Std("f"), Mean("v"), Sum("f")]  # noqa: E731


@pytest.mark.parametrize("sched", ["fused", "pull", "push"])
def test_groupby_equals_reference(sched, cuda_device):
    a, m = G.arrays(), G.meta()
    B, n = m["sort_shape"]["blocks"], m["sort_shape"]["rows_per_block"]
    _set_scheduler(sched)
    ds = rd.from_numpy_columns({"k": a["agg_k"], "v": a["agg_v"], "f": a["agg_f"],
                                "unused": np.arange(B * n, dtype=np.int64)}, override_num_blocks=B)
    out = ds.groupby("k").aggregate(*AGGS())
    assert out.num_blocks() == B
    got = out.to_numpy()
    tag = "agg_pull"
    assert list(got.keys()) == m[tag + "_columns"]
    for c in ("k", "sum(v)", "count()", "min(v)"):
        np.testing.assert_array_equal(got[c], a[f"{tag}_{c}"])
    np.testing.assert_array_equal(got["max(f)"], a[f"{tag}_max(f)"])
    for c in ("mean(f)", "std(f)", "mean(v)", "sum(f)"):
        np.testing.assert_allclose(got[c], a[f"{tag}_{c}"], rtol=FLOAT_RTOL)


@pytest.mark.parametrize("sched", ["fused", "pull", "push"])
def test_small_groupby_goldens(sched, cuda_device):
    """python/ray/data/tests/test_all_to_all.py:187-317, :508-580: A = x % 3, B = x."""
    _set_scheduler(sched)
    m = G.meta()
    xs = m["small_groupby_input"]
    ds = rd.from_items([{"A": x % 3, "B": x} for x in xs], override_num_blocks=4)
    out = ds.groupby("A").aggregate(Sum("B"), Mean("B"), Count(), Min("B"), Max("B"), Std("B")).to_numpy()
    assert list(out["A"]) == [0, 1, 2]
    assert list(out["sum(B)"]) == [1683, 1617, 1650]
    assert list(out["mean(B)"]) == [49.5, 49.0, 50.0]
    assert list(out["count()"]) == [34, 33, 33]
    assert list(out["min(B)"]) == m["small_groupby"]["min(B)"] and list(out["max(B)"]) == m["small_groupby"]["max(B)"]
    np.testing.assert_allclose(out["std(B)"], m["small_groupby"]["std(B)"], rtol=FLOAT_RTOL)
    assert [r["count()"] for r in ds.groupby("A").count().take_all()] == [34, 33, 33]
    assert [r["sum(B)"] for r in ds.groupby("A").sum("B").take_all()] == [1683, 1617, 1650]
    # global aggregation (key=None) and the Dataset shortcuts
    g = m["small_global_agg"]
    assert ds.sum("B") == g["sum(B)"][0] == 4950 and ds.min("B") == 0 and ds.max("B") == 99
    assert ds.mean("B") == g["mean(B)"][0] and abs(ds.std("B") - g["std(B)"][0]) < 1e-9
    assert ds.aggregate(Sum("B"), Count()) == {"sum(B)": 4950, "count()": 100}
    # duplicate names
    cols = ds.groupby("A").aggregate(Sum("B"), Sum("B"), Count(), Count()).columns()
    assert cols == m["dup_names_columns"]
    assert sorted(ds.unique("A")) == [0, 1, 2]


def test_groupby_skewed_and_single_group(cuda_device):
    """Zipf-like skew: one key holds most rows (spans many reduction tiles)."""
    rng = np.random.default_rng(3)
    n = 400_000
    k = np.where(rng.random(n) < 0.7, 7, rng.integers(0, 5000, n)).astype(np.int64)
    v = rng.integers(-2**31, 2**31, n, dtype=np.int64)
    f = rng.random(n)
    out = rd.from_numpy_columns({"k": k, "v": v, "f": f}, override_num_blocks=8).groupby("k").aggregate(
        Sum("v"), Mean("f"), Count(), Std("f")).to_numpy()
    want = pa.table({"k": k, "v": v, "f": f}).group_by("k").aggregate(
        [("v", "sum"), ("f", "mean"), ("v", "count"), ("f", "stddev", pac.VarianceOptions(ddof=1))]).sort_by("k")
    np.testing.assert_array_equal(out["k"], want["k"].to_numpy())
    np.testing.assert_array_equal(out["sum(v)"], want["v_sum"].to_numpy())
    np.testing.assert_array_equal(out["count()"], want["v_count"].to_numpy())
    np.testing.assert_allclose(out["mean(f)"], want["f_mean"].to_numpy(), rtol=FLOAT_RTOL)
    std_want = np.nan_to_num(want["f_stddev"].to_numpy(zero_copy_only=False).astype(np.float64), nan=0.0)
    np.testing.assert_allclose(out["std(f)"], std_want, rtol=FLOAT_RTOL, atol=1e-12)


def test_aggregate_task_spec_map_reduce_partials(cuda_device):
    """SortAggregateTaskSpec.map / reduce with partial_reduce, as the push-based merge stage uses them."""
    rng = np.random.default_rng(4)
    blocks = [{"k": rng.integers(0, 50, 3000, dtype=np.int64), "f": rng.random(3000)} for _ in range(3)]
    aggs = [Sum("f"), Mean("f"), Count(), Std("f"), Min("f")]
    bounds = [(10,), (30,)]
    maps = [SortAggregateTaskSpec.map(i, DeviceBlock.from_numpy(b), 3, bounds, "k", aggs)[:-1] for i, b in enumerate(blocks)]
    finals = []
    for j in range(3):
        partial, _ = SortAggregateTaskSpec.reduce("k", aggs, *[m[j] for m in maps[:2]], partial_reduce=True)
        final, meta = SortAggregateTaskSpec.reduce("k", aggs, partial, maps[2][j])
        finals.append(final)
        assert meta.num_rows == final.num_rows
    got = DeviceBlock.concat(finals).to_numpy()
    want = orc.concat(orc.aggregate_dataset(blocks, "k", [("sum", "f"), ("mean", "f"), ("count", None), ("std", "f"), ("min", "f")], bounds))
    np.testing.assert_array_equal(got["k"], want["k"])
    np.testing.assert_array_equal(got["count()"], want["count()"])
    np.testing.assert_array_equal(got["min(f)"], want["min(f)"])
    for c in ("sum(f)", "mean(f)", "std(f)"):
        np.testing.assert_allclose(got[c], np.asarray(want[c], dtype=np.float64), rtol=FLOAT_RTOL)


# --------------------------------------------------------------------- random_shuffle
@pytest.mark.parametrize("sched", ["fused", "pull", "push"])
def test_random_shuffle_properties(sched, cuda_device):
    """python/ray/data/tests/test_all_to_all.py:1050-1142."""
    _set_scheduler(sched)
    B, n = 8, 2500
    ds = rd.range(B * n, override_num_blocks=B)
    r0, r0b, r1 = (ds.random_shuffle(seed=s) for s in (0, 0, 1))
    a, b, c = (np.concatenate([blk["id"] for blk in r.iter_batches(batch_size=None)]) for r in (r0, r0b, r1))
    np.testing.assert_array_equal(a, b)                       # same seed, same output
    assert (a != c).mean() > 0.9                              # different seed
    np.testing.assert_array_equal(np.sort(a), np.arange(B * n))  # exact multiset permutation
    assert (a != np.arange(B * n)).mean() > 0.9               # input order not preserved
    assert r0.num_blocks() == B and sum(r0._block_num_rows()) == B * n
    for blk in r0.iter_batches(batch_size=None):              # every output draws from every input
        assert len(set(blk["id"] // n)) == B
        runs = np.diff(blk["id"]) == 1
        assert runs.sum() < 0.01 * len(runs) + 10             # no long contiguous runs
    assert ds.random_shuffle().count() == B * n               # unseeded works
    assert rd.from_items([]).random_shuffle(seed=0).count() == 0
    with pytest.raises(DeprecationWarning):
        ds.random_shuffle(seed=0, num_blocks=3)


def test_random_shuffle_pull_equals_oracle_restatement(cuda_device):
    """Task-structured shuffle (per-block permutation, slices, shuffled slice order, per-reducer
    permutation) is restated in the oracle with the same Feistel stream: exact equality."""
    B, n = 4, 1003
    row = np.arange(B * n, dtype=np.int64)
    pay = np.random.default_rng(0).integers(0, 256, (B * n, 24)).astype(np.uint8)
    blocks = [{"row": row[i * n:(i + 1) * n], "pay": pay[i * n:(i + 1) * n]} for i in range(B)]
    outs, _ = _run("pull", ShuffleTaskSpec(1 << 30, random_shuffle=True, random_seed=5), blocks, B)
    want = orc.shuffle_dataset(blocks, 5)
    for got, w in zip(outs, want):
        host = got.to_numpy()
        np.testing.assert_array_equal(host["row"], w["row"])
        np.testing.assert_array_equal(host["pay"], w["pay"])


def test_random_shuffle_fused_is_one_global_bijection(cuda_device):
    n = 100_000
    pay = np.random.default_rng(1).integers(0, 256, (n, 64)).astype(np.uint8)
    ds = rd.from_numpy_columns({"id": np.arange(n, dtype=np.int64), "payload": pay}, override_num_blocks=8)
    out = ds.random_shuffle(seed=0).to_numpy()
    p = orc.feistel_permutation(n, 0)
    np.testing.assert_array_equal(out["id"], p.astype(np.int64))
    np.testing.assert_array_equal(out["payload"], pay[p])


def test_repartition(cuda_device):
    """Split repartition keeps row order; shuffle repartition gives output j the j-th ceil(n/R)-row
    slice of every input block (ShuffleTaskSpec without random_shuffle, shuffle_task_spec.py:86-93)."""
    ds = rd.range(1000, override_num_blocks=7).repartition(3)
    assert ds.num_blocks() == 3 and ds.count() == 1000
    np.testing.assert_array_equal(ds.to_numpy()["id"], np.arange(1000))
    src = rd.range(1000, override_num_blocks=7)
    out = src.repartition(3, shuffle=True)
    assert out.num_blocks() == 3 and out.count() == 1000
    ins = [b["id"] for b in src.iter_batches(batch_size=None)]
    for j, blk in enumerate(out.iter_batches(batch_size=None)):
        want = np.concatenate([x[j * -(-len(x) // 3):(j + 1) * -(-len(x) // 3)] for x in ins])
        np.testing.assert_array_equal(blk["id"], want)


def test_product_fails_loudly_without_cuda_tensors():
    from ray_project_ray_b200 import ops

    with pytest.raises(RuntimeError, match="no CPU fallback"):
        ops.argsort(torch.zeros(4, dtype=torch.int64))


@pytest.mark.parametrize("sched", ["fused", "pull", "push"])
def test_groupby_two_key_columns_and_absmax(sched, cuda_device):
    """groupby(["a", "b"]) (dataset.py:1788-1836) and AbsMax (_aggregate.py:323-357) against pyarrow's group_by."""
    from ray_project_ray_b200.aggregate import AbsMax

    _set_scheduler(sched)
    rng = np.random.default_rng(12)
    n = 60_000
    a = rng.integers(0, 40, n, dtype=np.int64)
    b = rng.integers(-3, 3, n).astype(np.float64) / 2          # float second key with -0.0 / 0.0 ties
    b[rng.integers(0, n, 500)] = -0.0
    v = rng.integers(-10**9, 10**9, n, dtype=np.int64)
    f = rng.standard_normal(n)
    ds = rd.from_numpy_columns({"a": a, "b": b, "v": v, "f": f, "junk": np.arange(n)}, override_num_blocks=6)
    out = ds.groupby(["a", "b"]).aggregate(Sum("v"), Count(), Mean("f"), AbsMax("v"), AbsMax("f"), Max("f")).to_numpy()
    assert list(out.keys()) == ["a", "b", "sum(v)", "count()", "mean(f)", "abs_max(v)", "abs_max(f)", "max(f)"]
    t = pa.table({"a": a, "b": b + 0.0, "v": v, "f": f, "av": np.abs(v), "af": np.abs(f)})
    want = t.group_by(["a", "b"]).aggregate([("v", "sum"), ("v", "count"), ("f", "mean"), ("av", "max"), ("af", "max"),
                                             ("f", "max")]).sort_by([("a", "ascending"), ("b", "ascending")])
    np.testing.assert_array_equal(out["a"], want["a"].to_numpy())
    np.testing.assert_array_equal(out["b"] + 0.0, want["b"].to_numpy())
    np.testing.assert_array_equal(out["sum(v)"], want["v_sum"].to_numpy())
    np.testing.assert_array_equal(out["count()"], want["v_count"].to_numpy())
    np.testing.assert_array_equal(out["abs_max(v)"], want["av_max"].to_numpy())
    np.testing.assert_array_equal(out["abs_max(f)"], want["af_max"].to_numpy())
    np.testing.assert_array_equal(out["max(f)"], want["f_max"].to_numpy())
    np.testing.assert_allclose(out["mean(f)"], want["f_mean"].to_numpy(), rtol=FLOAT_RTOL)
    assert ds.aggregate(AbsMax("v")) == {"abs_max(v)": int(np.abs(v).max())}


@pytest.mark.parametrize("sched", ["fused", "pull", "push"])
def test_quantile_equals_reference_outputs(sched, cuda_device):
    """Quantile (aggregate/_aggregate.py:360-431) against what the reference's own code returned
    (tests/golden/quantile_outputs.json) and its test data (tests/test_all_to_all.py:720-775)."""
    from ray_project_ray_b200.aggregate import Quantile

    _set_scheduler(sched)
    q = G.quantile()
    k, vi, vf = np.array(q["k"], np.int64), np.array(q["vi"], np.int64), np.array(q["vf"], np.float64)
    ds = rd.from_numpy_columns({"k": k, "vi": vi, "vf": vf, "junk": np.arange(len(k))}, override_num_blocks=q["blocks"])
    aggs = [Quantile("vi", q=qq, alias_name=f"qi{j}") for j, qq in enumerate(q["qs"])]
    aggs += [Quantile("vf", q=qq, alias_name=f"qf{j}") for j, qq in enumerate(q["qs"])]
    out = ds.groupby("k").aggregate(*aggs, Count(), Sum("vi"), Quantile("vf")).to_numpy()
    ref = q["grouped"]["pull"]
    assert list(out.keys()) == list(ref.keys())
    np.testing.assert_array_equal(out["k"], ref["k"])
    np.testing.assert_array_equal(out["count()"], ref["count()"])
    np.testing.assert_array_equal(out["sum(vi)"], ref["sum(vi)"])
    for name in ref:
        if name.startswith("q"):
            assert out[name].dtype == np.float64
            np.testing.assert_allclose(out[name], np.array(ref[name], np.float64), rtol=1e-12, atol=0)
    glob = ds.aggregate(*aggs)
    for name, v in q["global"].items():
        assert abs(glob[name] - v[0]) <= 1e-12 * abs(v[0])
    xs = q["small_input"]
    small = rd.from_items([{"A": x % 3, "B": x} for x in xs], override_num_blocks=4)
    got = small.groupby("A").aggregate(Count(), Quantile("B"), Quantile("B", q=0.9, alias_name="q9")).to_numpy()
    assert list(got["quantile(B)"]) == [49.5, 49.0, 50.0] and list(got["q9"]) == q["small_grouped"]["q9"]
    assert small.aggregate(Quantile("B")) == {"quantile(B)": 49.5}
    # one row per group, and a group spanning many blocks
    one = rd.from_numpy_columns({"k": np.arange(1000), "v": np.arange(1000) * 1.5}, override_num_blocks=3)
    np.testing.assert_array_equal(one.groupby("k").aggregate(Quantile("v", q=0.3)).to_numpy()["quantile(v)"],
                                  np.arange(1000) * 1.5)
    rng = np.random.default_rng(5)
    v = rng.standard_normal(300_001)
    big = rd.from_numpy_columns({"k": np.zeros(v.size, np.int64), "v": v}, override_num_blocks=8)
    got = big.groupby("k").aggregate(Quantile("v", q=0.5), Quantile("v", q=0.123, alias_name="p")).to_numpy()
    assert got["quantile(v)"][0] == np.sort(v)[150_000]
    assert abs(got["p"][0] - orc.percentile(v.tolist(), 0.123)) <= 1e-12


@pytest.mark.parametrize("sched", ["fused", "pull", "push"])
@pytest.mark.parametrize("desc", [False, True])
def test_sort_by_byte_string_key(sched, desc, cuda_device):
    """Fixed-width byte-string keys (gensort's 10-byte key) sort in pyarrow's lexicographic
    order, stably, through every scheduler; the packed chunk columns do not appear in the output."""
    from ray_project_ray_b200 import ops

    _set_scheduler(sched)
    rng = np.random.default_rng(21)
    n = 40_000
    key = rng.integers(0, 256, (n, 10)).astype(np.uint8)
    key[:, :8] = key[rng.integers(0, 300, n), :8]           # many ties on the first 8 bytes: the last 2 decide
    key[rng.integers(0, n, 200)] = key[0]                    # exact duplicates: stability
    payload = rng.integers(0, 256, (n, 90)).astype(np.uint8)
    row = np.arange(n, dtype=np.int64)
    ds = rd.from_numpy_columns({"key": key, "payload": payload, "row": row}, override_num_blocks=5)
    out = ds.sort("key", descending=desc)
    assert out.columns() == ["key", "payload", "row"] and out.num_blocks() == 5
    got = out.to_numpy()
    want = orc.stable_argsort_bytes(key, desc)
    np.testing.assert_array_equal(got["row"], row[want])
    np.testing.assert_array_equal(got["key"], key[want])
    np.testing.assert_array_equal(got["payload"], payload[want])
    # mixed key: an integer column first, the byte string second
    a = rng.integers(0, 7, n, dtype=np.int64)
    ds2 = rd.from_numpy_columns({"a": a, "key": key[:, :3].copy(), "row": row}, override_num_blocks=3)
    got2 = ds2.sort(["a", "key"], descending=desc).to_numpy()
    k3 = key[:, :3]
    order = np.lexsort([(255 - k3[:, 2]) if desc else k3[:, 2], (255 - k3[:, 1]) if desc else k3[:, 1],
                        (255 - k3[:, 0]) if desc else k3[:, 0], -a if desc else a])
    np.testing.assert_array_equal(got2["row"], row[order])
    # the packing kernel itself
    t = torch.from_numpy(key).to(cuda_device)
    hi = ops.pack_key_bytes(t, 0, 8)
    lo = ops.pack_key_bytes(t, 8, 2)
    assert hi.dtype == torch.uint64 and lo.dtype == torch.uint16
    np.testing.assert_array_equal(hi.view(torch.int64).cpu().numpy().view(np.uint64),
                                  key[:, :8].copy().view(">u8").astype(np.uint64).ravel())
    np.testing.assert_array_equal(lo.view(torch.int16).cpu().numpy().view(np.uint16),
                                  (key[:, 8].astype(np.uint16) << 8) | key[:, 9].astype(np.uint16))
    np.testing.assert_array_equal(ops.unpack_key_bytes([hi, lo], 10).cpu().numpy(), key)
    k3 = torch.from_numpy(key[:, :3].copy()).to(cuda_device)
    c3 = ops.byte_key_chunks(k3)
    assert len(c3) == 1 and c3[0].dtype == torch.uint32
    np.testing.assert_array_equal(ops.unpack_key_bytes(c3, 3).cpu().numpy(), key[:, :3])
    with pytest.raises(ValueError):
        ds.sort("key", boundaries=[1, 2])


def test_iter_batches_across_blocks_and_local_shuffle(cuda_device):
    """Dataset.iter_batches / iter_torch_batches (dataset.py:3640-3800): batches of batch_size rows cut
    across block boundaries; with a local shuffle buffer every row still appears exactly once, the
    result is deterministic per seed, and rows leave the stream order (tests/test_iterator... local_shuffle)."""
    n = 10_000
    ds = rd.range(n, override_num_blocks=7)
    sizes = [len(b["id"]) for b in ds.iter_batches(batch_size=512)]
    assert sizes == [512] * (n // 512) + [n % 512]
    assert np.array_equal(np.concatenate([b["id"] for b in ds.iter_batches(batch_size=512)]), np.arange(n))
    assert sum(len(b["id"]) for b in ds.iter_batches(batch_size=512, drop_last=True)) == (n // 512) * 512
    tb = list(ds.iter_torch_batches(batch_size=3000))
    assert all(t["id"].is_cuda for t in tb) and [t["id"].numel() for t in tb] == [3000, 3000, 3000, 1000]

    def run(seed):
        return [b["id"] for b in ds.iter_batches(batch_size=256, local_shuffle_buffer_size=1000, local_shuffle_seed=seed)]

    a, b, c = run(5), run(5), run(6)
    flat = np.concatenate(a)
    assert [len(x) for x in a][:-1] == [256] * (len(a) - 1)
    assert np.array_equal(np.sort(flat), np.arange(n)) and not np.array_equal(flat, np.arange(n))
    assert all(np.array_equal(x, y) for x, y in zip(a, b))
    assert not np.array_equal(flat, np.concatenate(c))
    # a local buffer shuffles locally: early batches only hold rows from the first blocks
    assert a[0].max() < 4 * (n // 7 + 1)
    with pytest.raises(ValueError):
        next(iter(ds.iter_batches(batch_size=None, local_shuffle_buffer_size=10)))
    order = [int(b["id"][0]) for b in rd.range(90, override_num_blocks=9).randomize_block_order(seed=42).iter_batches(batch_size=None)]
    assert order == [10 * i for i in G.batcher()["randomize_block_order"]["42"]]


def test_multi_column_sort_first_column_distinct_shortcut(cuda_device):
    """Multi-column keys whose first column has no duplicates need only the sort by that column; one
    duplicate anywhere must send the sort down the full chain. Both against the oracle's lexicographic order."""
    rng = np.random.default_rng(31)
    n = 300_000
    a = rng.permutation(n).astype(np.int64) * 7 - 1000          # all distinct
    b = rng.integers(0, 5, n, dtype=np.int64)
    row = np.arange(n, dtype=np.int64)
    for desc in (False, True):
        got = rd.from_numpy_columns({"a": a, "b": b, "row": row}, override_num_blocks=4).sort(["a", "b"], descending=desc).to_numpy()
        np.testing.assert_array_equal(got["row"], row[orc.stable_argsort_multi([a, b], desc)])
        a2 = a.copy()
        a2[123_456] = a2[17]                                      # one tie: b decides it, the sample does not see it
        b2 = b.copy()
        b2[123_456], b2[17] = (0, 4)
        got = rd.from_numpy_columns({"a": a2, "b": b2, "row": row}, override_num_blocks=4).sort(["a", "b"], descending=desc).to_numpy()
        np.testing.assert_array_equal(got["row"], row[orc.stable_argsort_multi([a2, b2], desc)])
        f = rng.standard_normal(n)                                # float first column (codes are probed)
        got = rd.from_numpy_columns({"f": f, "b": b, "row": row}, override_num_blocks=4).sort(["f", "b"], descending=desc).to_numpy()
        np.testing.assert_array_equal(got["row"], row[orc.stable_argsort_multi([f, b], desc)])


def test_map_groups_boundaries_match_reference(cuda_device):
    """GroupedData.map_groups (grouped_data.py:105-236): groups = the reference's own get_key_boundaries on the
    sorted rows (tests/golden/map_groups_outputs.json), for one and for several key columns; one output block per
    sorted block; fn_args / fn_kwargs reach fn."""
    import json
    import os

    with open(os.path.join(os.path.dirname(__file__), "golden", "map_groups_outputs.json")) as f:
        cases = json.load(f)["cases"]
    for c in cases:
        cols = {k: np.asarray(v, dtype=np.float64 if k == "c" else np.int64) for k, v in c["columns"].items()}
        n = len(cols["row"])
        shuffled = np.random.default_rng(1).permutation(n)
        ds = rd.from_numpy_columns({k: v[shuffled] for k, v in cols.items()}, override_num_blocks=3)

        def fn(g, scale, offset=0):
            return {"first": g["row"][:1].copy(), "size": np.array([len(g["row"])], dtype=np.int64) * scale + offset,
                    "rows": np.array([int(np.sort(g["row"]).astype(np.int64).sum())], dtype=np.int64)}

        out = ds.groupby(c["keys"]).map_groups(fn, fn_args=(2,), fn_kwargs={"offset": 1})
        assert out.num_blocks() <= 3
        got = out.to_numpy()
        ends = np.asarray(c["boundaries"])
        starts = np.concatenate(([0], ends[:-1]))
        np.testing.assert_array_equal((got["size"] - 1) // 2, ends - starts)
        want_rows = [int(cols["row"][lo:hi].sum()) for lo, hi in zip(starts, ends)]   # same row SET per group
        np.testing.assert_array_equal(got["rows"], want_rows)
    # pandas batches, no key
    ds = rd.from_numpy_columns({"a": np.arange(10, dtype=np.int64)}, override_num_blocks=2)
    tot = ds.groupby(None).map_groups(lambda df: {"s": np.array([int(df["a"].sum())])}, batch_format="pandas").to_numpy()
    assert tot["s"].tolist() == [45]


def test_boundary_above_integer_range_and_nan_boundary(cuda_device):
    """find_partition_index (util.py:719-758) with a boundary no int8 value reaches: ascending every row lies below
    it, descending none lies at or above it (numpy.searchsorted on the reference's side); clamping the boundary to
    127 would misplace the rows equal to 127. A NaN boundary on an integer column is refused."""
    from ray_project_ray_b200.block import find_partition_indices

    col = np.array([-128, -5, 0, 3, 127, 127, 127], dtype=np.int8)
    asc = torch.from_numpy(col).to(cuda_device)
    assert find_partition_indices(asc, [300, 127, -129, 2.5], False) == [7, 4, 0, 3]
    assert find_partition_indices(asc, [300], False) == [int(np.searchsorted(col.astype(np.int64), 300, side="left"))]
    desc = torch.from_numpy(col[::-1].copy()).to(cuda_device)
    assert find_partition_indices(desc, [300, 127, -129], True) == [0, 3, 7]
    with pytest.raises(ValueError):
        find_partition_indices(asc, [float("nan")], False)


@pytest.mark.parametrize("sched", ["fused", "pull", "push"])
@pytest.mark.parametrize("ignore", [True, False])
def test_null_policy_matches_reference_outputs(sched, ignore, cuda_device):
    """Aggregations over nullable value columns equal the outputs of the reference's own generate_aggregate_fn
    (tests/golden/null_outputs.json): nulls skipped or poisoning per `ignore_nulls`, all-null groups null, Count counts
    rows, Mean / Std divide by the non-null count - grouped under every scheduler, and global."""
    import json
    import os

    from ray_project_ray_b200.aggregate import AbsMax

    with open(os.path.join(os.path.dirname(__file__), "golden", "null_outputs.json")) as f:
        g = json.load(f)
    _set_scheduler(sched)
    N, B = len(g["k"]), g["blocks"]
    n = N // B
    tables = []
    for i in range(B):
        sl = slice(i * n, (i + 1) * n)
        tables.append(pa.table({"k": np.asarray(g["k"][sl], dtype=np.int64),
                                "vi": pa.array(np.asarray(g["vi"][sl], dtype=np.int64), mask=np.asarray(g["null_i"][sl])),
                                "vf": pa.array(np.asarray(g["vf"][sl], dtype=np.float64), mask=np.asarray(g["null_f"][sl])),
                                "w": np.asarray(g["vi"][sl], dtype=np.int64)}))
    ds = rd.from_arrow(tables)
    assert ds.columns() == ["k", "vi", "vf", "w"]                    # validity columns are hidden
    back = ds.to_numpy()
    assert np.isnan(back["vi"][np.asarray(g["null_i"])]).all() and not np.isnan(back["vi"][~np.asarray(g["null_i"])]).any()
    aggs = [Count(), Sum("vi", ignore_nulls=ignore), Min("vi", ignore_nulls=ignore), Max("vi", ignore_nulls=ignore),
            Mean("vi", ignore_nulls=ignore), Std("vi", ignore_nulls=ignore), AbsMax("vi", ignore_nulls=ignore),
            Sum("vf", ignore_nulls=ignore), Min("vf", ignore_nulls=ignore), Max("vf", ignore_nulls=ignore),
            Mean("vf", ignore_nulls=ignore), Std("vf", ignore_nulls=ignore), Sum("w", ignore_nulls=ignore), Mean("w")]
    out = ds.groupby("k").aggregate(*aggs)
    want = g["grouped"][f"{'push' if sched == 'push' else 'pull'}_ignore{int(ignore)}"]
    assert out.columns() == list(want)
    rows = out.take_all()
    assert [r["k"] for r in rows] == want["k"] and [r["count()"] for r in rows] == want["count()"]
    for c in want:
        got = [r[c] for r in rows]
        assert [x is None for x in got] == [x is None for x in want[c]], c
        a = np.array([x for x in got if x is not None], dtype=np.float64)
        b = np.array([x for x in want[c] if x is not None], dtype=np.float64)
        if c.startswith(("sum(vi", "min(vi", "max(vi", "abs_max(vi", "sum(w")):
            np.testing.assert_array_equal(a, b, err_msg=c)             # integer results are exact
        else:
            np.testing.assert_allclose(a, b, rtol=FLOAT_RTOL, atol=0, err_msg=c)
    glob = ds.aggregate(*aggs)
    gw = g["global"][f"ignore{int(ignore)}"]
    for c, v in gw.items():
        if v[0] is None:
            assert glob[c] is None, c
        else:
            assert glob[c] is not None and abs(glob[c] - v[0]) <= FLOAT_RTOL * abs(v[0]), c
    # through Arrow and back: nulls where they were
    t = out.get_internal_block_refs()[0].to_arrow() if out.num_blocks() else None
    if t is not None and t.num_rows:
        assert t["sum(vi)"].null_count == sum(x is None for x in [r["sum(vi)"] for r in out.get_internal_block_refs()[0:1] for r in
                                                                  BlockAccessor.for_block(r).iter_rows()])
    # nullable keys are refused, not silently grouped under 0
    bad = rd.from_arrow(pa.table({"k": pa.array([1, None, 2]), "v": [1, 2, 3]}))
    with pytest.raises(NotImplementedError):
        bad.groupby("k").sum("v").take_all()
    # a nullable SORT key: nulls last in both directions, after NaN, in input order (pyarrow's sort_indices)
    rng = np.random.default_rng(3)
    kv = np.round(rng.standard_normal(4000), 1)
    kv[::97] = np.nan
    kmask = rng.random(4000) < 0.15
    t = pa.table({"k": pa.array(kv, mask=kmask), "row": np.arange(4000, dtype=np.int64)})
    for desc in (False, True):
        got = rd.from_arrow(t, override_num_blocks=None).sort("k", descending=desc).to_numpy()["row"]
        want = pac.sort_indices(t, sort_keys=[("k", "descending" if desc else "ascending")], null_placement="at_end").to_numpy()
        np.testing.assert_array_equal(got, want)


def test_from_pandas_missing_values_are_nulls(cuda_device):
    """pandas NaN / None in a numeric column = nulls in ray.data (pa.Table.from_pandas): ingested as a nullable column,
    so Count / Mean / Min see them as the reference does; non-numeric columns with missing values are refused."""
    import pandas as pd

    ds = rd.from_pandas(pd.DataFrame({"a": [1.0, np.nan, 3.0], "b": [1, 2, 3]}))
    assert ds.columns() == ["a", "b"]
    assert ds.sum("a") == 4.0 and ds.mean("a") == 2.0 and ds.min("a") == 1.0
    assert ds.sum("a", ignore_nulls=False) is None
    assert ds.aggregate(Count()) == {"count()": 3}
    ok = rd.from_pandas(pd.DataFrame({"a": [1.0, 2.0, 3.0]}))
    assert ok.sum("a") == 6.0
    with pytest.raises(ValueError, match="missing values"):
        rd.from_pandas(pd.DataFrame({"s": ["x", None, "z"]}))
    # nulls ride along through sort and shuffle as payload
    rng = np.random.default_rng(2)
    k = rng.permutation(5000).astype(np.int64)
    v = rng.random(5000)
    mask = rng.random(5000) < 0.3
    t = pa.table({"k": k, "v": pa.array(v, mask=mask)})
    s = rd.from_arrow(t).sort("k").to_numpy()
    order = np.argsort(k)
    np.testing.assert_array_equal(np.isnan(s["v"]), mask[order])
    np.testing.assert_array_equal(s["v"][~mask[order]], v[order][~mask[order]])
    sh = rd.from_arrow(t).random_shuffle(seed=1).to_numpy()
    p = orc.feistel_permutation(5000, 1)
    np.testing.assert_array_equal(np.isnan(sh["v"]), mask[p])


def test_int64_sum_wraps_within_and_across_blocks(cuda_device):
    """Documented divergence: group sums are int64 and wrap like pyarrow.compute.sum inside a block;
    the reference merges the per-block sums as Python ints (aggregate/_aggregate.py:119-120), which never wrap.
    Pinned here: the device result equals the exact sum modulo 2^64, whatever the block structure."""
    big = 2 ** 62
    k = np.array([1, 1, 1, 1, 2, 2], dtype=np.int64)
    v = np.array([big, big, big, 5, -big, 7], dtype=np.int64)
    exact = {1: 3 * big + 5, 2: -big + 7}
    wrap = {g: ((s + 2 ** 63) % 2 ** 64) - 2 ** 63 for g, s in exact.items()}
    for nblocks in (1, 3):
        got = rd.from_numpy_columns({"k": k, "v": v}, override_num_blocks=nblocks).groupby("k").sum("v").to_numpy()
        assert dict(zip(got["k"].tolist(), got["sum(v)"].tolist())) == wrap


@pytest.mark.parametrize("dist_kind", ["uniform", "zipf", "hot", "negative_i32"])
def test_dense_range_aggregation_equals_sorted_path(dist_kind, cuda_device):
    """Small-range integer keys take the sort-free path (csrc/groupby.cu "Dense-range aggregation"): the same groups,
    keys ascending, and bit-identical integer results as the sorted path and the oracle - also when one key holds a
    large share of the rows (combined per tile in shared memory before it reaches the table)."""
    from ray_project_ray_b200 import aggregate as A
    from ray_project_ray_b200.aggregate import AbsMax

    rng = np.random.default_rng(17)
    n = 400_000
    if dist_kind == "uniform":
        k = rng.integers(0, 50_000, n, dtype=np.int64)
    elif dist_kind == "zipf":
        k = (rng.zipf(1.1, n) % 30_000).astype(np.int64) * 3 - 7
    elif dist_kind == "hot":
        k = np.where(rng.random(n) < 0.6, 12345, rng.integers(0, 100_000, n)).astype(np.int64)
    else:
        k = rng.integers(-2000, 2000, n).astype(np.int32)
    v = rng.integers(-2**40, 2**40, n, dtype=np.int64)
    w = rng.integers(-5, 5, n, dtype=np.int64)
    ds = rd.from_numpy_columns({"k": k, "v": v, "w": w}, override_num_blocks=4)
    aggs = [Sum("v"), Mean("v"), Count(), Min("v"), Max("w"), AbsMax("w"), Sum("w", alias_name="sw")]
    calls = []
    orig = A.ops.dense_group_accumulate
    A.ops.dense_group_accumulate = lambda *a, **kw: (calls.append(1), orig(*a, **kw))[1]
    try:
        got = ds.groupby("k").aggregate(*aggs).to_numpy()
    finally:
        A.ops.dense_group_accumulate = orig
    assert calls, "the dense path did not run"
    saved = A.DENSE_MAX_RANGE
    A.DENSE_MAX_RANGE = 0
    try:
        ref = ds.groupby("k").aggregate(*aggs).to_numpy()
    finally:
        A.DENSE_MAX_RANGE = saved
    assert list(got) == list(ref)
    for c in got:
        assert got[c].dtype == ref[c].dtype
        np.testing.assert_array_equal(got[c], ref[c])        # Mean = the same int64 sum / count: identical bits
    keys, counts = np.unique(k, return_counts=True)
    np.testing.assert_array_equal(got["k"], keys)
    np.testing.assert_array_equal(got["count()"], counts)
    order = np.argsort(k, kind="stable")
    bounds = np.concatenate(([0], np.cumsum(counts)))[:-1]
    np.testing.assert_array_equal(got["sum(v)"], np.add.reduceat(v[order], bounds))
    np.testing.assert_array_equal(got["min(v)"], np.minimum.reduceat(v[order], bounds))
    np.testing.assert_array_equal(got["max(w)"], np.maximum.reduceat(w[order], bounds))
    # not applicable: float value column, wide key range -> the sorted path, same API
    wide = rd.from_numpy_columns({"k": k.astype(np.int64) * (2**40), "v": v}, override_num_blocks=2).groupby("k").sum("v").to_numpy()
    np.testing.assert_array_equal(wide["sum(v)"], np.add.reduceat(v[order], bounds))


def test_fused_scheduler_hands_over_when_the_dataset_is_too_large(cuda_device):
    """More rows than one fused pass takes (`fused_max_rows`, by default the argsort's 2^30-1): the default scheduler
    runs the task-structured one with sampled boundaries instead of failing; same global result."""
    ctx = DataContext.get_current()
    rng = np.random.default_rng(8)
    k = rng.integers(0, 1000, 50_000, dtype=np.int64)
    v = rng.integers(-9, 9, 50_000, dtype=np.int64)
    ds = rd.from_numpy_columns({"k": k, "v": v}, override_num_blocks=5)
    ctx.set_config("fused_max_rows", 10_000)
    try:
        got = ds.sort("k", descending=True).to_numpy()
        agg = ds.groupby("k").sum("v").to_numpy()
    finally:
        ctx.set_config("fused_max_rows", (1 << 30) - 1)
    order = orc.stable_argsort(k, True)
    np.testing.assert_array_equal(got["k"], k[order])
    np.testing.assert_array_equal(got["v"], v[order])
    keys = np.unique(k)
    np.testing.assert_array_equal(agg["k"], keys)
    np.testing.assert_array_equal(agg["sum(v)"], [v[k == x].sum() for x in keys])
