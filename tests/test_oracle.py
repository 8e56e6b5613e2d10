"""Pins the CPU oracle (oracle/exchange_oracle.py) against
  (a) outputs of the reference's own code (tests/golden/, generated under the stub-ray shim),
  (b) the golden vectors written in the reference's tests,
  (c) the installed pyarrow, whose sort_indices / group_by the reference calls.
Runs without a GPU."""
import math

import numpy as np
import pyarrow as pa
import pyarrow.compute as pac
import pytest

from oracle import exchange_oracle as orc
from tests import _golden as G


@pytest.mark.parametrize("case", [c["tag"] for c in G.meta()["sort_cases"]])
def test_sort_matches_reference_per_block(case):
    info = next(c for c in G.meta()["sort_cases"] if c["tag"] == case)
    a = G.arrays()
    blocks = G.sort_case_blocks(info["case"])
    bounds = [(b,) for b in a[case + "_bounds"]]
    out = orc.sort_dataset(blocks, "k", info["descending"], bounds)
    want = G.split_by_sizes(a[case + "_rows"], a[case + "_sizes"])
    assert [orc.num_rows(b) for b in out] == [len(w) for w in want]
    for got, w in zip(out, want):
        if len(w):
            np.testing.assert_array_equal(got["row"], w)


@pytest.mark.parametrize("desc", [False, True])
def test_sort_nan_keys_global_order_matches_reference(desc):
    a = G.arrays()
    key = a["sort_f64_nan_key"]
    want = a[f"sort_f64_nan_{'desc' if desc else 'asc'}_rows"]
    np.testing.assert_array_equal(orc.stable_argsort(key, desc), want)


@pytest.mark.parametrize("desc", [False, True])
def test_two_column_sort_matches_reference(desc):
    a = G.arrays()
    got = orc.stable_argsort_multi([a["sort2_k1"], a["sort2_k2"]], desc)
    np.testing.assert_array_equal(got, a[f"sort2_{'desc' if desc else 'asc'}_rows"])


def test_specified_boundaries_golden_of_the_reference_tests():
    # python/ray/data/tests/test_sort.py:23-51, reproduced by the reference run in the fixtures
    ids = np.arange(1000, dtype=np.int64)
    blocks = [{"id": ids[i * 50:(i + 1) * 50]} for i in range(20)]
    for case in G.meta()["specified_boundaries"]:
        desc, bounds = case["descending"], case["boundaries"]
        out = orc.sort_dataset(blocks, "id", desc, orc.user_boundaries(bounds))
        # what the reference test asserts, computed independently
        cuts = [0] + sorted(round(b) for b in bounds) + [1000]
        expected = [list(range(cuts[i], cuts[i + 1])) for i in range(len(cuts) - 1)]
        if desc:
            expected = [list(reversed(b)) for b in reversed(expected)]
        assert [list(b["id"]) if b else [] for b in out] == expected
        assert expected == case["blocks"]  # and what the reference itself produced


def test_map_partitions_match_reference():
    a = G.arrays()
    n = G.meta()["sort_shape"]["rows_per_block"]
    t0 = {"k": a["sort_i64_ties_key"][:n], "row": np.arange(n, dtype=np.int64)}
    parts = orc.sort_and_partition(t0, [(-100,), (0,), (100,)], ["k"], False)
    np.testing.assert_array_equal(np.concatenate([p["row"] for p in parts]), a["map_asc_rows"])
    assert [orc.num_rows(p) for p in parts] == list(a["map_asc_sizes"])
    parts = orc.sort_and_partition(t0, [(100,), (0,), (-100,)], ["k"], True)
    np.testing.assert_array_equal(np.concatenate([p["row"] for p in parts]), a["map_desc_rows"])
    assert [orc.num_rows(p) for p in parts] == list(a["map_desc_sizes"])
    # empty block -> len(boundaries)+1 empty partitions (tests/test_sort.py:142-191)
    assert [orc.num_rows(p) for p in orc.sort_and_partition({}, [1, 5, 10], ["A"], False)] == G.meta()["empty_sort_and_partition"] == [0, 0, 0, 0]
    assert orc.num_rows(orc.merge_sorted_blocks([{}], ["A"], False)) == 0


def test_ties_land_in_one_block():
    # tests/test_sort.py:79-94: all-equal keys over 100 blocks -> exactly one non-empty output
    blocks = [{"item": np.ones(1, dtype=np.int64)} for _ in range(100)]
    bounds = orc.boundaries_from_samples(orc.concat(blocks), ["item"], 100)
    out = orc.sort_dataset(blocks, "item", False, bounds)
    sizes = [orc.num_rows(b) for b in out]
    assert len(sizes) == 100 and sorted(sizes)[-1] == 100 and sum(1 for s in sizes if s) == 1


def test_sample_boundaries_quantiles_match_reference():
    for case in G.meta()["sample_boundaries"]:
        got = orc.boundaries_from_samples({"k": np.asarray(case["values"], dtype=np.int64)}, ["k"], case["R"])
        assert [int(b[0]) for b in got] == case["boundaries"]
    assert len(orc.boundaries_from_samples({}, ["id"], 3)) == G.meta()["sample_boundaries_empty"] == 2


@pytest.mark.parametrize("kind", ["int", "float", "float_special"])
@pytest.mark.parametrize("desc", [False, True])
def test_key_order_equals_pyarrow_sort_indices(kind, desc):
    rng = np.random.default_rng(5)
    n = 20_000
    if kind == "int":
        k = rng.integers(-100, 100, n)
    elif kind == "float":
        k = np.round(rng.standard_normal(n), 1)
    else:
        k = np.round(rng.standard_normal(n), 1)
        k[::7] = np.nan
        k[::11] = 0.0
        k[::13] = -0.0
        k[5], k[6] = np.inf, -np.inf
        k[8] = -np.nan
    want = pac.sort_indices(pa.table({"k": k}), [("k", "descending" if desc else "ascending")]).to_numpy()
    np.testing.assert_array_equal(orc.stable_argsort(k, desc), want)


@pytest.mark.parametrize("sched", ["pull", "push"])
def test_groupby_matches_reference(sched):
    a, m = G.arrays(), G.meta()
    B, n = m["sort_shape"]["blocks"], m["sort_shape"]["rows_per_block"]
    blocks = [{"k": a["agg_k"][i * n:(i + 1) * n], "v": a["agg_v"][i * n:(i + 1) * n], "f": a["agg_f"][i * n:(i + 1) * n],
               "unused": np.arange(i * n, (i + 1) * n)} for i in range(B)]
    aggs = [("sum", "v"), ("mean", "f"), ("count", None), ("min", "v"), ("max", "f"), ("std", "f"), ("mean", "v"), ("sum", "f")]
    bounds = orc.boundaries_from_samples({"k": a["agg_k"][::97]}, ["k"], B)
    out = orc.concat(orc.aggregate_dataset(blocks, "k", aggs, bounds))
    tag = f"agg_{sched}"
    assert list(out.keys()) == m[tag + "_columns"] == ["k", "sum(v)", "mean(f)", "count()", "min(v)", "max(f)", "std(f)", "mean(v)", "sum(f)"]
    for c in ("k", "sum(v)", "count()", "min(v)"):
        np.testing.assert_array_equal(np.asarray(out[c], dtype=np.int64), a[f"{tag}_{c}"])
    np.testing.assert_array_equal(out["max(f)"], a[f"{tag}_max(f)"])
    for c in ("mean(f)", "std(f)", "mean(v)", "sum(f)"):
        np.testing.assert_allclose(np.asarray(out[c], dtype=np.float64), a[f"{tag}_{c}"], rtol=1e-9)
    assert (np.diff(out["k"]) > 0).all()  # ascending, one row per key


def test_small_groupby_goldens_of_the_reference_tests():
    m = G.meta()
    xs = m["small_groupby_input"]
    blocks = [{"A": np.array([x % 3 for x in xs[i::4]]), "B": np.array(xs[i::4])} for i in range(4)]
    aggs = [("sum", "B"), ("mean", "B"), ("count", None), ("min", "B"), ("max", "B"), ("std", "B")]
    out = orc.concat(orc.aggregate_dataset(blocks, "A", aggs, [(1,), (2,), (2,)]))
    # python/ray/data/tests/test_all_to_all.py:265-271, :529-535, :208-214
    assert list(out["sum(B)"]) == [1683, 1617, 1650] == m["small_groupby"]["sum(B)"]
    assert list(out["mean(B)"]) == [49.5, 49.0, 50.0] == m["small_groupby"]["mean(B)"]
    assert list(out["count()"]) == [34, 33, 33] == m["small_groupby"]["count()"]
    assert list(out["min(B)"]) == m["small_groupby"]["min(B)"] and list(out["max(B)"]) == m["small_groupby"]["max(B)"]
    np.testing.assert_allclose(out["std(B)"], m["small_groupby"]["std(B)"], rtol=1e-12)
    assert m["dup_names_columns"] == ["A", "sum(B)", "sum(B)_2", "count()", "count()_2"]


def test_groupby_equals_pyarrow_group_by():
    rng = np.random.default_rng(0)
    n = 30_000
    k = rng.integers(0, 300, n)
    v = rng.integers(-2**31, 2**31, n)
    f = rng.random(n)
    want = pa.table({"k": k, "v": v, "f": f}).group_by("k").aggregate([("v", "sum"), ("f", "mean"), ("v", "count")]).sort_by("k")
    out = orc.concat(orc.aggregate_dataset([{"k": k, "v": v, "f": f}], "k", [("sum", "v"), ("mean", "f"), ("count", None)], []))
    np.testing.assert_array_equal(out["k"], want["k"].to_numpy())
    np.testing.assert_array_equal(out["sum(v)"], want["v_sum"].to_numpy())
    np.testing.assert_array_equal(out["count()"], want["v_count"].to_numpy())
    np.testing.assert_allclose(out["mean(f)"], want["f_mean"].to_numpy(), rtol=1e-12)


def test_shuffle_structure_matches_reference_properties():
    m = G.meta()
    B, n = m["sort_shape"]["blocks"], m["sort_shape"]["rows_per_block"]
    row = np.arange(B * n, dtype=np.int64)
    blocks = [{"row": row[i * n:(i + 1) * n]} for i in range(B)]
    a = orc.shuffle_dataset(blocks, 0)
    b = orc.shuffle_dataset(blocks, 0)
    c = orc.shuffle_dataset(blocks, 1)
    ref = m["random_shuffle"]["pull"]
    assert ref["deterministic_per_seed"] and all(np.array_equal(x["row"], y["row"]) for x, y in zip(a, b))
    assert ref["seed_changes_output"] and not all(np.array_equal(x["row"], y["row"]) for x, y in zip(a, c))
    assert ref["is_permutation"] and np.array_equal(np.sort(np.concatenate([x["row"] for x in a])), row)
    # which reducer gets each mapper's short slice is random: sizes lie between B*309 and B*313 in both
    for sizes in ([len(x["row"]) for x in a], ref["sizes"]):
        assert sum(sizes) == B * n and all(B * (n - 7 * 313) <= s <= B * 313 for s in sizes)
    assert ref["every_output_sees_every_input"] and all(len(set(x["row"] // n)) == B for x in a)
    # slice sizes of one map task (ceil(n/R) rows, last slices short / empty)
    sizes = sorted(orc.num_rows(s) for s in orc.shuffle_map(3, {"row": row[:1003]}, 8, 5))
    assert sizes == m["shuffle_map_slice_sizes_sorted"]
    sizes = sorted(orc.num_rows(s) for s in orc.shuffle_map(0, {"row": row[:5]}, 8, 5))
    assert sizes == m["shuffle_map_tiny_slice_sizes_sorted"]


def test_feistel_is_a_bijection_with_inverse():
    for n in (1, 2, 3, 5, 64, 1000, 4097):
        for seed in (0, 7):
            p = orc.feistel_permutation(n, seed)
            assert np.array_equal(np.sort(p), np.arange(n))
            q = orc.feistel_permutation(n, seed, inverse=True)
            assert np.array_equal(q[p], np.arange(n))


def test_quantile_restatement_equals_reference_outputs():
    """Quantile (aggregate/_aggregate.py:360-431): the value at an exact rank, the interpolation rounded
    to 5 decimals between ranks, equal to what the reference's code returned. (The reference's column type
    is int64 or double depending on which groups share an output block; values are what is pinned.)"""
    q = G.quantile()
    k, vi, vf = np.array(q["k"], np.int64), np.array(q["vi"], np.int64), np.array(q["vf"], np.float64)
    B = q["blocks"]
    n = len(k) // B
    blocks = [{"k": k[i * n:(i + 1) * n], "vi": vi[i * n:(i + 1) * n], "vf": vf[i * n:(i + 1) * n]} for i in range(B)]
    assert q["grouped"]["pull"] == q["grouped"]["push"]
    ref = q["grouped"]["pull"]
    for j, qq in enumerate(q["qs"]):
        for col, name in (("vi", f"qi{j}"), ("vf", f"qf{j}")):
            got = orc.quantile_dataset(blocks, "k", col, qq)
            assert list(got["k"]) == ref["k"]
            assert list(got[f"quantile({col})"]) == ref[name]
            assert list(orc.quantile_dataset(blocks, None, col, qq)[f"quantile({col})"]) == q["global"][name]
    xs = q["small_input"]
    small = [{"A": np.array([x % 3 for x in xs[i::4]]), "B": np.array(xs[i::4])} for i in range(4)]
    assert list(orc.quantile_dataset(small, "A", "B")["quantile(B)"]) == q["small_grouped"]["quantile(B)"] == [49.5, 49, 50]
    assert list(orc.quantile_dataset(small, "A", "B", 0.9)["quantile(B)"]) == q["small_grouped"]["q9"]
    assert list(orc.quantile_dataset(small, None, "B")["quantile(B)"]) == q["small_global"]["quantile(B)"]


def test_byte_string_key_order_equals_pyarrow():
    """fixed_size_binary keys: the oracle's order is what pyarrow.compute.sort_indices gives (asc, desc, ties)."""
    import pyarrow as pa
    import pyarrow.compute as pac

    rng = np.random.default_rng(8)
    for w in (1, 3, 8, 10, 17):
        col = rng.integers(0, 4 if w > 3 else 256, (5000, w)).astype(np.uint8)   # few symbols: many ties and shared prefixes
        col[rng.integers(0, 5000, 50)] = 0
        col[rng.integers(0, 5000, 50)] = 255
        arr = pa.FixedSizeBinaryArray.from_buffers(pa.binary(w), len(col), [None, pa.py_buffer(col.tobytes())])
        for desc in (False, True):
            want = pac.sort_indices(pa.table({"k": arr}), [("k", "descending" if desc else "ascending")]).to_numpy()
            np.testing.assert_array_equal(orc.stable_argsort_bytes(col, desc), want)


def test_key_boundaries_match_reference_map_groups():
    """oracle.key_boundaries == the reference's own get_key_boundaries (tests/golden/make_golden.py run_map_groups)."""
    import json
    import os

    with open(os.path.join(os.path.dirname(__file__), "golden", "map_groups_outputs.json")) as f:
        cases = json.load(f)["cases"]
    assert len(cases) >= 5
    for c in cases:
        keys = [c["keys"]] if isinstance(c["keys"], str) else c["keys"]
        cols = [np.asarray(c["columns"][k]) for k in keys]
        assert orc.key_boundaries(cols) == c["boundaries"]


def _null_golden():
    import json
    import os

    with open(os.path.join(os.path.dirname(__file__), "golden", "null_outputs.json")) as f:
        return json.load(f)


def test_null_policy_oracle_matches_reference_outputs():
    """oracle.null_group_aggregate == the reference's own generate_aggregate_fn on nullable columns (pull and push,
    ignore_nulls True / False, grouped and global): tests/golden/make_golden.py run_nulls."""
    g = _null_golden()
    k = np.asarray(g["k"])
    cols = {"vi": (np.asarray(g["vi"], dtype=np.int64), np.asarray(g["null_i"])),
            "vf": (np.asarray(g["vf"], dtype=np.float64), np.asarray(g["null_f"])),
            "w": (np.asarray(g["vi"], dtype=np.int64), np.zeros(len(k), dtype=bool))}
    checked = 0
    for tag, out in list(g["grouped"].items()) + [("global_" + t, o) for t, o in g["global"].items()]:
        ignore = tag.endswith("1")
        keys = None if tag.startswith("global") else k
        assert out["count()"] == ([len(k)] if keys is None else np.unique(k, return_counts=True)[1].tolist())
        for name, got in out.items():
            if "(" not in name or name == "count()":
                continue
            kind, col = name[:-1].split("(")
            want = orc.null_group_aggregate(keys, cols[col][0], cols[col][1], kind, ignore or name == "mean(w)")
            assert [x is None for x in got] == [x is None for x in want], (tag, name)
            a = np.array([x for x in got if x is not None], dtype=np.float64)
            b = np.array([x for x in want if x is not None], dtype=np.float64)
            np.testing.assert_allclose(a, b, rtol=1e-12, atol=1e-9, err_msg=f"{tag} {name}")
            checked += 1
    assert checked >= 6 * 13
