"""CPU tests of the host-side mirror: scheduler arithmetic against reference outputs, argument
validation, and that libb200x.so loads and exports every symbol include/b200x.h declares."""
import ctypes
import os
import re

import numpy as np
import pytest

from ray_project_ray_b200 import _lib
from ray_project_ray_b200.aggregate import AggregateFn, Count, Mean, Sum, resolved_names
from ray_project_ray_b200.planner import generate_aggregate_fn
from ray_project_ray_b200.schedulers import (PushBasedShuffleTaskScheduler, _MergeTaskSchedule,
                                             shuffle_output_sizes)
from ray_project_ray_b200.sort_task_spec import SortKey
from tests import _golden as G


def test_library_exports_every_declared_symbol():
    header = open(_lib.HEADER_PATH).read()
    declared = set(re.findall(r"\b(b200x_[a-z0-9_]+)\s*\(", header))
    assert declared, "no declarations parsed"
    lib = _lib.load()
    for name in declared:
        assert hasattr(lib, name), f"{name} declared in include/b200x.h but not exported"
    assert declared == set(_lib.SIGNATURES), declared ^ set(_lib.SIGNATURES)
    assert lib.b200x_version() == 100
    assert isinstance(lib.b200x_device_count(), int)


def test_argument_errors_do_not_need_a_gpu():
    lib = _lib.load()
    need = ctypes.c_size_t(0)
    # size query works on the host alone
    assert lib.b200x_radix_argsort(None, 0, 0, None, None, None, None, 1000, 0, 0, None, ctypes.byref(need), None) == 0
    assert need.value >= 1000 * 24
    # too many rows / bad key type are argument errors with a message
    assert lib.b200x_radix_argsort(None, 0, 0, None, None, None, None, 1 << 30, 0, 0, None, ctypes.byref(need), None) == 2
    assert b"2^30" in lib.b200x_last_error_string()
    assert lib.b200x_radix_argsort(None, 99, 0, None, None, None, None, 10, 0, 0, None, ctypes.byref(need), None) == 1
    with pytest.raises(_lib.B200XError):
        _lib.call("b200x_perm_indices_host", 10, 0, 0, 5, 10, None)


@pytest.mark.parametrize("case", range(len(G.meta()["push_schedules"])))
def test_push_based_schedule_equals_reference(case):
    g = G.meta()["push_schedules"][case]
    st = PushBasedShuffleTaskScheduler._compute_shuffle_schedule(g["cpus"], g["num_input_blocks"], g["merge_factor"],
                                                                 g["num_output_blocks"])
    ms = st.merge_schedule
    assert st.num_rounds == g["num_rounds"]
    assert st.num_map_tasks_per_round == g["num_map_tasks_per_round"]
    assert st._merge_task_placement == g["placement"]
    assert ms.num_merge_tasks_per_round == g["num_merge_tasks_per_round"]
    assert ms.num_reducers_per_merger == g["num_reducers_per_merger"]
    assert [ms.get_num_reducers_per_merge_idx(i) for i in range(ms.num_merge_tasks_per_round)] == g["reducers_per_merge_idx"]
    assert [ms.get_merge_idx_for_reducer_idx(j) for j in range(g["num_output_blocks"])] == g["merge_idx_for_reducer"]
    assert list(ms.round_robin_reduce_idx_iterator()) == g["round_robin"]
    assert st.get_estimated_num_refs() == g["estimated_num_refs"]


@pytest.mark.parametrize("num_input_blocks,merge_factor,cpus", [
    (100, 2, {"a": 100}), (100, 3, {"a": 10, "b": 10, "c": 10}), (1000, 2, {f"n{i}": 16 for i in range(20)}),
    (260, 2, {"a": 128}), (1, 2, {"a": 128}), (100, 1.5, {"a": 7, "b": 9}), (64, 2, {f"gpu{i}": 3 for i in range(8)})])
def test_push_based_schedule_invariants(num_input_blocks, merge_factor, cpus):
    """The invariants python/ray/data/tests/test_sort.py:305-419 checks on the schedule."""
    num_output_blocks = num_input_blocks
    st = PushBasedShuffleTaskScheduler._compute_shuffle_schedule(cpus, num_input_blocks, merge_factor, num_output_blocks)
    ms = st.merge_schedule
    total_cpus = sum(cpus.values())
    # all inputs are covered
    assert st.num_rounds * st.num_map_tasks_per_round >= num_input_blocks
    # CPUs are not oversubscribed
    assert st.num_map_tasks_per_round + ms.num_merge_tasks_per_round <= max(total_cpus, 2)
    # every reducer is fed by exactly one merger, reducers per merger differ by <= 1
    per = [ms.get_num_reducers_per_merge_idx(i) for i in range(ms.num_merge_tasks_per_round)]
    assert sum(per) == num_output_blocks and max(per) - min(per) <= 1
    seen = list(ms.round_robin_reduce_idx_iterator())
    assert sorted(seen) == list(range(num_output_blocks))
    for j in range(num_output_blocks):
        m = ms.get_merge_idx_for_reducer_idx(j)
        assert ms._first_reducer(m) <= j < ms._first_reducer(m) + per[m]
    # mergers are spread over nodes in proportion to their CPUs
    if num_input_blocks >= total_cpus:
        for node, c in cpus.items():
            assert abs(st._merge_task_placement.count(node) - c / (merge_factor + 1)) <= 1


def test_merge_schedule_fewer_reducers_than_mergers():
    ms = _MergeTaskSchedule(3, 8)
    assert ms.num_merge_tasks_per_round == 3 and ms.num_reducers_per_merger == 1
    assert list(ms.round_robin_reduce_idx_iterator()) == [0, 1, 2]


def test_sort_key_validation_matches_reference_errors():
    # sort_task_spec.py:31-54
    with pytest.raises(ValueError, match="Key must be a string"):
        SortKey(3)
    with pytest.raises(ValueError, match="Length of `descending`"):
        SortKey(["a", "b"], [True])
    with pytest.raises(ValueError, match="mixed key orders"):
        SortKey(["a", "b"], [True, False])
    with pytest.raises(ValueError, match="int or float"):
        SortKey("a", boundaries=["x"])
    k = SortKey("a", True, [5, 1, 5, 3.5])
    assert k.boundaries == [1, 3.5, 5] and k.get_descending() and k.get_columns() == ["a"]
    assert k.to_arrow_sort_args() == [("a", "descending")] and k.to_pandas_sort_args() == (["a"], False)
    assert SortKey(None).get_columns() == []

    class S:
        names = ["a", "b"]

    SortKey("a").validate_schema(S())
    with pytest.raises(ValueError, match="does not exist"):
        SortKey("c").validate_schema(S())


def test_aggregate_names_and_validation():
    assert [a.name for a in (Sum("B"), Mean("B"), Count())] == ["sum(B)", "mean(B)", "count()"]
    assert Sum("B", alias_name="total").name == "total"
    assert resolved_names([Sum("B"), Sum("B"), Count(), Count()]) == G.meta()["dup_names_columns"][1:]
    with pytest.raises(ValueError, match="at least one aggregation"):
        generate_aggregate_fn("A", [])
    with pytest.raises(ValueError, match="Exactly one of"):
        AggregateFn(init=lambda k: 0, merge=lambda a, b: a + b)


def test_shuffle_output_sizes_follow_reference_slicing():
    # 8 blocks x 2500 rows into 8 outputs: slices of ceil(2500/8)=313 rows, the last one 309
    sizes = shuffle_output_sizes([2500] * 8, 8)
    assert sum(sizes) == 20000 and sizes == [8 * 313] * 7 + [8 * 309]
    assert shuffle_output_sizes([5], 8) == [1, 1, 1, 1, 1, 0, 0, 0]


def test_product_never_imports_the_oracle():
    root = os.path.dirname(_lib._HERE)
    pkg = os.path.join(root, "ray-project-ray_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith(".py"):
                src = open(os.path.join(dirpath, f)).read()
                assert "oracle" not in src, f"{f} mentions the oracle"


def test_batchers_equal_reference_batch_boundaries_and_shuffle_buffers(monkeypatch):
    """Batcher / ShufflingBatcher host logic (_internal/batcher.py) against the reference's own classes:
    same batches, and - with the shuffle replaced by a recorder on both sides - the same rows in the
    shuffle buffer at every re-shuffle with the same seeds. Runs on CPU tensors (no kernels involved)."""
    import torch

    from ray_project_ray_b200.batcher import blocks_to_batches
    from ray_project_ray_b200.block import DeviceBlock, DeviceBlockAccessor
    from tests import _golden as G

    shuffles = []

    def recorder(self, random_seed):
        shuffles.append({"seed": random_seed, "rows": self._block.columns["row"].tolist()})
        return self._block

    monkeypatch.setattr(DeviceBlockAccessor, "random_shuffle", recorder)
    for case in G.batcher()["cases"]:
        del shuffles[:]
        lo, blocks = 0, []
        for sz in case["block_sizes"]:
            blocks.append(DeviceBlock({"row": torch.arange(lo, lo + sz)}, device=torch.device("cpu")))
            lo += sz
        out = [b.columns["row"].tolist() for b in blocks_to_batches(
            iter(blocks), batch_size=case["batch_size"], drop_last=case["drop_last"],
            shuffle_buffer_min_size=case["shuffle_buffer_min_size"], shuffle_seed=11)]
        assert out == case["batches"], case["block_sizes"]
        assert shuffles == case["shuffles"], case["block_sizes"]


def test_randomize_block_order_equals_reference_order():
    """planner/randomize_blocks.py: random.seed(seed) + random.shuffle of the block list - the same calls,
    so the same block order as the reference's own function returned (tests/golden/batcher_outputs.json)."""
    import torch

    from ray_project_ray_b200.block import DeviceBlock, DeviceBlockAccessor
    from ray_project_ray_b200.interfaces import RefBundle
    from ray_project_ray_b200.planner import RandomizeBlocks, generate_randomize_blocks_fn
    from tests import _golden as G

    for seed, want in G.batcher()["randomize_block_order"].items():
        blocks = [DeviceBlock({"row": torch.tensor([i])}, device=torch.device("cpu")) for i in range(9)]
        refs = [RefBundle([(b, DeviceBlockAccessor(b).get_metadata(None, None))], owns_blocks=True) for b in blocks]
        out, stats = generate_randomize_blocks_fn(RandomizeBlocks(None, seed=int(seed)))(refs, None)
        assert [int(r.blocks[0][0].columns["row"][0]) for r in out] == want
        assert list(stats) == ["RandomizeBlockOrder"] and len(stats["RandomizeBlockOrder"]) == 9
    assert generate_randomize_blocks_fn(RandomizeBlocks(None, seed=1))([], None) == ([], {"RandomizeBlockOrder": []})


def test_record_layout_helpers_on_host_tensors():
    """Pitch detection, member views and the copy-free re-join of record slices are plain stride arithmetic."""
    import torch

    from ray_project_ray_b200.block import _inner, _join_views, _row_pitch, auto_pitch, member_view

    assert [auto_pitch(x) for x in (1, 16, 17, 33, 64, 65, 100, 128, 129, 300)] == [16, 16, 32, 64, 64, 128, 128, 128, 256, 384]
    rec = torch.arange(10 * 128, dtype=torch.int64).to(torch.uint8).reshape(10, 128)
    key = member_view(rec, 0, torch.uint8, (10,))
    pay = member_view(rec, 10, torch.uint8, (90,))
    val = member_view(rec, 104, torch.int64, ())
    assert key.shape == (10, 10) and pay.shape == (10, 90) and val.shape == (10,)
    assert _row_pitch(key) == 128 and _row_pitch(pay) == 128 and _row_pitch(val) == 128 and _row_pitch(rec) is None
    assert _inner(pay) == (90, True) and _inner(val) == (8, True)
    assert torch.equal(pay, rec[:, 10:100]) and torch.equal(val, rec[:, 104:112].contiguous().view(torch.int64).reshape(-1))
    parts = [pay[0:3], pay[3:4], pay[4:10]]
    joined = _join_views(parts)
    assert joined.data_ptr() == pay.data_ptr() and joined.shape == (10, 90) and torch.equal(joined, pay)
    other = pay.clone()                                        # a separate allocation is copied, not aliased
    cat = _join_views([pay[0:3], other[3:10]])
    assert cat.data_ptr() != pay.data_ptr() and torch.equal(cat, pay)
    transposed = torch.zeros(4, 6).t()                         # not dense inside a row: neither a record member nor joined
    assert _row_pitch(transposed) is None


def test_hidden_validity_columns_on_host_blocks():
    """Nullable columns at the edges of the path (no kernel involved): Arrow nulls in, hidden `::valid` column, schema
    and column names without it, `select` and `concat` keep it, Arrow nulls / NaN / None out."""
    import pyarrow as pa
    import torch

    from ray_project_ray_b200.block import DeviceBlock, DeviceBlockAccessor, is_hidden, valid_name

    cpu = torch.device("cpu")
    t = pa.table({"k": [1, 2, 3, 4], "v": pa.array([1.5, None, 3.5, None]), "i": pa.array([7, 8, None, 9])})
    b = DeviceBlock.from_arrow(t, cpu)
    assert b.visible_names() == ["k", "v", "i"] and b.schema().names == ["k", "v", "i"]
    assert [c for c in b.column_names if is_hidden(c)] == [valid_name("v"), valid_name("i")]
    assert b.validity("k") is None and b.validity("v").tolist() == [1, 0, 1, 0]
    host = b.to_numpy()
    assert np.isnan(host["v"][[1, 3]]).all() and host["v"][0] == 1.5
    assert host["i"].dtype == np.float64 and np.isnan(host["i"][2]) and host["i"][3] == 9      # as pyarrow's to_numpy
    back = b.to_arrow()
    assert back["v"].to_pylist() == [1.5, None, 3.5, None] and back["i"].to_pylist() == [7, 8, None, 9]
    rows = list(DeviceBlockAccessor(b).iter_rows())
    assert rows[1] == {"k": 2, "v": None, "i": 8.0} and rows[2]["i"] is None
    sel = b.select(["v"])
    assert sel.column_names == ["v", valid_name("v")] and sel.schema().names == ["v"]
    plain = DeviceBlock.from_arrow(pa.table({"k": [5], "v": [9.0], "i": [1]}), cpu)            # no nulls: no hidden columns
    assert not any(is_hidden(c) for c in plain.column_names)
    both = DeviceBlock.concat([b, plain])
    assert both.num_rows == 5 and both.validity("v").tolist() == [1, 0, 1, 0, 1] and both.validity("i").tolist() == [1, 1, 0, 1, 1]
    with pytest.raises(ValueError):
        DeviceBlock.from_arrow(pa.table({"s": pa.array([b"ab", None], type=pa.binary(2))}), cpu)
