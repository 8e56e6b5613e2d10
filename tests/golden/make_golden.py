"""Generate tests/golden/*.npz|json by running the REFERENCE's own hot-path code.

Runs only in the build container (needs /root/reference; `ray` itself cannot be imported because
its Cython/C++ core is not built). The reference's pure-Python exchange code is imported
unmodified under a stub top-level `ray` package, and `ray.remote/get/wait/nodes` are replaced by
an eager in-process fake so that BOTH of the reference's schedulers run as written
(SURVEY.md §8c). Nothing here is shipped or imported by the product; the outputs are committed
so that tests can run on the GPU box, where /root/reference does not exist.

    python tests/golden/make_golden.py            # everything
    python tests/golden/make_golden.py quantile   # only quantile_outputs.json
    python tests/golden/make_golden.py batcher    # only batcher_outputs.json
"""
import importlib
import inspect
import json
import os
import pickle
import random
import sys
import types

import numpy as np
import pyarrow as pa

REF = "/root/reference/python/ray"
OUT = os.path.dirname(os.path.abspath(__file__))
sys.dont_write_bytecode = True


def _mod(name, path=None, **attrs):
    m = types.ModuleType(name)
    if path is not None:
        m.__path__ = [path] if path else []
    m.__dict__.update(attrs)
    sys.modules[name] = m
    return m


class Ref:
    def __init__(self, v):
        self.v = v

    def __class_getitem__(cls, k):
        return cls


class RemoteFn:
    def __init__(self, fn, opts=None):
        self.fn, self.opts = fn, dict(opts or {})

    def options(self, **kw):
        return RemoteFn(self.fn, {**self.opts, **kw})

    def remote(self, *a, **kw):
        a = [x.v if isinstance(x, Ref) else x for x in a]
        kw = {k: (v.v if isinstance(v, Ref) else v) for k, v in kw.items()}
        out = self.fn(*a, **kw)
        n = self.opts.get("num_returns", 1)
        if inspect.isgenerator(out):
            out = list(out)
        if n == 1:
            return Ref(out)
        out = list(out)
        assert len(out) == n, (len(out), n)
        return [Ref(o) for o in out]


def install_shim(num_nodes=8, cpus_per_node=3):
    class _RC:
        def get_node_id(self):
            return "node0"

    ray = _mod("ray", REF, ObjectRef=Ref, DynamicObjectRefGenerator=object, cloudpickle=pickle)
    ray.runtime_context = _mod("ray.runtime_context", get_runtime_context=lambda: _RC())
    priv = _mod("ray._private", "")
    rc = _mod("ray._private.ray_constants",
              env_bool=lambda k, d: bool(int(os.environ.get(k, d))), env_integer=lambda k, d: int(os.environ.get(k, d)),
              LOGGER_FORMAT="%(message)s", LOGGER_LEVEL="info", CALLER_MEMORY_USAGE_PER_OBJECT_REF=3000)
    ut = _mod("ray._private.utils", _get_pyarrow_version=lambda: pa.__version__)
    wk = _mod("ray._private.worker", global_worker=types.SimpleNamespace(
        core_worker=types.SimpleNamespace(get_local_memory_store_bytes_used=lambda: 0)))
    priv.ray_constants, priv.utils, priv.worker = rc, ut, wk
    ray._private = priv
    u = _mod("ray.util", REF + "/util")
    ann = importlib.import_module("ray.util.annotations")
    u.PublicAPI, u.DeveloperAPI, u.Deprecated = ann.PublicAPI, ann.DeveloperAPI, ann.Deprecated

    class _Client:
        class ray:
            @staticmethod
            def is_connected():
                return True

    u.client = _Client
    ray.util = u
    importlib.import_module("ray.types")
    sys.modules["ray.types"].ObjectRef = Ref
    ray.put = lambda x: x

    def remote(*a, **kw):
        if a and callable(a[0]) and not kw:
            return RemoteFn(a[0])
        return lambda fn: RemoteFn(fn, kw)

    ray.remote = remote
    ray.get = lambda r: [x.v for x in r] if isinstance(r, list) else r.v
    ray.wait = lambda refs, **kw: (list(refs), [])
    ray.nodes = lambda: [{"NodeID": f"gpu{i}", "Resources": {"CPU": cpus_per_node}} for i in range(num_nodes)]

    _mod("ray.air", REF + "/air")
    _mod("ray.air.util", REF + "/air/util")
    _mod("ray.data", REF + "/data")
    _mod("ray.data._internal", REF + "/data/_internal")
    _mod("ray.data._internal.execution", REF + "/data/_internal/execution")

    class PB:
        def __init__(self, *a, **k):
            pass

        def fetch_until_complete(self, refs):
            return ray.get(list(refs))

        def close(self):
            pass

        def set_description(self, *a):
            pass

        def update(self, *a, **k):
            pass

    _mod("ray.data._internal.progress_bar", ProgressBar=PB)
    cache = {}

    def cached_remote_fn(fn, **kw):
        if fn not in cache:
            cache[fn] = ray.remote(**kw)(fn) if kw else ray.remote(fn)
        return cache[fn]

    _mod("ray.data._internal.remote_fn", cached_remote_fn=cached_remote_fn)
    _mod("ray.data._internal.stats", StatsDict=dict)

    class _L:
        def __init__(self, n):
            pass

        def get_logger(self, **k):
            import logging

            return logging.getLogger("ray.data")

    _mod("ray.data._internal.dataset_logger", DatasetLogger=_L)
    _mod("ray.data._internal.memory_tracing", trace_deallocation=lambda *a, **k: None)
    from typing import Callable

    ei = _mod("ray.data._internal.execution.interfaces", REF + "/data/_internal/execution/interfaces",
              ExecutionOptions=dict, AllToAllTransformFn=Callable)
    rb = importlib.import_module("ray.data._internal.execution.interfaces.ref_bundle")
    tc = importlib.import_module("ray.data._internal.execution.interfaces.task_context")
    ei.RefBundle, ei.TaskContext = rb.RefBundle, tc.TaskContext

    class NodeAffinitySchedulingStrategy:
        def __init__(self, node_id, soft):
            self.node_id, self.soft = node_id, soft

    _mod("ray.util.scheduling_strategies", NodeAffinitySchedulingStrategy=NodeAffinitySchedulingStrategy,
         SchedulingStrategyT=object)
    _mod("ray.data._internal.execution.operators", "")
    _mod("ray.data._internal.execution.operators.map_transformer", MapTransformer=object)
    return ray, rb, tc, PB


def run_all():
    ray, rb, tc, PB = install_shim()
    from ray.data.block import BlockAccessor
    from ray.data.context import DataContext
    from ray.data._internal.planner.exchange.sort_task_spec import SortKey, SortTaskSpec
    from ray.data._internal.planner.exchange.shuffle_task_spec import ShuffleTaskSpec
    from ray.data._internal.planner.exchange.push_based_shuffle_task_scheduler import PushBasedShuffleTaskScheduler
    from ray.data.aggregate import Count, Max, Mean, Min, Std, Sum
    sort_mod = importlib.import_module("ray.data._internal.planner.sort")
    agg_mod = importlib.import_module("ray.data._internal.planner.aggregate")
    shuf_mod = importlib.import_module("ray.data._internal.planner.random_shuffle")

    def bundles(tables):
        return [rb.RefBundle([(Ref(t), BlockAccessor.for_block(t).get_metadata(None, None))], owns_blocks=True)
                for t in tables]

    def ctx():
        return tc.TaskContext(task_idx=0, target_max_block_size=1 << 30,
                              sub_progress_bar_dict={"Sort Sample": PB(), "Shuffle Map": PB(), "Shuffle Reduce": PB()})

    def blocks_of(out_bundles):
        return [ray.get(b.blocks[0][0]) for b in out_bundles]

    captured = {}
    orig_sample = SortTaskSpec.sample_boundaries

    def capture(blocks, sort_key, num_reducers):
        b = orig_sample(blocks, sort_key, num_reducers)
        captured["boundaries"] = list(b)  # copy: generate_sort_fn reverses the list in place for descending
        return b

    SortTaskSpec.sample_boundaries = staticmethod(capture)

    gold = {"pyarrow": pa.__version__, "numpy": np.__version__}
    arrays = {}

    # ---------------------------------------------------------------- sort goldens
    rng = np.random.default_rng(20240601)
    B, n = 8, 2500
    N = B * n
    cases = {
        "i64_ties": rng.integers(-250, 250, N, dtype=np.int64),
        "i64_wide": rng.integers(np.iinfo(np.int64).min, np.iinfo(np.int64).max, N, dtype=np.int64),
        "f64_special": None,
    }
    f = rng.standard_normal(N)
    f[rng.integers(0, N, N // 10)] = 0.0
    f[rng.integers(0, N, N // 10)] = -0.0
    f[rng.integers(0, N, 7)] = np.inf
    f[rng.integers(0, N, 7)] = -np.inf
    f = np.round(f, 2)  # ties
    cases["f64_special"] = f
    row = np.arange(N, dtype=np.int64)
    sort_index = []
    for name, key in cases.items():
        arrays[f"sort_{name}_key"] = key
        tables = [pa.table({"k": key[i * n:(i + 1) * n], "row": row[i * n:(i + 1) * n]}) for i in range(B)]
        for desc in (False, True):
            for push in (False, True):
                DataContext.get_current().use_push_based_shuffle = push
                random.seed(12345)
                out, stats = sort_mod.generate_sort_fn(SortKey("k", descending=desc))(bundles(tables), ctx())
                ob = blocks_of(out)
                tag = f"sort_{name}_{'desc' if desc else 'asc'}_{'push' if push else 'pull'}"
                arrays[tag + "_rows"] = np.concatenate([t["row"].to_numpy() if t.num_rows else np.zeros(0, np.int64) for t in ob])
                arrays[tag + "_sizes"] = np.array([t.num_rows for t in ob], dtype=np.int64)
                bnd = [b[0] for b in captured["boundaries"]]
                arrays[tag + "_bounds"] = np.array(bnd, dtype=key.dtype)
                sort_index.append({"tag": tag, "case": name, "descending": desc, "push": push,
                                   "stats_keys": sorted(stats.keys())})
    gold["sort_cases"] = sort_index
    gold["sort_shape"] = {"blocks": B, "rows_per_block": n}

    # NaN keys: the reference's sampled boundaries degenerate (np.quantile propagates NaN), so pin only
    # the global row order, which is what the contract covers.
    fn = f.copy()
    fn[rng.integers(0, N, N // 20)] = np.nan
    arrays["sort_f64_nan_key"] = fn
    tables = [pa.table({"k": fn[i * n:(i + 1) * n], "row": row[i * n:(i + 1) * n]}) for i in range(B)]
    for desc in (False, True):
        DataContext.get_current().use_push_based_shuffle = False
        random.seed(7)
        out, _ = sort_mod.generate_sort_fn(SortKey("k", descending=desc))(bundles(tables), ctx())
        arrays[f"sort_f64_nan_{'desc' if desc else 'asc'}_rows"] = np.concatenate(
            [t["row"].to_numpy() for t in blocks_of(out) if t.num_rows])

    # user-specified boundaries (the reference's own golden, tests/test_sort.py:23-51)
    spec_cases = []
    ids = np.arange(1000, dtype=np.int64)
    tables = [pa.table({"id": ids[i * 50:(i + 1) * 50]}) for i in range(20)]
    for desc, bounds in [(True, list(range(100, 1000, 200))), (False, list(range(100, 1000, 200))),
                         (True, [1, 998]), (False, [1, 998]), (True, [501.5]), (False, [501.5])]:
        DataContext.get_current().use_push_based_shuffle = False
        out, _ = sort_mod.generate_sort_fn(SortKey("id", desc, bounds))(bundles(tables), ctx())
        ob = blocks_of(out)
        spec_cases.append({"descending": desc, "boundaries": bounds,
                           "blocks": [t["id"].to_pylist() if t.num_rows else [] for t in ob]})
    gold["specified_boundaries"] = spec_cases

    # two-column key
    k1 = rng.integers(0, 40, N, dtype=np.int64)
    k2 = rng.integers(-5, 5, N, dtype=np.int64)
    arrays["sort2_k1"], arrays["sort2_k2"] = k1, k2
    tables = [pa.table({"a": k1[i * n:(i + 1) * n], "b": k2[i * n:(i + 1) * n], "row": row[i * n:(i + 1) * n]}) for i in range(B)]
    for desc in (False, True):
        random.seed(99)
        out, _ = sort_mod.generate_sort_fn(SortKey(["a", "b"], descending=desc))(bundles(tables), ctx())
        arrays[f"sort2_{'desc' if desc else 'asc'}_rows"] = np.concatenate(
            [t["row"].to_numpy() for t in blocks_of(out) if t.num_rows])

    # ---------------------------------------------------------------- block-level map / reduce
    t0 = pa.table({"k": cases["i64_ties"][:n], "row": row[:n]})
    parts = SortTaskSpec.map(0, t0, 4, [(-100,), (0,), (100,)], SortKey("k"))[:-1]
    arrays["map_asc_rows"] = np.concatenate([p["row"].to_numpy() for p in parts])
    arrays["map_asc_sizes"] = np.array([p.num_rows for p in parts])
    parts = SortTaskSpec.map(0, t0, 4, [(100,), (0,), (-100,)], SortKey("k", descending=True))[:-1]
    arrays["map_desc_rows"] = np.concatenate([p["row"].to_numpy() for p in parts])
    arrays["map_desc_sizes"] = np.array([p.num_rows for p in parts])
    empties = BlockAccessor.for_block(pa.Table.from_pydict({})).sort_and_partition([1, 5, 10], SortKey("A"))
    gold["empty_sort_and_partition"] = [p.num_rows for p in empties]

    # ---------------------------------------------------------------- groupby goldens
    G = 200
    gk = rng.integers(0, G, N, dtype=np.int64)
    gv = rng.integers(-2**31, 2**31, N, dtype=np.int64)
    gf = rng.random(N) * 100
    arrays["agg_k"], arrays["agg_v"], arrays["agg_f"] = gk, gv, gf
    tables = [pa.table({"k": gk[i * n:(i + 1) * n], "v": gv[i * n:(i + 1) * n], "f": gf[i * n:(i + 1) * n],
                        "unused": row[i * n:(i + 1) * n]}) for i in range(B)]
    for push in (False, True):
        DataContext.get_current().use_push_based_shuffle = push
        random.seed(4242)
        aggs = [Sum("v"), Mean("f"), Count(), Min("v"), Max("f"), Std("f"), Mean("v"), Sum("f")]
        out, stats = agg_mod.generate_aggregate_fn("k", aggs)(bundles(tables), ctx())
        ob = [t for t in blocks_of(out)]
        tag = f"agg_{'push' if push else 'pull'}"
        full = pa.concat_tables([t for t in ob if t.num_rows])
        gold[tag + "_columns"] = full.column_names
        for c in full.column_names:
            arrays[f"{tag}_{c}"] = full[c].to_numpy()
        arrays[tag + "_sizes"] = np.array([t.num_rows for t in ob])
    # the reference's own small golden: A = x % 3, B = x (tests/test_all_to_all.py:244-317, :508-580)
    xs = list(range(100))
    random.seed(1)
    random.shuffle(xs)
    small = [pa.table({"A": [x % 3 for x in xs[i::4]], "B": xs[i::4]}) for i in range(4)]
    DataContext.get_current().use_push_based_shuffle = False
    out, _ = agg_mod.generate_aggregate_fn("A", [Sum("B"), Mean("B"), Count(), Min("B"), Max("B"), Std("B")])(bundles(small), ctx())
    full = pa.concat_tables([t for t in blocks_of(out) if t.num_rows])
    gold["small_groupby"] = {c: full[c].to_pylist() for c in full.column_names}
    gold["small_groupby_input"] = xs
    out, _ = agg_mod.generate_aggregate_fn(None, [Sum("B"), Mean("B"), Count(), Min("B"), Max("B"), Std("B")])(bundles(small), ctx())
    full = pa.concat_tables([t for t in blocks_of(out) if t.num_rows])
    gold["small_global_agg"] = {c: full[c].to_pylist() for c in full.column_names}
    # duplicate aggregation names get a _2 suffix (arrow_block.py:537-552)
    out, _ = agg_mod.generate_aggregate_fn("A", [Sum("B"), Sum("B"), Count(), Count()])(bundles(small), ctx())
    gold["dup_names_columns"] = pa.concat_tables([t for t in blocks_of(out) if t.num_rows]).column_names

    # ---------------------------------------------------------------- random_shuffle structure
    tables = [pa.table({"row": row[i * n:(i + 1) * n]}) for i in range(B)]
    shuf = {}
    for push in (False, True):
        DataContext.get_current().use_push_based_shuffle = push
        outs = []
        for seed in (0, 0, 1):
            out, stats = shuf_mod.generate_random_shuffle_fn(seed)(bundles(tables), ctx())
            outs.append([t["row"].to_numpy() for t in blocks_of(out)])
        same = all(np.array_equal(a, b) for a, b in zip(outs[0], outs[1]))
        diff = not all(np.array_equal(a, b) for a, b in zip(outs[0], outs[2]))
        allrows = np.sort(np.concatenate(outs[0]))
        shuf["push" if push else "pull"] = {
            "deterministic_per_seed": bool(same), "seed_changes_output": bool(diff),
            "is_permutation": bool(np.array_equal(allrows, row)), "sizes": [len(a) for a in outs[0]],
            "every_output_sees_every_input": bool(all(len(set(a // n)) == B for a in outs[0])),
            "stats_keys": sorted(stats.keys())}
    gold["random_shuffle"] = shuf
    m = ShuffleTaskSpec.map(3, pa.table({"row": row[:1003]}), 8, 1 << 30, None, True, 5)[:-1]
    gold["shuffle_map_slice_sizes_sorted"] = sorted(p.num_rows for p in m)
    m = ShuffleTaskSpec.map(0, pa.table({"row": row[:5]}), 8, 1 << 30, None, True, 5)[:-1]
    gold["shuffle_map_tiny_slice_sizes_sorted"] = sorted(p.num_rows for p in m)

    # ---------------------------------------------------------------- push-based schedules
    sched = []
    combos = [({"n0": 3}, 10, 2, 10), ({"n0": 3, "n1": 3}, 10, 2, 10), ({f"n{i}": 16 for i in range(20)}, 1000, 2, 1000),
              ({"n0": 128}, 260, 2, 260), ({"n0": 128}, 1, 2, 1), ({"n0": 2, "n1": 1}, 4, 2, 7),
              ({f"gpu{i}": 3 for i in range(8)}, 8, 2, 8), ({f"gpu{i}": 3 for i in range(8)}, 64, 2, 64),
              ({"a": 8, "b": 4, "c": 1}, 37, 2, 11), ({"a": 8, "b": 4, "c": 1}, 37, 1.5, 50), ({"a": 1}, 5, 2, 3),
              ({"a": 5, "b": 7}, 100, 3, 13)]
    for cpus, nin, mf, nout in combos:
        st = PushBasedShuffleTaskScheduler._compute_shuffle_schedule(cpus, nin, mf, nout)
        ms = st.merge_schedule
        sched.append({
            "cpus": cpus, "num_input_blocks": nin, "merge_factor": mf, "num_output_blocks": nout,
            "num_rounds": st.num_rounds, "num_map_tasks_per_round": st.num_map_tasks_per_round,
            "placement": [o["scheduling_strategy"].node_id for o in st._merge_task_options],
            "num_merge_tasks_per_round": ms.num_merge_tasks_per_round,
            "num_reducers_per_merger": ms.num_reducers_per_merger,
            "reducers_per_merge_idx": [ms.get_num_reducers_per_merge_idx(i) for i in range(ms.num_merge_tasks_per_round)],
            "merge_idx_for_reducer": [ms.get_merge_idx_for_reducer_idx(j) for j in range(nout)],
            "round_robin": list(ms.round_robin_reduce_idx_iterator()),
            "estimated_num_refs": st.get_estimated_num_refs()})
    gold["push_schedules"] = sched

    # ---------------------------------------------------------------- sample_boundaries on fixed samples
    # (the deterministic half: lexsort + nearest-rank quantiles), by feeding blocks whose size equals n_samples
    sb = []
    for R, nb in [(4, 4), (8, 8), (8, 2), (3, 5)]:
        ns = int(R * 10 / nb)
        vals = rng.integers(0, 1000, nb * ns, dtype=np.int64)
        tbls = [Ref(pa.table({"k": vals[i * ns:(i + 1) * ns]})) for i in range(nb)]
        b = orig_sample(tbls, SortKey("k"), R)
        sb.append({"R": R, "num_blocks": nb, "values": vals.tolist(), "boundaries": [int(x[0]) for x in b]})
    gold["sample_boundaries"] = sb
    gold["sample_boundaries_empty"] = len(orig_sample([Ref(pa.Table.from_pydict({}))] * 3, SortKey("id"), 3))

    np.savez_compressed(os.path.join(OUT, "reference_outputs.npz"), **arrays)
    with open(os.path.join(OUT, "reference_outputs.json"), "w") as fjson:
        json.dump(gold, fjson, indent=1, default=lambda o: o.item() if hasattr(o, "item") else str(o))
    print("wrote", len(arrays), "arrays;", os.path.getsize(os.path.join(OUT, "reference_outputs.npz")) // 1024, "KiB")


def run_quantile(shim=None):
    """Quantile goldens (aggregate/_aggregate.py:360-431) -> quantile_outputs.json (inputs included)."""
    ray, rb, tc, PB = shim or install_shim()
    from ray.data.block import BlockAccessor
    from ray.data.context import DataContext
    from ray.data.aggregate import Count, Quantile, Sum
    agg_mod = importlib.import_module("ray.data._internal.planner.aggregate")

    def bundles(tables):
        return [rb.RefBundle([(Ref(t), BlockAccessor.for_block(t).get_metadata(None, None))], owns_blocks=True)
                for t in tables]

    def ctx():
        return tc.TaskContext(task_idx=0, target_max_block_size=1 << 30,
                              sub_progress_bar_dict={"Sort Sample": PB(), "Shuffle Map": PB(), "Shuffle Reduce": PB()})

    def run(tables, key, aggs, push):
        DataContext.get_current().use_push_based_shuffle = push
        random.seed(777)
        out, _ = agg_mod.generate_aggregate_fn(key, aggs)(bundles(tables), ctx())
        cols = {}
        for b in out:
            t = ray.get(b.blocks[0][0])
            if t.num_rows:
                for c in t.column_names:
                    cols.setdefault(c, []).extend(t[c].to_pylist())
        return cols

    rng = np.random.default_rng(20240602)
    N, B, G = 2000, 4, 37
    k = rng.integers(0, G, N, dtype=np.int64)
    vi = rng.integers(-10**6, 10**6, N, dtype=np.int64)
    vf = np.round(rng.standard_normal(N) * 1000, 3)
    n = N // B
    tables = [pa.table({"k": k[i * n:(i + 1) * n], "vi": vi[i * n:(i + 1) * n], "vf": vf[i * n:(i + 1) * n]})
              for i in range(B)]
    qs = [0.0, 0.1, 0.25, 0.5, 0.9, 0.99, 1.0]
    gold = {"pyarrow": pa.__version__, "numpy": np.__version__, "k": k.tolist(), "vi": vi.tolist(), "vf": vf.tolist(),
            "blocks": B, "qs": qs, "grouped": {}, "global": {}}
    for push in (False, True):
        aggs = [Quantile("vi", q=q, alias_name=f"qi{j}") for j, q in enumerate(qs)]
        aggs += [Quantile("vf", q=q, alias_name=f"qf{j}") for j, q in enumerate(qs)]
        aggs += [Count(), Sum("vi"), Quantile("vf")]
        gold["grouped"]["push" if push else "pull"] = run(tables, "k", aggs, push)
    aggs = [Quantile("vi", q=q, alias_name=f"qi{j}") for j, q in enumerate(qs)]
    aggs += [Quantile("vf", q=q, alias_name=f"qf{j}") for j, q in enumerate(qs)]
    gold["global"] = run(tables, None, aggs, False)
    # the reference's own test data (tests/test_all_to_all.py:720-775): A = x % 3, B = x
    xs = list(range(100))
    random.seed(1)
    random.shuffle(xs)
    small = [pa.table({"A": [x % 3 for x in xs[i::4]], "B": xs[i::4]}) for i in range(4)]
    gold["small_input"] = xs
    gold["small_grouped"] = run(small, "A", [Count(), Quantile("B"), Quantile("B", q=0.9, alias_name="q9")], False)
    gold["small_global"] = run(small, None, [Quantile("B")], False)
    with open(os.path.join(OUT, "quantile_outputs.json"), "w") as fjson:
        json.dump(gold, fjson, default=lambda o: o.item() if hasattr(o, "item") else str(o))
    print("wrote quantile_outputs.json", os.path.getsize(os.path.join(OUT, "quantile_outputs.json")) // 1024, "KiB")


def run_nulls(shim=None):
    """Aggregations over nullable value columns (null policy of _internal/null_aggregate.py:11-17) -> null_outputs.json
    (inputs included): the reference's own generate_aggregate_fn, pull and push, ignore_nulls True and False, grouped
    and global; groups with no null, some nulls, only nulls."""
    ray, rb, tc, PB = shim or install_shim()
    from ray.data.block import BlockAccessor
    from ray.data.context import DataContext
    from ray.data.aggregate import AbsMax, Count, Max, Mean, Min, Std, Sum
    agg_mod = importlib.import_module("ray.data._internal.planner.aggregate")

    def bundles(tables):
        return [rb.RefBundle([(Ref(t), BlockAccessor.for_block(t).get_metadata(None, None))], owns_blocks=True)
                for t in tables]

    def ctx():
        return tc.TaskContext(task_idx=0, target_max_block_size=1 << 30,
                              sub_progress_bar_dict={"Sort Sample": PB(), "Shuffle Map": PB(), "Shuffle Reduce": PB()})

    def run(tables, key, aggs, push):
        DataContext.get_current().use_push_based_shuffle = push
        random.seed(777)
        out, _ = agg_mod.generate_aggregate_fn(key, aggs)(bundles(tables), ctx())
        cols = {}
        for b in out:
            t = ray.get(b.blocks[0][0])
            if t.num_rows:
                for c in t.column_names:
                    cols.setdefault(c, []).extend(t[c].to_pylist())
        return cols

    rng = np.random.default_rng(20240711)
    N, B, G = 1200, 4, 23
    k = rng.integers(0, G, N, dtype=np.int64)
    vi = rng.integers(-1000, 1000, N, dtype=np.int64)
    vf = np.round(rng.standard_normal(N) * 10, 3)
    null_i = rng.random(N) < 0.2
    null_f = rng.random(N) < 0.3
    null_i[k == 3] = True           # a group with only nulls
    null_f[k == 3] = True
    null_i[k == 5] = False          # a group without nulls
    null_f[k == 5] = False
    n = N // B
    tables = [pa.table({"k": k[i * n:(i + 1) * n],
                        "vi": pa.array(vi[i * n:(i + 1) * n], mask=null_i[i * n:(i + 1) * n]),
                        "vf": pa.array(vf[i * n:(i + 1) * n], mask=null_f[i * n:(i + 1) * n]),
                        "w": vi[i * n:(i + 1) * n]}) for i in range(B)]
    gold = {"pyarrow": pa.__version__, "k": k.tolist(), "vi": vi.tolist(), "vf": vf.tolist(), "null_i": null_i.tolist(),
            "null_f": null_f.tolist(), "blocks": B, "grouped": {}, "global": {}}

    def aggs(ignore):
        return [Count(), Sum("vi", ignore_nulls=ignore), Min("vi", ignore_nulls=ignore), Max("vi", ignore_nulls=ignore),
                Mean("vi", ignore_nulls=ignore), Std("vi", ignore_nulls=ignore), AbsMax("vi", ignore_nulls=ignore),
                Sum("vf", ignore_nulls=ignore), Min("vf", ignore_nulls=ignore), Max("vf", ignore_nulls=ignore),
                Mean("vf", ignore_nulls=ignore), Std("vf", ignore_nulls=ignore), Sum("w", ignore_nulls=ignore), Mean("w")]

    for push in (False, True):
        for ignore in (True, False):
            gold["grouped"][f"{'push' if push else 'pull'}_ignore{int(ignore)}"] = run(tables, "k", aggs(ignore), push)
    for ignore in (True, False):
        gold["global"][f"ignore{int(ignore)}"] = run(tables, None, aggs(ignore), False)
    with open(os.path.join(OUT, "null_outputs.json"), "w") as fjson:
        json.dump(gold, fjson, default=lambda o: o.item() if hasattr(o, "item") else str(o))
    print("wrote null_outputs.json", os.path.getsize(os.path.join(OUT, "null_outputs.json")) // 1024, "KiB")


def run_batcher(shim=None):
    """Batcher / ShufflingBatcher goldens (_internal/batcher.py) -> batcher_outputs.json. The classes are
    the reference's own; the driver loop is block_batching/util.py:116-139 restated (that module needs
    ray.actor). random_shuffle is replaced by a recorder that keeps the row order, so that which rows
    are in the buffer at every re-shuffle, the seeds used and the batch boundaries are all pinned."""
    install = shim or install_shim()
    bmod = importlib.import_module("ray.data._internal.batcher")
    from ray.data._internal.arrow_block import ArrowBlockAccessor

    shuffles = []
    orig = ArrowBlockAccessor.random_shuffle

    def recorder(self, random_seed):
        shuffles.append({"seed": random_seed, "rows": self._table["row"].to_pylist()})
        return self._table

    ArrowBlockAccessor.random_shuffle = recorder
    cases = []
    combos = [([5, 0, 12, 1, 22], 7, None, False), ([5, 0, 12, 1, 22], 7, None, True), ([5, 0, 12, 1, 22], None, None, False),
              ([100], 32, None, False), ([3, 3, 3, 3], 10, None, False), ([3, 3, 3, 3], 10, None, True),
              ([50, 50, 50, 50, 50], 16, 40, False), ([50, 50, 50, 50, 50], 16, 40, True), ([10] * 30, 25, 60, False),
              ([500], 64, 100, False), ([7, 0, 9, 200, 1, 1, 1, 64], 32, 10, False), ([20, 20], 50, 5, False)]
    for sizes, batch_size, min_size, drop_last in combos:
        del shuffles[:]
        lo, blocks = 0, []
        for sz in sizes:
            blocks.append(pa.table({"row": np.arange(lo, lo + sz, dtype=np.int64)}))
            lo += sz
        if min_size is not None:
            batcher = bmod.ShufflingBatcher(batch_size=batch_size, shuffle_buffer_min_size=min_size, shuffle_seed=11)
        else:
            batcher = bmod.Batcher(batch_size=batch_size)
        out = []
        for blk in blocks:
            batcher.add(blk)
            while batcher.has_batch():
                out.append(batcher.next_batch()["row"].to_pylist())
        batcher.done_adding()
        while batcher.has_batch():
            out.append(batcher.next_batch()["row"].to_pylist())
        if not drop_last and batcher.has_any():
            out.append(batcher.next_batch()["row"].to_pylist())
        cases.append({"block_sizes": sizes, "batch_size": batch_size, "shuffle_buffer_min_size": min_size,
                      "drop_last": drop_last, "batches": out, "shuffles": [dict(x) for x in shuffles]})
    ArrowBlockAccessor.random_shuffle = orig
    # randomize_block_order (planner/randomize_blocks.py): the reference's own function on 9 one-row blocks
    ray, rb, tc, PB = install
    from ray.data.block import BlockAccessor
    rmod = importlib.import_module("ray.data._internal.planner.randomize_blocks")
    orders = {}
    for seed in (0, 1, 42, 2024):
        op = types.SimpleNamespace(_seed=seed, _name="RandomizeBlockOrder")
        tables = [pa.table({"row": [i]}) for i in range(9)]
        refs = [rb.RefBundle([(Ref(t), BlockAccessor.for_block(t).get_metadata(None, None))], owns_blocks=True)
                for t in tables]
        out, _ = rmod.generate_randomize_blocks_fn(op)(refs, None)
        orders[str(seed)] = [ray.get(b.blocks[0][0])["row"][0].as_py() for b in out]
    with open(os.path.join(OUT, "batcher_outputs.json"), "w") as fjson:
        json.dump({"cases": cases, "randomize_block_order": orders}, fjson)
    print("wrote batcher_outputs.json", os.path.getsize(os.path.join(OUT, "batcher_outputs.json")) // 1024, "KiB")


def run_map_groups(shim=None):
    """GroupedData.map_groups' group-boundary finder (grouped_data.py:181-200) -> map_groups_outputs.json.
    `grouped_data.py` cannot be imported under the stub (it pulls in ray.data.dataset), so the class
    `_MultiColumnSortedKey` (:13-33) and the nested function `get_key_boundaries` are cut out of the reference's
    own source text with `ast` and executed as they stand, on the reference's real ArrowBlockAccessor."""
    import ast
    import textwrap

    install_shim() if shim is None else shim
    from ray.data.block import BlockAccessor

    path = os.path.join(REF, "data", "grouped_data.py")
    src = open(path).read()
    tree = ast.parse(src)
    cls = next(n for n in tree.body if isinstance(n, ast.ClassDef) and n.name == "_MultiColumnSortedKey")
    inner = next(n for n in ast.walk(tree) if isinstance(n, ast.FunctionDef) and n.name == "get_key_boundaries")
    lines = src.splitlines()
    cases = []
    rng = np.random.default_rng(2024)
    for n, keys in ((240, "a"), (240, ["a", "b"]), (300, ["a", "b", "c"]), (1, ["a", "b"]), (60, ["b", "a"])):
        a = rng.integers(-3, 4, n).astype(np.int64)
        b = rng.integers(0, 3, n).astype(np.int64)
        c = np.round(rng.standard_normal(n), 0)
        row = np.arange(n, dtype=np.int64)
        t = pa.table({"a": a, "b": b, "c": c, "row": row})
        klist = [keys] if isinstance(keys, str) else keys
        t = t.take(pa.compute.sort_indices(t, sort_keys=[(k, "ascending") for k in klist]))
        ns = {"List": list, "BlockAccessor": BlockAccessor}
        exec("\n".join(lines[cls.lineno - 1:cls.end_lineno]), ns)
        fake_self = types.SimpleNamespace(_key=keys)
        ns["self"] = fake_self
        exec(textwrap.dedent("\n".join(lines[inner.lineno - 1:inner.end_lineno])), ns)
        bounds = [int(x) for x in ns["get_key_boundaries"](BlockAccessor.for_block(t))]
        cases.append({"keys": keys, "columns": {k: t[k].to_pylist() for k in t.column_names}, "boundaries": bounds})
    with open(os.path.join(OUT, "map_groups_outputs.json"), "w") as fjson:
        json.dump({"cases": cases}, fjson)
    print("wrote map_groups_outputs.json", os.path.getsize(os.path.join(OUT, "map_groups_outputs.json")) // 1024, "KiB")


if __name__ == "__main__":
    if not os.path.isdir(REF):
        sys.exit("needs /root/reference (build container only)")
    if sys.argv[1:] == ["quantile"]:
        run_quantile()
    elif sys.argv[1:] == ["batcher"]:
        run_batcher()
    elif sys.argv[1:] == ["map_groups"]:
        run_map_groups()
    elif sys.argv[1:] == ["nulls"]:
        run_nulls()
    else:
        run_all()
        run_quantile()
        run_batcher()
        run_map_groups()
        run_nulls()
