"""Loader of the committed reference outputs (tests/golden/, made by make_golden.py)."""
import json
import os

import numpy as np

_DIR = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
_cache = {}


def arrays():
    if "a" not in _cache:
        _cache["a"] = dict(np.load(os.path.join(_DIR, "reference_outputs.npz")))
    return _cache["a"]


def meta():
    if "m" not in _cache:
        with open(os.path.join(_DIR, "reference_outputs.json")) as f:
            _cache["m"] = json.load(f)
    return _cache["m"]


def sort_case_blocks(case):
    """Input blocks (dict of numpy) of a sort golden case."""
    a, m = arrays(), meta()
    B, n = m["sort_shape"]["blocks"], m["sort_shape"]["rows_per_block"]
    key = a[f"sort_{case}_key"]
    row = np.arange(B * n, dtype=np.int64)
    return [{"k": key[i * n:(i + 1) * n], "row": row[i * n:(i + 1) * n]} for i in range(B)]


def split_by_sizes(flat, sizes):
    out, lo = [], 0
    for s in sizes:
        out.append(flat[lo:lo + int(s)])
        lo += int(s)
    return out


def quantile():
    """Outputs of the reference's Quantile aggregation (golden/quantile_outputs.json, inputs included)."""
    if "q" not in _cache:
        with open(os.path.join(_DIR, "quantile_outputs.json")) as f:
            _cache["q"] = json.load(f)
    return _cache["q"]


def batcher():
    """Batch boundaries and shuffle-buffer contents of the reference's Batcher / ShufflingBatcher."""
    if "b" not in _cache:
        with open(os.path.join(_DIR, "batcher_outputs.json")) as f:
            _cache["b"] = json.load(f)
    return _cache["b"]
