"""GPU test of b200x_radix_argsort_static, the sort entry point for callers that cannot allocate.

Some users of libb200x.so cannot give the sort a scratch buffer: they run inside a process whose
memory pool already owns the GPU, on a stream loop that must not hit cudaMalloc's implicit sync,
or behind an FFI shim without a device allocator. For them the scratch is a fixed array inside
the library. This test calls that entry point directly through the C ABI (ctypes), twice in a
row on different inputs, and requires bit-identical results to b200x_radix_argsort with a
caller-provided temp buffer; the second call is smaller than the first so that stale contents
left in the shared array by the first call would show. It also checks the size limit error.

The file name sorts last so that `pytest -x` has run every other test before this one.
"""
import ctypes

import numpy as np
import pytest
import torch

from ray_project_ray_b200 import _lib

pytestmark = pytest.mark.gpu

MAX_ROWS = 1 << 20


def _sort_with_temp(keys, descending, dev):
    n = keys.numel()
    perm = torch.empty(n, dtype=torch.int32, device=dev)
    codes = torch.empty(n, dtype=torch.int64, device=dev)
    skeys = torch.empty(n, dtype=keys.dtype, device=dev)
    st = torch.cuda.current_stream(dev).cuda_stream
    need = _lib.query_temp("b200x_radix_argsort", keys.data_ptr(), 0, int(descending), None, perm.data_ptr(),
                           codes.data_ptr(), skeys.data_ptr(), n, 0, 0, stream=st)
    temp = torch.empty(max(need, 256), dtype=torch.uint8, device=dev)
    nb = ctypes.c_size_t(temp.numel())
    _lib.call("b200x_radix_argsort", keys.data_ptr(), 0, int(descending), None, perm.data_ptr(), codes.data_ptr(),
              skeys.data_ptr(), n, 0, 0, temp.data_ptr(), ctypes.byref(nb), st)
    return perm, codes, skeys


def _sort_static(keys, descending, dev):
    n = keys.numel()
    # the caller owns inputs and outputs; the entry point itself allocates nothing
    perm = torch.empty(n, dtype=torch.int32, device=dev)
    codes = torch.empty(n, dtype=torch.int64, device=dev)
    skeys = torch.empty(n, dtype=keys.dtype, device=dev)
    st = torch.cuda.current_stream(dev).cuda_stream
    _lib.call("b200x_radix_argsort_static", keys.data_ptr(), 0, int(descending), None, perm.data_ptr(),
              codes.data_ptr(), skeys.data_ptr(), n, st)
    return perm, codes, skeys


def test_static_scratch_sort_matches_the_temp_buffer_entry_point(cuda_device):
    dev = cuda_device
    rng = np.random.default_rng(2024)
    # call 1: the largest accepted input, full-range keys (all eight digit passes run)
    # call 2: a smaller input with few distinct keys and descending order (passes are skipped, ties matter)
    inputs = [(rng.integers(np.iinfo(np.int64).min, np.iinfo(np.int64).max, MAX_ROWS, dtype=np.int64), False),
              (rng.integers(-300, 300, 300_001, dtype=np.int64), True)]
    with torch.cuda.device(dev):
        for host_keys, descending in inputs:
            keys = torch.from_numpy(host_keys).to(dev)
            got = _sort_static(keys, descending, dev)
            want = _sort_with_temp(keys, descending, dev)
            torch.cuda.synchronize()
            for g, w in zip(got, want):
                assert torch.equal(g, w)  # integer outputs: exact
            # and it really is the stable sort
            order = np.argsort(-host_keys if descending else host_keys, kind="stable")
            np.testing.assert_array_equal(got[0].cpu().numpy().view(np.uint32), order.astype(np.uint32))
            np.testing.assert_array_equal(got[2].cpu().numpy(), host_keys[order])


def test_static_scratch_sort_rejects_inputs_beyond_its_limit(cuda_device):
    dev = cuda_device
    with torch.cuda.device(dev):
        keys = torch.zeros(MAX_ROWS + 1, dtype=torch.int64, device=dev)
        perm = torch.empty(MAX_ROWS + 1, dtype=torch.int32, device=dev)
        with pytest.raises(_lib.B200XError) as err:
            _lib.call("b200x_radix_argsort_static", keys.data_ptr(), 0, 0, None, perm.data_ptr(), None, None,
                      MAX_ROWS + 1, torch.cuda.current_stream(dev).cuda_stream)
        assert err.value.code == 2 and "1048576" in err.value.message
