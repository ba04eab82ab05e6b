"""GPU: melissa_impute_noalloc against melissa_impute, both through the C ABI.

melissa_impute_noalloc is for hosts that cannot allocate device memory when they
want predictions (all GPU memory already owned by another runtime's pool, or a
request handler that must stay out of the CUDA allocator); it stages everything
in one fixed array inside the library instead of cudaMalloc'ed buffers.  Because
that array is shared by all calls and never cleared, the test calls it twice in
a row with different batches: the second result must not be contaminated by what
the first left behind.  Same kernel, same inputs: the outputs must be identical.
This file sorts last on purpose, so a `pytest -x` run has finished every other
test before it starts; it must also pass when run alone in a fresh process.
"""
import ctypes as C

import numpy as np
import pytest

from melissa_b200 import _lib
from conftest import subset

pytestmark = pytest.mark.gpu


def _call(fn, c, W, r, test):
    M, D, K = W.shape
    N = r.shape[0]
    mus = np.ascontiguousarray(c["basis"]["mus"], np.float64)
    Wf = np.asfortranarray(W).ravel(order="F").copy()
    rf = np.asfortranarray(r).ravel(order="F").copy()
    pp = np.ascontiguousarray(test["pair_ptr"], np.int64)
    pc = np.ascontiguousarray(test["pair_cell"], np.int32)
    pr = np.ascontiguousarray(test["pair_region"], np.int32)
    x = np.ascontiguousarray(test["x"], np.float64)
    pred = np.full(x.shape[0], np.nan)
    rc = fn(N, M, K, 1, D - 1, _lib.ptr(mus, _lib.c_dbl), float(c["basis"]["gamma"]), _lib.ptr(Wf, _lib.c_dbl),
            _lib.ptr(rf, _lib.c_dbl), pc.shape[0], _lib.ptr(pp, _lib.c_i64), _lib.ptr(pc, _lib.c_i32),
            _lib.ptr(pr, _lib.c_i32), _lib.ptr(x, _lib.c_dbl), _lib.ptr(pred, _lib.c_dbl), 0)
    return rc, pred


def test_noalloc_imputation_matches_the_allocating_entry_point_twice(c1):
    lib = _lib.load()
    W, r = c1["k2_W"], c1["k2_r_nk"]
    rng = np.random.RandomState(7)
    batches = [subset(c1, rng.rand(c1["x"].shape[0]) < 0.5), subset(c1, rng.rand(c1["x"].shape[0]) < 0.2)]
    for test in batches:                                   # two calls in a row on the shared array
        rc0, want = _call(lib.melissa_impute, c1, W, r, test)
        rc1, got = _call(lib.melissa_impute_noalloc, c1, W, r, test)
        assert rc0 == 0 and rc1 == 0, lib.melissa_last_error()
        assert not np.isnan(got).any()
        np.testing.assert_array_equal(got, want)           # same kernel, same inputs
        assert got.min() >= 0.0 and got.max() <= 1.0


def test_noalloc_leaves_free_device_memory_and_the_current_device_alone(c1):
    """The point of the entry point: between its first call on the device (which has
    already happened or happens in the warm-up below) and any later call, the amount of
    free device memory does not move and the caller's current device stays put."""
    import torch
    lib = _lib.load()
    W, r = c1["k2_W"], c1["k2_r_nk"]
    test = subset(c1, np.arange(c1["x"].shape[0]) % 3 == 0)
    rc, _ = _call(lib.melissa_impute_noalloc, c1, W, r, test)          # warm-up: context, module load, address look-up
    assert rc == 0, lib.melissa_last_error()
    torch.cuda.synchronize()
    dev_before = torch.cuda.current_device()
    free_before, _tot = torch.cuda.mem_get_info(0)
    for _ in range(3):
        rc, got = _call(lib.melissa_impute_noalloc, c1, W, r, test)
        assert rc == 0, lib.melissa_last_error()
    free_after, _tot = torch.cuda.mem_get_info(0)
    assert free_after == free_before
    assert torch.cuda.current_device() == dev_before


@pytest.mark.parametrize("entry", ["melissa_impute", "melissa_impute_noalloc"])
def test_impute_rejects_a_pair_list_that_does_not_start_at_zero(c1, entry):
    """pair_ptr[0] = c > 0 would leave pred[0..c) unwritten -- in the no-alloc variant that
    means handing the PREVIOUS caller's bytes back; a negative value would read before x."""
    lib = _lib.load()
    W, r = c1["k2_W"], c1["k2_r_nk"]
    test = subset(c1, np.ones(c1["x"].shape[0], bool))
    rc, first = _call(getattr(lib, entry), c1, W, r, test)              # leaves its pred in the shared array
    assert rc == 0
    for shift in (3, -3):
        bad = dict(test)
        bad["pair_ptr"] = test["pair_ptr"].copy()
        bad["pair_ptr"][0] = shift
        rc, got = _call(getattr(lib, entry), c1, W, r, bad)
        assert rc == 1 and "pair_ptr[0]" in lib.melissa_last_error().decode()
        assert np.isnan(got).all()                                      # nothing was written to the caller's buffer


def test_noalloc_refuses_a_batch_larger_than_its_fixed_array(c2):
    lib = _lib.load()
    cap = 8 << 20
    test = subset(c2, np.ones(c2["x"].shape[0], bool))     # 399,654 CpGs; doubled, x + pred alone need 12.8 MB
    big = dict(test)
    reps = 2
    big["x"] = np.tile(test["x"], reps)
    big["pair_ptr"] = np.concatenate([test["pair_ptr"], test["pair_ptr"][1:] + test["pair_ptr"][-1]])
    big["pair_cell"] = np.tile(test["pair_cell"], reps)
    big["pair_region"] = np.tile(test["pair_region"], reps)
    rc, _ = _call(lib.melissa_impute_noalloc, c2, c2["W"], c2["r_nk"], big)
    assert rc == 1
    msg = lib.melissa_last_error().decode()
    assert "fixed scratch" in msg and str(cap) in msg
