"""ssppcu_bspline_eval_batch_noalloc: the allocation-free twin of ssppcu_bspline_eval_batch (include/sspp_cuda.h).

No counterpart in the reference: the entry point was added for callers that cannot
call cudaMalloc.  Its device staging is a file-scope `static __device__` array of SSPPCU_NOALLOC_MAX_FLOATS floats (4 MiB)
in sspp_b200/csrc/sspp_kernels.cu whose address is looked up by cudaGetSymbolAddress once per device.

Collected last on purpose.  Called through the C ABI (ctypes), compared with the
existing entry point on the same inputs: same kernel, same arithmetic -> the outputs must be bit-identical, and both
within the 1e-5 the spline tests use against the FP64 row table (tests/test_splines_product.py)."""
import ctypes as C

import numpy as np
import pytest

from sspp_b200 import capi
from sspp_b200 import splines as bs

F = C.POINTER(C.c_float)
MAX_FLOATS = 1 << 20


def _call(name, rows, ctrl, device=0):
    """out[n_curves][n_theta][d] through the C ABI entry `name`; returns (rc, out, message)"""
    rows = np.ascontiguousarray(rows, np.float32)
    ctrl = np.ascontiguousarray(ctrl, np.float32)
    out = np.full((ctrl.shape[0], rows.shape[0], ctrl.shape[2]), np.nan, np.float32)
    err = C.create_string_buffer(512)
    rc = getattr(capi.lib, name)(C.c_int(device), rows.ctypes.data_as(F), C.c_int(rows.shape[0]), C.c_int(rows.shape[1]),
                                 ctrl.ctypes.data_as(F), C.c_int(ctrl.shape[0]), C.c_int(ctrl.shape[2]), out.ctypes.data_as(F),
                                 err, C.c_size_t(len(err)))
    return rc, out, err.value.decode()


def _problem(seed, n_curves, n_theta, n, d, k):
    rng = np.random.default_rng(seed)
    t = bs.knot_vector(n, k)
    theta = np.concatenate(([0.0, 1.0, 0.5], rng.uniform(0, 1, n_theta - 3)))
    rows64 = np.array([bs.bspline(x, t, np.eye(n), k) for x in theta])          # [n_theta][n], FP64 (BSplines.py:44-51)
    ctrl = rng.normal(size=(n_curves, n, d))
    return rows64, ctrl


@pytest.mark.gpu
def test_noalloc_entry_matches_the_allocating_one_on_two_calls_in_a_row():
    # two different shapes back to back, the second smaller: it runs on top of the first call's leftovers in the shared
    # static array; then the first again
    shapes = [(300, 265, 10, 7, 3), (17, 40, 5, 4, 2), (300, 265, 10, 7, 3)]
    for call, (B, T, n, d, k) in enumerate(shapes):
        rows64, ctrl = _problem(100 + call % 2, B, T, n, d, k)
        rc1, want, msg1 = _call("ssppcu_bspline_eval_batch", rows64, ctrl)
        assert rc1 == capi.OK, msg1
        rc2, got, msg2 = _call("ssppcu_bspline_eval_batch_noalloc", rows64, ctrl)
        assert rc2 == capi.OK, msg2
        assert not np.isnan(got).any(), "every output element must have been written"
        assert np.array_equal(got, want), f"call {call}: the two entry points differ"
        ref = np.einsum("ti,bid->btd", rows64, ctrl)
        np.testing.assert_allclose(got, ref, rtol=1e-5, atol=1e-5)
    # the Python wrapper takes the same path
    t = bs.knot_vector(10, 3)
    th = np.linspace(0, 1, 33)
    cb = np.random.default_rng(7).normal(size=(9, 10, 7))
    assert np.array_equal(bs.bspline_batch_noalloc(th, t, cb, 3), bs.bspline_batch(th, t, cb, 3))
    assert np.array_equal(bs.bspline_batch_noalloc(th, t, cb, 3, derivative=True), bs.bspline_batch(th, t, cb, 3, derivative=True))


@pytest.mark.gpu
def test_noalloc_entry_at_its_size_limit():
    # the largest batch of this shape that fits: regions are rounded up to 64 floats
    T, n, d = 64, 10, 7
    up = lambda x: (x + 63) // 64 * 64
    B = (MAX_FLOATS - up(T * n)) // (n * d + T * d)
    while up(T * n) + up(B * n * d) + up(B * T * d) > MAX_FLOATS:
        B -= 1
    rows64, ctrl = _problem(3, B, T, n, d, 3)
    rc, got, msg = _call("ssppcu_bspline_eval_batch_noalloc", rows64, ctrl)
    assert rc == capi.OK, msg
    rc, want, msg = _call("ssppcu_bspline_eval_batch", rows64, ctrl)
    assert rc == capi.OK, msg
    assert np.array_equal(got, want)


def test_noalloc_entry_rejects_sizes_beyond_the_static_scratch():
    """refused before any CUDA call, so this runs with or without a GPU"""
    T, n, d = 64, 10, 7
    B = MAX_FLOATS // (n * d + T * d) + 1
    rows = np.zeros((T, n), np.float32)
    ctrl = np.zeros((B, n, d), np.float32)
    rc, _, msg = _call("ssppcu_bspline_eval_batch_noalloc", rows, ctrl)
    assert rc == capi.ELIMIT
    assert "static device scratch" in msg and str(MAX_FLOATS) in msg
    rc, _, msg = _call("ssppcu_bspline_eval_batch_noalloc", np.zeros((600, 100), np.float32), np.zeros((1, 100, 1), np.float32))
    assert rc == capi.ELIMIT and "shared memory" in msg
    with pytest.raises(capi.SsppError) as ei:
        bs.bspline_batch_noalloc(np.linspace(0, 1, T), bs.knot_vector(n, 3), np.zeros((B, n, d)), 3)
    assert ei.value.code == capi.ELIMIT
    err = C.create_string_buffer(64)
    assert capi.lib.ssppcu_bspline_eval_batch_noalloc(0, None, 1, 1, None, 1, 1, None, err, C.c_size_t(64)) == capi.EINVAL


@pytest.mark.gpu
def test_noalloc_entry_on_a_second_device():
    """every GPU holds its own copy of the static array and its address is looked up with that device current"""
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    rows64, ctrl = _problem(9, 50, 40, 8, 3, 3)
    for dev in (1, 0, 1):
        rc, want, msg = _call("ssppcu_bspline_eval_batch", rows64, ctrl, device=dev)
        assert rc == capi.OK, msg
        rc, got, msg = _call("ssppcu_bspline_eval_batch_noalloc", rows64, ctrl, device=dev)
        assert rc == capi.OK, msg
        assert np.array_equal(got, want)
    rc, _, msg = _call("ssppcu_bspline_eval_batch_noalloc", rows64, ctrl, device=torch.cuda.device_count())
    assert rc == capi.EINVAL and "device ordinal" in msg
