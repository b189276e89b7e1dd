"""sspp_b200.splines — the reference's Python spline twins (sspp/BSplines.py, sspp/CubicPath.py) as PRODUCT functions over the
C ABI — against the golden outputs recorded by importing the reference modules (tests/golden/spline_golden.npz), the stored
artefact scripts/bspline_params.npy (KAT #1) and the cases of sspp/tests/test_BSplines.py (KAT #2-#4)."""
import numpy as np
import pytest

from sspp_b200 import splines as bs


def _shapes(golden):
    return [tuple(int(x) for x in s) for s in golden["shapes"]]


def test_product_functions_match_reference_outputs(golden):
    th = golden["thetas"]
    for s, (n, d, k) in enumerate(_shapes(golden)):
        t = golden[f"s{s}_knots"]
        assert np.array_equal(bs.knot_vector(n, k), golden[f"s{s}_knot_vector"])
        ctrl, t2 = bs.compute_control_points(golden[f"s{s}_via"], k)
        assert np.array_equal(t2, t)
        np.testing.assert_allclose(ctrl, golden[f"s{s}_ctrl"], rtol=0, atol=1e-12)
        for i in range(n):                                   # array theta and scalar theta, same values as the reference's recursion
            assert np.array_equal(bs.B(th, k, i, t), golden[f"s{s}_B"][:, i])
            assert np.array_equal(bs.dB(th, k, i, t), golden[f"s{s}_dB"][:, i])
        assert bs.B(float(th[7]), k, 1, t) == golden[f"s{s}_B"][7, 1]
        c = golden[f"s{s}_ctrl"]
        np.testing.assert_allclose(bs.bspline(th, t, c, k), golden[f"s{s}_curve"], rtol=0, atol=1e-14)
        np.testing.assert_allclose(bs.bspline_derivative(th, t, c, k), golden[f"s{s}_dcurve"], rtol=0, atol=1e-12)
        np.testing.assert_allclose(bs.bspline(float(th[20]), t, c, k), golden[f"s{s}_curve"][20], rtol=0, atol=1e-14)


def test_cubic_path_matches_reference(golden):
    us = golden["cubic_us"]
    for i, (a, b, c) in enumerate(golden["cubic_triples"]):
        p = bs.CubicPath()
        assert p.plan(a, b, c) is True
        pos, vel, acc = p.evaluate_with_derivatives(us)
        assert np.array_equal(pos, golden["cubic_pos"][i]) and np.array_equal(vel, golden["cubic_vel"][i]) and np.array_equal(acc, golden["cubic_acc"][i])
        assert np.array_equal(p.evaluate(float(us[3])), golden["cubic_pos"][i, 3])
        np.testing.assert_allclose(p.evaluate(0.5), b, atol=1e-14)
        assert np.array_equal(p.evaluate(-3.0), p.evaluate(0.0)) and np.array_equal(p.evaluate(7.0), p.evaluate(1.0))   # u is clipped
    with pytest.raises(RuntimeError):
        bs.CubicPath().evaluate(0.1)


def test_kat1_reproduces_bspline_params(golden):
    """cfg 1 (i): scripts/main_bspline.py:176-209 - CubicPath through start / via / end, 7 samples, fit with k = 2 - reproduces
    the stored scripts/bspline_params.npy from its own end points"""
    t, c, k = golden["kat_knot_vec"], golden["kat_ctr_pts"], int(golden["kat_k"])
    A = np.array([[bs.B(i / 6, k, j, t) for j in range(7)] for i in range(7)])
    A[0, 0] = A[6, 6] = 1.0
    via = A @ c
    path = bs.CubicPath(); path.plan(via[0], via[3], via[6])
    samples = path.evaluate(np.arange(7) / 6)
    ctrl, t2 = bs.compute_control_points(samples, k)
    np.testing.assert_allclose(ctrl, c, atol=1e-13)
    assert np.array_equal(t2, t)


def test_reference_unit_test_cases(golden):
    for n, k in ((5, 2), (7, 1), (6, 3)):                    # test_BSplines.py:49-54
        t = bs.knot_vector(n, k)
        assert len(t) == n + k + 1 and np.all(t[:k] == 0) and np.all(t[-k:] == 1)
    ramp = bs.bspline(np.linspace(0, 1, 25), bs.knot_vector(7, 1), np.arange(7.0), 1)          # :85-94
    np.testing.assert_allclose(ramp, np.linspace(0, 6, 25), atol=1e-6)
    assert np.array_equal(ramp, golden["ramp_curve"])
    const = bs.bspline(np.linspace(0, 0.999, 25), bs.knot_vector(6, 3), np.ones(6), 3)        # :73-82
    np.testing.assert_allclose(const, golden["const_curve"], atol=1e-15)
    with pytest.raises(ValueError):
        bs.knot_vector(2, 3)
    with pytest.raises(ValueError):
        bs.compute_control_points(np.zeros((1, 3)), 2)


@pytest.mark.gpu
def test_batched_evaluation_on_the_gpu(golden):
    rng = np.random.default_rng(0)
    th = np.concatenate((golden["thetas"], rng.uniform(0, 1, 200)))
    for s, (n, d, k) in enumerate(_shapes(golden)):
        t = golden[f"s{s}_knots"]
        cb = rng.normal(size=(300, n, d))
        cb[0] = golden[f"s{s}_ctrl"]
        out = bs.bspline_batch(th, t, cb, k)
        ref = np.einsum("ti,bid->btd", np.array([bs.bspline(x, t, np.eye(n), k) for x in th]), cb)
        np.testing.assert_allclose(out, ref, rtol=1e-5, atol=1e-5)
        np.testing.assert_allclose(out[0, :65], golden[f"s{s}_curve"], rtol=1e-5, atol=1e-5)
        dout = bs.bspline_batch(th, t, cb[:4], k, derivative=True)
        np.testing.assert_allclose(dout[0, :65], golden[f"s{s}_dcurve"], rtol=1e-4, atol=1e-4)


def test_batched_evaluation_fails_loudly_without_a_gpu(golden):
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    from sspp_b200 import capi
    with pytest.raises(capi.SsppError, match="no CUDA device"):
        bs.bspline_batch([0.5], golden["s0_knots"], np.zeros((2, 7, 9)), 2)
