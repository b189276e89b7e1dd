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


@pytest.mark.skipif(not __import__("pathlib").Path("/root/reference/sspp/BSplines.py").exists(),
                    reason="the reference checkout is only present in the development container")
def test_random_inputs_bit_identical_to_the_live_reference_module():
    """sspp/BSplines.py and sspp/CubicPath.py imported from the reference checkout (casadi / matplotlib stubbed: they are only used
    by its plotting and casadi variants) against the product functions on random orders, sizes, control points and parameters,
    inside and outside [0, 1]: every value must carry the same bits (fit: 1e-12, it goes through a linear solve)"""
    import importlib
    import sys
    import types
    saved = {n: sys.modules.get(n) for n in ("casadi", "matplotlib", "matplotlib.pyplot", "sspp", "sspp.BSplines", "sspp.CubicPath")}
    for name in ("casadi", "matplotlib", "matplotlib.pyplot"):
        sys.modules.setdefault(name, types.ModuleType(name))
    sys.path.insert(0, "/root/reference")
    try:
        ref = importlib.import_module("sspp.BSplines")
        refc = importlib.import_module("sspp.CubicPath")
        rng = np.random.default_rng(0)
        for trial in range(150):
            k = int(rng.integers(1, 5)); n = int(rng.integers(max(2, k + 1), 14)); d = int(rng.integers(1, 8))
            t_ref, t = ref.knot_vector(n, k), bs.knot_vector(n, k)
            assert np.array_equal(t_ref, t)
            c = rng.normal(size=(n, d))
            for theta in list(rng.uniform(-0.1, 1.1, 5)) + [0.0, 1.0, 0.5]:
                assert np.array_equal(np.asarray(ref.bspline(theta, t_ref, c, k)), np.asarray(bs.bspline(theta, t, c, k)))
                assert np.array_equal(np.asarray(ref.bspline_derivative(theta, t_ref, c, k)), np.asarray(bs.bspline_derivative(theta, t, c, k)))
                i = int(rng.integers(0, n))
                assert ref.B(theta, k, i, t_ref) == bs.B(theta, k, i, t) and ref.dB(theta, k, i, t_ref) == bs.dB(theta, k, i, t)
            via = rng.normal(size=(n, d))
            cr, tr = ref.compute_control_points(via, k)
            cp, tp = bs.compute_control_points(via, k)
            assert np.allclose(cr, cp, rtol=0, atol=1e-12) and np.array_equal(tr, tp)
            s, v, e = rng.normal(size=(3, d))
            pr, pp = refc.CubicPath(), bs.CubicPath()
            pr.plan(s, v, e); pp.plan(s, v, e)
            for u in rng.uniform(-0.2, 1.2, 4):
                for fn in ("evaluate", "evaluate_derivative", "evaluate_second_derivative"):
                    if hasattr(pr, fn):
                        assert np.array_equal(np.asarray(getattr(pr, fn)(u)), np.asarray(getattr(pp, fn)(u))), fn
    finally:
        sys.path.remove("/root/reference")
        for n, m in saved.items():
            if m is None:
                sys.modules.pop(n, None)
            else:
                sys.modules[n] = m
