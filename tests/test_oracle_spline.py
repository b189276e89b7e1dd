"""Pins the spline oracles against the reference's own artefacts:
golden outputs produced by importing sspp/BSplines.py and sspp/CubicPath.py, the stored
scripts/bspline_params.npy (KAT #1) and the cases of sspp/tests/test_BSplines.py (KAT #2-#4)."""
import numpy as np
import pytest

from oracle import bsplines_np as bn
from tests.helpers import FIX


@pytest.fixture(scope="module")
def o1():
    from sspp_b200 import capi
    from oracle import pyoracle
    return pyoracle.O1(capi.Model(path=str(FIX / "robocrane7.xml")), 0, 7)


def _shapes(golden):
    return [tuple(int(x) for x in s) for s in golden["shapes"]]


def test_restatement_matches_reference_outputs(golden):
    th = golden["thetas"]
    for s, (n, d, k) in enumerate(_shapes(golden)):
        t = golden[f"s{s}_knots"]
        assert np.array_equal(bn.uniform_knots(n, k), golden[f"s{s}_knot_vector"])
        ctrl, t2 = bn.fit_control_points(golden[f"s{s}_via"], k)
        assert np.array_equal(t2, t)
        np.testing.assert_allclose(ctrl, golden[f"s{s}_ctrl"], rtol=0, atol=1e-12)
        B = np.array([bn.basis(x, k, t) for x in th])
        assert np.array_equal(B, golden[f"s{s}_B"])          # same two-term blend per entry: bit exact
        dB = np.array([bn.basis_deriv(x, k, t) for x in th])
        assert np.array_equal(dB, golden[f"s{s}_dB"])
        c = golden[f"s{s}_ctrl"]
        cur = np.array([bn.curve(x, t, c, k) for x in th])
        np.testing.assert_allclose(cur, golden[f"s{s}_curve"], rtol=0, atol=1e-14)
        dcur = np.array([bn.curve_deriv(x, t, c, k) for x in th])
        np.testing.assert_allclose(dcur, golden[f"s{s}_dcurve"], rtol=0, atol=1e-12)


def test_cubic_path_matches_reference(golden):
    for i, (a, b, c) in enumerate(golden["cubic_triples"]):
        coef = bn.cubic_plan(a, b, c)
        for j, u in enumerate(golden["cubic_us"]):
            p, v, acc = bn.cubic_eval(coef, u)
            assert np.array_equal(p, golden["cubic_pos"][i, j])
            assert np.array_equal(v, golden["cubic_vel"][i, j])
            assert np.array_equal(acc, golden["cubic_acc"][i, j])
        np.testing.assert_allclose(bn.cubic_eval(coef, 0.5)[0], b, atol=1e-14)   # passes through via at u = 1/2


def test_kat1_bspline_params(golden):
    """scripts/bspline_params.npy: 7 via points on the CubicPath through start/via/end, fitted with k = 2."""
    t, c, k = golden["kat_knot_vec"], golden["kat_ctr_pts"], int(golden["kat_k"])
    assert k == 2 and c.shape == (7, 9) and np.array_equal(t, bn.uniform_knots(7, 2))
    A = bn.fit_matrix(7, 2)
    via = A @ c
    assert np.allclose(via[:, 7:], 0)                      # passive joints padded with zeros (main_bspline.py:188)
    coef = bn.cubic_plan(via[0], via[3], via[6])           # start, via(u=1/2), end
    for i in range(7):
        np.testing.assert_allclose(bn.cubic_eval(coef, i / 6)[0], via[i], atol=1e-13)
    refit, _ = bn.fit_control_points(via, 2)
    np.testing.assert_allclose(refit, c, atol=1e-13)
    from tests.helpers import Q_START, Q_END
    np.testing.assert_allclose(c[0, :7], Q_START, atol=5e-5)   # the configurations the 7-DoF cases use
    np.testing.assert_allclose(c[6, :7], Q_END, atol=5e-5)


def test_kat_reference_unit_tests(golden):
    """sspp/tests/test_BSplines.py:49-54,73-94"""
    for n, k in ((5, 2), (7, 1), (6, 3)):
        t = bn.uniform_knots(n, k)
        assert len(t) == n + k + 1 and np.all(t[:k] == 0) and np.all(t[-k:] == 1)
    t1 = bn.uniform_knots(7, 1)
    ramp = np.array([bn.curve(x, t1, np.arange(7.0), 1) for x in np.linspace(0, 1, 25)])
    np.testing.assert_allclose(ramp, np.linspace(0, 6, 25), atol=1e-6)
    assert np.array_equal(ramp, golden["ramp_curve"])
    t3 = bn.uniform_knots(6, 3)
    const = np.array([bn.curve(x, t3, np.ones(6), 3) for x in np.linspace(0, 0.999, 25)])
    np.testing.assert_allclose(const, 1.0, atol=1e-6)
    np.testing.assert_allclose(const, golden["const_curve"], atol=1e-15)


def test_o1_cox_de_boor_matches_reference_curve(golden, o1):
    """O1's C++ basis on explicit knots == the reference's bspline() for theta in [0, 1]."""
    th = golden["thetas"]
    for s, (n, d, k) in enumerate(_shapes(golden)):
        t, c = golden[f"s{s}_knots"], golden[f"s{s}_ctrl"]
        for j, x in enumerate(th):
            if x < 0 or x > 1:
                continue
            np.testing.assert_allclose(o1.spline_eval_knots(c, k, t, x), golden[f"s{s}_curve"][j], rtol=0, atol=1e-13)
            if x < 1:
                np.testing.assert_allclose(o1.basis_row(n, k, t, x), golden[f"s{s}_B"][j], rtol=0, atol=1e-15)


def test_eigen_semantics_interpolation(o1):
    """KnotAveraging + Interpolate as Eigen defines them (SURVEY 8a row A3'): known knot vectors, the
    n=3,p=2 collocation matrix, and the interpolation property."""
    np.testing.assert_allclose(o1.knots(5, 2), [0, 0, 0, .375, .625, 1, 1, 1], atol=1e-15)
    k10 = o1.knots(10, 3)
    np.testing.assert_allclose(k10[4:10], np.arange(2, 8) / 9, atol=1e-15)
    np.testing.assert_allclose(o1.basis_row(3, 2, o1.knots(3, 2), 0.5), [0.25, 0.5, 0.25], atol=1e-15)
    rng = np.random.default_rng(3)
    for K, p, D in ((10, 3, 7), (3, 2, 4), (5, 2, 4), (7, 3, 7), (4, 3, 3)):
        V = rng.normal(size=(K, D))
        ctrl = o1.interpolate(V, p)
        for i in range(K):
            np.testing.assert_allclose(o1.spline_eval(ctrl, p, i / (K - 1)), V[i], atol=1e-12)
        np.testing.assert_allclose(ctrl[0], V[0], atol=1e-14)
        np.testing.assert_allclose(ctrl[-1], V[-1], atol=1e-14)


def test_product_spline_algebra_matches_o1(o1):
    """the host algebra behind ssppcu_create (collocation matrices, Ainv) against the independent O1 statement"""
    from sspp_b200 import capi
    rng = np.random.default_rng(4)
    for K, p in ((10, 3), (3, 2), (5, 2), (7, 3), (16, 3), (4, 3), (2, 1)):
        knots, ainv = capi.spline_setup(K, p)
        np.testing.assert_allclose(knots, o1.knots(K, p), atol=1e-15)
        V = rng.normal(size=(K, 3))
        np.testing.assert_allclose(ainv @ V, o1.interpolate(V, p), atol=1e-11)
        grid = np.concatenate((np.linspace(0, 1, 33), rng.uniform(0, 1, 10)))
        psi_c = capi.collocation_matrix(K, p, grid, False)
        psi_v = capi.collocation_matrix(K, p, grid, True)
        np.testing.assert_allclose(psi_c.sum(1), 1.0, atol=1e-13)        # partition of unity incl. u = 1
        assert ((psi_c != 0).sum(1) <= p + 1).all()
        ctrl = o1.interpolate(V, p)
        for m, u in enumerate(grid):
            np.testing.assert_allclose(psi_c[m] @ ctrl, o1.spline_eval(ctrl, p, u), atol=1e-12)
            np.testing.assert_allclose(psi_v[m] @ V, o1.spline_eval(ctrl, p, u), atol=1e-11)
    with pytest.raises(capi.SsppError):
        capi.spline_setup(2, 3)
