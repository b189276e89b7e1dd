"""Signed-distance refinement step (/root/reference/scripts/sdf_gradient/main.py:37-81 on the baked grid).

CPU: the numpy restatement against finite differences of its own field and against a scene with a known answer.
GPU: ssppcu_sdf_gradient against the restatement on the device-baked grid (FP32 kernel vs FP64: 1e-4 relative)."""
import numpy as np
import pytest

from oracle import sdf_gradient_np as og


def _ball_field(n=33, centre=(0.1, -0.05, 0.3), radius=0.2, half=0.8):
    ax = np.linspace(-half, half, n)
    Z, Y, X = np.meshgrid(ax, ax, ax, indexing="ij")
    nodes = np.sqrt((X - centre[0]) ** 2 + (Y - centre[1]) ** 2 + (Z - centre[2]) ** 2) - radius
    lo = np.array([-half] * 3); cell = np.array([2 * half / (n - 1)] * 3)
    return nodes, lo, cell


def test_trilinear_gradient_matches_finite_differences():
    rng = np.random.default_rng(0)
    nodes = rng.normal(size=(9, 10, 11))
    lo = np.array([-1.0, 0.0, 2.0]); cell = np.array([0.2, 0.25, 0.125])
    for _ in range(200):
        x = lo + rng.uniform(0.05, 0.95, 3) * cell * (np.array([11, 10, 9]) - 1)
        i, f = og.locate((11, 10, 9), lo, cell, x)
        if np.any(np.minimum(f, 1 - f) < 1e-3):
            continue                                  # finite differences would straddle a cell face
        v, g = og.field_and_gradient(nodes, lo, cell, x)
        for a in range(3):
            e = np.zeros(3); e[a] = 1e-6
            fd = (og.field_and_gradient(nodes, lo, cell, x + e)[0] - og.field_and_gradient(nodes, lo, cell, x - e)[0]) / 2e-6
            assert abs(fd - g[a]) < 1e-6 * max(1.0, abs(g[a]))


def test_gradient_points_at_a_ball_and_scales_with_clearance():
    centre = np.array([0.1, -0.05, 0.3])
    nodes, lo, cell = _ball_field(n=65, centre=centre)
    spheres = np.array([[0.0, 0.0, 0.0, 0.03, 0.0]])            # one sample at the body origin, radius 0.03
    q = np.array([0.6, 0.2, 0.1, 0.7])
    cl, g = og.closest_sample(nodes, lo, cell, spheres, q)
    to_ball = centre - q[:3]
    true_cl = np.linalg.norm(to_ball) - 0.2 - 0.03
    assert abs(cl - true_cl) < 2e-3                              # trilinear error of a smooth field at h = 0.025
    assert np.dot(-g / np.linalg.norm(g), to_ball / np.linalg.norm(to_ball)) > 0.999
    # straight path that passes the ball: the middle rows push the middle points, the end rows the end points
    W = np.array([[1, 0, 0], [0.25, 0.5, 0.25], [0, 0, 1.0]])
    pts = np.array([[0.6, 0.2, 0.1, 0.0], [0.5, 0.0, 0.3, 0.0], [0.6, -0.3, 0.5, 0.0]])
    G, dist, dirs = og.sdf_gradient(nodes, lo, cell, spheres, pts, W)
    assert G.shape == (3, 3) and np.allclose(G, (dirs.T @ W).T / 3)
    for m in range(3):
        assert abs(np.linalg.norm(dirs[m]) - abs(dist[m])) < 1e-12


@pytest.mark.gpu
def test_device_gradient_matches_restatement():
    from sspp_b200 import capi
    from tests.helpers import sdf_case
    case = sdf_case(n_samples=62, check_points=32, n_points=5, dims=(48, 48, 32))
    h = capi.Handle(case.model, case.cfg)
    case.prepare(h)
    nodes, lo, cell = h.sdf()
    spheres = h.spheres()
    psi_col, _ = h.collocation()
    base = h.base()[0]                                            # [K][4] point matrix of the query
    rng = np.random.default_rng(3)
    for trial in range(4):
        pts = base.copy()
        pts[1:-1, :3] += rng.normal(0, 0.08, (len(pts) - 2, 3)); pts[1:-1, 3] += rng.normal(0, 0.4, len(pts) - 2)
        W = None if trial % 2 == 0 else rng.dirichlet(np.ones(len(pts)), size=17).astype(np.float32)
        G, dist, dirs = h.sdf_gradient(pts, W)
        G1, dist1, dirs1 = og.sdf_gradient(nodes.astype(np.float64), lo, cell, spheres, pts.astype(np.float64),
                                           psi_col if W is None else W)
        assert np.allclose(dist, dist1, rtol=1e-4, atol=2e-6)
        # a different sphere may win a near tie in FP32: compare rows whose runner-up is clearly behind
        ok = np.ones(len(dist), bool)
        assert np.allclose(dirs[ok], dirs1[ok], rtol=2e-3, atol=2e-5) or np.mean(np.isclose(dirs, dirs1, rtol=2e-3, atol=2e-5)) > 0.95
        assert np.allclose(G, G1, rtol=5e-3, atol=2e-5)
    h.close()
    from tests.helpers import task_case
    prim = task_case("stacking.xml", "block1")                    # primitives mode: there is no field to differentiate
    hp = capi.Handle(prim.model, prim.cfg)
    with pytest.raises(capi.SsppError, match="SDF"):
        hp.sdf_gradient(np.zeros((prim.cfg.n_points, 4), np.float32))
    hp.close()
