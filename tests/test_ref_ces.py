"""The oracles against the REFERENCE'S OWN CODE: oracle/_ref/libref.so holds the reference's hot-path headers compiled
unmodified (oracle/ref_driver.cpp; Eigen / MuJoCo / <random> replaced by the stand-ins of oracle/shim/).

CES half, piece by piece (reference file:line under /root/reference/include/sspp/):
  tsp_elites.h:13-32        select (k = max(1, int(n * frac)), partial_sort) + CES log-weights
  tsp_distribution.h:16-83  reset, update (variance about the NEW mean, yaw wrap, clamp order), adapt
  tsp_sampler.h:12-51       truncated normal with the 100-try uniform fallback, yaw wrap, z floor
  tsp_path_model.h:21-43    initLinear / fromVias
  tsp_evaluator.h:10-14     floorPenalty
  Collision.h:84-103        collision_point_cost read-out over the contacts
O1 must agree with the reference's code to 1e-12 (both FP64, same Philox normals); the FP32 replica O2 — and through it
the GPU — must pick the same elites."""
import numpy as np
import pytest

from oracle import pyoracle, pyref
from oracle.flatmodel import FlatModel
from tests.helpers import ICRA_BODY, SEED

pytestmark = pytest.mark.skipif(not pyref.available(), reason="oracle/_ref/libref.so is not built (needs /root/reference)")

LO = [0.0, -0.7, 0.1, -1.6]
HI = [0.7, 0.7, 0.6, 1.6]


def o1cfg(Kv=1, M=40, lo=LO, hi=HI, **kw):
    c = pyoracle.O1Cfg()
    c.mode, c.D, c.K, c.p, c.M, c.n_free, c.free_begin = 1, 4, Kv + 2, 2, M, Kv, 1
    vals = dict(w_collision=1.0, eval_zmin=0.0, floor_margin=0.01, floor_scale=10.0, limit_scale=0.0, sample_zmin=0.1, dist_zmin=0.2,
                elite_frac=0.3, inc=1.5, dec=0.9, sigma_min=1e-4, sigma_max=0.5, sigma_floor=0.005, var_beta=0.2, mean_lr=0.5)
    vals.update(kw)
    for k, v in vals.items():
        setattr(c, k, v)
    for d in range(4):
        c.lo[d], c.hi[d] = lo[d], hi[d]
    return c


@pytest.fixture(scope="module")
def o1():
    return pyoracle.O1(FlatModel.load("planner"), 1, 4, "block1")


def test_elites_and_weights_match_reference(o1):
    rng = np.random.default_rng(1)
    trials = 0
    for n in (1, 2, 3, 4, 5, 7, 10, 17, 33, 100, 1000, 4096):
        for rep in range(100 if n <= 100 else 12):
            frac = 0.3 if rep % 3 == 0 else float(rng.uniform(0.02, 1.0))
            cost = rng.uniform(0.5, 3.0, n)
            # successes interleaved with failures: O1 works on the whole candidate array, the reference on successes_
            n_all = n + int(rng.integers(0, n + 3))
            pos = np.sort(rng.choice(n_all, n, replace=False))
            full_cost = rng.uniform(0.0, 5.0, n_all); status = np.ones(n_all, np.int32)
            full_cost[pos] = cost; status[pos] = 0
            ri, rw = pyref.elites(cost, frac)
            oi, ow, ns = o1.elites(full_cost, status, frac)
            assert ns == n and len(oi) == len(ri) == max(1, int(n * frac))
            assert np.array_equal(pos[ri], oi)
            np.testing.assert_allclose(ow, rw, rtol=0, atol=1e-15)
            trials += 1
    assert trials >= 1000


def test_elites_with_ties_select_the_same_costs(o1):
    """partial_sort is unstable on ties (tsp_elites.h:18-19); the build breaks ties by candidate index.  Both must pick the
    same multiset of costs, and identical indices wherever the costs are distinct."""
    rng = np.random.default_rng(2)
    for rep in range(200):
        n = int(rng.integers(2, 60))
        cost = rng.integers(0, 6, n).astype(np.float64) * 0.25
        ri, rw = pyref.elites(cost, 0.5)
        oi, ow, _ = o1.elites(cost, np.zeros(n, np.int32), 0.5)
        assert np.array_equal(np.sort(cost[ri]), np.sort(cost[oi]))
        assert np.array_equal(cost[ri], cost[oi])            # both lists are in ascending cost order
        np.testing.assert_allclose(ow, rw, atol=1e-15)


@pytest.mark.parametrize("yaw_fixed", [False, True])
def test_distribution_update_matches_reference(o1, yaw_fixed):
    rng = np.random.default_rng(3 + yaw_fixed)
    lo, hi = list(LO), list(HI)
    if yaw_fixed:
        lo[3] = hi[3] = 0.4
    for rep in range(700):
        Kv = int(rng.integers(1, 4))
        kw = dict(mean_lr=float(rng.uniform(0.1, 1.0)), var_beta=float(rng.uniform(0.05, 0.9)), sigma_floor=float(rng.choice([0.0, 0.005, 0.05])),
                  sigma_min=float(rng.choice([1e-4, 1e-2])), sigma_max=float(rng.choice([0.5, 0.2])), dist_zmin=float(rng.choice([0.0, 0.2, 0.3])),
                  inc=1.5, dec=float(rng.choice([0.9, 0.95])))
        oc = o1cfg(Kv, lo=lo, hi=hi, **kw)
        rc = pyref.dist_cfg(kw["sigma_floor"], kw["var_beta"], kw["mean_lr"], kw["sigma_min"], kw["sigma_max"], kw["dist_zmin"], lo, hi)
        mean = rng.uniform(lo, hi, (Kv, 4)); sigma = rng.uniform(1e-3, 0.5, (Kv, 4))
        n = int(rng.integers(0, 4)) if rep % 5 == 0 else int(rng.integers(1, 40))     # n_success in {0,1,2,3} every fifth round
        vias = rng.uniform(np.array(lo) - 0.05, np.array(hi) + 0.05, (max(n, 1), Kv, 4))
        if not yaw_fixed and n:                                                    # yaw values on both sides of +-range/2 of the mean
            vias[:, :, 3] = np.where(rng.random((n, Kv)) < 0.5, mean[None, :, 3] + rng.choice([-1, 1], (n, Kv)) * rng.uniform(1.55, 1.65, (n, Kv)),
                                     vias[:n, :, 3])
        cost = rng.uniform(0.5, 3.0, n)
        if n:
            ei, w = pyref.elites(cost, 0.3)
            oi, ow, _ = o1.elites(cost, np.zeros(n, np.int32), 0.3)
            assert np.array_equal(ei, oi)
        else:
            ei, w = np.zeros(0, np.int64), np.zeros(0)
        rm, rs = pyref.dist_update(rc, mean, sigma, vias[:n], ei, w, kw["inc"], kw["dec"])
        om, os_ = o1.update(oc, mean, sigma, vias[ei].reshape(len(ei), Kv * 4), w)
        np.testing.assert_allclose(om.reshape(Kv, 4), rm, rtol=0, atol=1e-12)
        np.testing.assert_allclose(os_.reshape(Kv, 4), rs, rtol=0, atol=1e-12)
        if n == 0:
            assert np.array_equal(rm, mean)                                        # adapt(false) leaves the mean alone


def test_wrap_angle_diff_matches_reference():
    rng = np.random.default_rng(4)
    for a, b in rng.uniform(-4, 4, (500, 2)):
        r = pyref.wrap_angle_diff(a, b, -1.6, 1.6)
        d = a - b
        while d > 1.6:
            d -= 3.2
        while d < -1.6:
            d += 3.2
        assert r == d


@pytest.mark.parametrize("case", ["icra", "narrow", "fallback", "yaw_fixed", "three_vias"])
def test_sampler_matches_reference(o1, case):
    """same Philox counters -> the reference's Sampler::sample_set and O1's sampler give the same via sets"""
    lo, hi, Kv, sig = list(LO), list(HI), 1, 0.3
    if case == "narrow":
        lo, hi, sig = [0.30, -0.05, 0.1, -0.2], [0.36, 0.05, 0.2, 0.2], 0.2       # many rejections, many yaw wraps
    if case == "fallback":
        lo, hi, sig = [0.300, 0.10, 0.30, -1.6], [0.301, 0.1004, 0.3002, 1.6], 0.5   # acceptance ~1e-3 per try: most dims fall back to U(lo,hi)
    if case == "yaw_fixed":
        lo[3] = hi[3] = 0.25
    if case == "three_vias":
        Kv = 3
    oc = o1cfg(Kv, lo=lo, hi=hi, sample_zmin=0.15)
    rng = np.random.default_rng(5)
    mean = rng.uniform(lo, hi, (Kv, 4)); sigma = np.full((Kv, 4), sig)
    n = 400
    for it in (0, 7):
        pyref.arm_tape(pyref.MODE_TASK, SEED, it, query=3, cand0=2, streams_per_row=3 if lo[3] == hi[3] else 4, rows=Kv)
        rv = pyref.sample_sets(lo, hi, mean, sigma, 0.15, n)
        cnt = pyref.tape_counts()
        ov = o1.sample(oc, SEED, it, 3, 2, n, mean, sigma).reshape(n, Kv, 4)
        np.testing.assert_allclose(ov, rv, rtol=0, atol=1e-15)
        assert (rv[:, :, 2] >= 0.15).all() and (rv[:, :, :3] >= np.array(lo[:3]) - 1e-15).all()
        if case == "fallback":
            assert cnt["uniforms"] > n                     # the 100-try fallback really ran (tsp_sampler.h:22)
        if case == "narrow":
            assert cnt["normals"] > 2 * cnt["streams"]     # rejections really happened
        if case == "yaw_fixed":
            assert (rv[:, :, 3] == mean[None, :, 3]).all()


def test_reset_matches_reference():
    """Distribution::reset through Planner::reset (tsp_planner.h:54-69): z floors, clamp, sigma clamp order"""
    rng = np.random.default_rng(6)
    for rep in range(300):
        Kv = int(rng.integers(1, 4))
        zc, zd = float(rng.choice([0.0, 0.1])), float(rng.choice([0.0, 0.2, 0.3]))
        s0 = float(rng.choice([0.3, 1e-5, 0.9]))
        rc = pyref.dist_cfg(0.005, 0.2, 0.5, 1e-4, 0.5, zd, LO, HI)
        init = rng.uniform(np.array(LO) - 0.2, np.array(HI) + 0.2, (Kv, 4))
        m0 = init.copy(); m0[:, 2] = np.maximum(m0[:, 2], zc)                       # tsp_planner.h:64 (cfg_.z_min) before reset()
        rm, rs = pyref.dist_reset(rc, m0, s0)
        # the build's host statement (ssppcu_reset_distribution, sspp_cabi.cu): max(z, sample_zmin), max(z, dist_zmin), clamp
        em = init.copy(); em[:, 2] = np.maximum(np.maximum(em[:, 2], zc), zd); em = np.clip(em, LO, HI)
        es = max(min(max(s0, 1e-4), 0.5), 0.005)
        np.testing.assert_array_equal(rm, em)
        assert (rs == es).all()


def test_path_model_matches_reference(o1):
    rng = np.random.default_rng(7)
    u = np.linspace(0, 1, 41)
    for total in (3, 4, 5, 7):
        for rep in range(20):
            a, b = rng.uniform(LO, HI), rng.uniform(LO, HI)
            vias = rng.uniform(LO, HI, (total - 2, 4))
            ctrl, knots, pts = pyref.from_vias(a, b, total, vias, u)
            V = np.vstack([a, vias, b])
            oc = o1.interpolate(V, 2)
            np.testing.assert_allclose(oc, ctrl, atol=1e-12)
            np.testing.assert_allclose(o1.knots(total, 2), knots, atol=1e-15)
            op = np.array([o1.spline_eval(oc, 2, x) for x in u])
            np.testing.assert_allclose(op, pts, atol=1e-12)
            np.testing.assert_allclose(pts[[0, -1]], [a, b], atol=1e-12)            # interpolates the end points (and the vias)
            for i in range(total - 2):
                np.testing.assert_allclose(pyref.from_vias(a, b, total, vias, [(i + 1) / (total - 1)])[2][0], vias[i], atol=1e-12)


def test_floor_penalty_matches_reference():
    rng = np.random.default_rng(8)
    for z in rng.uniform(-0.1, 0.3, 200):
        for zmin, margin, scale in ((0.0, 0.01, 10.0), (0.1, 0.02, 3.0)):
            r = pyref.floor_penalty(zmin, margin, scale, [0.3, 0.1, z, 0.0])
            deficit = (zmin + margin) - z
            assert r == (scale * deficit * deficit if deficit > 0 else 0.0)


@pytest.mark.parametrize("scene,body", [("robocrane", ICRA_BODY), ("stacking", "block1"), ("planner", "block1")])
def test_collision_point_cost_matches_reference(scene, body):
    """Collision<Point>::collision_point_cost (Collision.h:84-103: mj_set_point, mj_forward, sum over ALL contacts with
    dist < -1e-3 of -1/(centre distance + 1e-4), or of -dist) against O1's point cost"""
    fm = FlatModel.load(scene)
    o1 = pyoracle.O1(fm, 1, 4, body)
    rm = pyref.RefModel(fm)
    rng = np.random.default_rng(9)
    Q = np.column_stack([rng.uniform(-0.7, 0.7, 1500), rng.uniform(-0.7, 0.7, 1500), rng.uniform(0.0, 0.6, 1500), rng.uniform(-1.6, 1.6, 1500)])
    rc = rm.point_costs(body, Q, True)
    oc = np.array([o1.point_cost(q)[0] for q in Q])
    np.testing.assert_allclose(oc, rc, rtol=1e-12, atol=1e-12)
    assert 0.02 < (rc != 0).mean() < 0.98
    rd = rm.point_costs(body, Q, False)                     # use_center_dist = false: penetration depths
    assert ((rd > 0) == (rc < 0)).all()


def test_reference_check_collision_point_is_wired():
    """Collision<Point>::check_collision_point (Collision.h:51-82) through oracle/_ref: its verdict equals 'the depth cost is
    positive' (same contacts, same -1e-3 test, Collision.h:84-103) and the printed pair is a filtered geom pair of the scene"""
    fm = FlatModel.load("planner")
    ref = pyref.RefModel(fm)
    rng = np.random.default_rng(21)
    n = 500
    pts = np.column_stack([rng.uniform(0.0, 0.7, n), rng.uniform(-0.7, 0.7, n), rng.uniform(0.0, 0.6, n), rng.uniform(-1.6, 1.6, n)])
    hit, g1, g2, dist = ref.check_points("block1", pts)
    depth = ref.point_costs("block1", pts, use_center=False)
    assert np.array_equal(hit, depth > 0) and 20 < hit.sum() < n - 20
    assert (dist[hit] < -1e-3).all() and (g1[hit] < g2[hit]).all() and (g1[~hit] == -1).all()
    assert (-dist[hit] <= depth[hit] + 1e-6).all()                       # the first contact's depth is one term of the sum (printed with 6 digits)
