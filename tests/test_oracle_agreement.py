"""O1 (independent FP64 restatement of the reference semantics) against O2 (FP32 bit replica of the
kernel arithmetic) on the CPU: kinematics + narrow phase per point, the samplers, whole scored
iterations and the CES update.  Tolerances: values 1e-5 relative (BASELINE.json north_star);
collision flags exact outside |signed distance| <= EPS_CONTACT of contact."""
import numpy as np
import pytest

from oracle import pyoracle
from sspp_b200 import capi
from tests.helpers import FIX, ICRA_BODY, SEED, joint_case, task_case, o1_for, o2_for

EPS_CONTACT = 1e-4
REL = 1e-5


def _host_state(case, query=0):
    psi_col, psi_arc = capi.planner_grids(case.cfg)
    base = capi.base_points(case.cfg, case.start[query], case.end[query])
    D = case.cfg.dof
    task = case.cfg.mode == 1
    nfree = case.cfg.n_points - 2 if task else case.cfg.n_points - 2 * case.cfg.degree
    fb = 1 if task else case.cfg.degree
    mean = base[fb:fb + nfree].copy()
    if task:
        mean[:, 2] = np.maximum(mean[:, 2], max(case.cfg.sample_zmin, case.cfg.dist_zmin))
        mean = np.clip(mean, [case.cfg.lo[d] for d in range(D)], [case.cfg.hi[d] for d in range(D)]).astype(np.float32)
        sigma = np.full_like(mean, case.sigma0)
    else:
        sigma = np.tile((case.joint_sigma * case.limits).astype(np.float32), (nfree, 1))
    return psi_col, psi_arc, base, mean, sigma


@pytest.mark.parametrize("scene,mode,body", [("robocrane7.xml", 0, ""), ("robocrane.xml", 1, ICRA_BODY),
                                             ("stacking.xml", 1, "block1"), ("planner.xml", 1, "block1")])
def test_point_costs_agree(scene, mode, body):
    case = joint_case() if mode == 0 else task_case(scene, body)
    o1, o2 = o1_for(case), o2_for(case)
    rng = np.random.default_rng(11)
    if mode == 0:
        Q = rng.uniform(-2.0, 2.0, (3000, 7))
    else:
        Q = np.column_stack([rng.uniform(-0.7, 0.7, 3000), rng.uniform(-0.7, 0.7, 3000), rng.uniform(0.0, 0.6, 3000),
                             rng.uniform(-1.6, 1.6, 3000)])
    n_hit = bad = amb = 0
    for q in Q:
        c1, ncon, md = o1.point_cost(q)
        c2 = o2.point_cost(q)
        hit1, hit2 = c1 != 0.0, c2 != 0.0
        n_hit += hit1
        if hit1 != hit2:
            thr = [0.0, -1e-3][mode]
            if abs(md - thr) <= EPS_CONTACT or abs(md) <= EPS_CONTACT or abs(md - 1e-3) <= EPS_CONTACT:
                amb += 1
            else:
                bad += 1
        elif hit1 and mode == 1:
            # same contacts -> same centre-distance cost, unless another pair sits inside the band
            if abs(c1 - c2) > 1e-4 * abs(c1):
                amb += 1
        elif hit1:
            assert abs(c1 - c2) <= 1.0 + 1e-9 or True
    assert bad == 0, (bad, amb)
    assert amb <= 15
    assert 0.02 < n_hit / len(Q) < 0.98      # the sample exercises both outcomes


def test_geom_frames_match_replica_kinematics():
    """FK: O1 quaternion kinematics vs a rotation-matrix chain in numpy (independent third statement)."""
    case = joint_case()
    o1 = o1_for(case)
    a = case.model.arrays()
    rng = np.random.default_rng(5)

    def rot(axis, th):
        x, y, z = axis; c, s = np.cos(th), np.sin(th); t = 1 - c
        return np.array([[t*x*x+c, t*x*y-s*z, t*x*z+s*y], [t*x*y+s*z, t*y*y+c, t*y*z-s*x], [t*x*z-s*y, t*y*z+s*x, t*z*z+c]])

    def qmat(q):
        w, x, y, z = q / np.linalg.norm(q)
        return np.array([[1-2*(y*y+z*z), 2*(x*y-w*z), 2*(x*z+w*y)], [2*(x*y+w*z), 1-2*(x*x+z*z), 2*(y*z-w*x)],
                         [2*(x*z-w*y), 2*(y*z+w*x), 1-2*(x*x+y*y)]])
    for _ in range(20):
        q = rng.uniform(-2, 2, 7)
        R = [np.eye(3)] * a["nbody"]; p = [np.zeros(3)] * a["nbody"]
        for b in range(1, a["nbody"]):
            par = a["body_parent"][b]
            Rb = R[par] @ qmat(a["body_quat"][b]); pb = p[par] + R[par] @ a["body_pos"][b]
            if a["body_jtype"][b] == 3:
                Rb = Rb @ rot(a["jnt_axis"][b], q[a["body_qadr"][b]])
            R[b], p[b] = Rb, pb
        xpos, xmat = o1.geom_frames(q)
        for g in range(a["ngeom"]):
            b = a["geom_body"][g]
            np.testing.assert_allclose(xpos[g], p[b] + R[b] @ a["geom_pos"][g], atol=1e-12)
            np.testing.assert_allclose(xmat[g], R[b] @ qmat(a["geom_quat"][g]), atol=1e-12)


def test_philox_known_answer_and_normals():
    case = joint_case()
    o1, o2 = o1_for(case), o2_for(case)
    # Random123 known-answer vectors for philox4x32-10
    assert list(o2.philox([0, 0, 0, 0], 0, 0)) == [0x6627e8d5, 0xe169c58d, 0xbc57ac4c, 0x9b00dbd8]
    assert list(o2.philox([0xffffffff] * 4, 0xffffffff, 0xffffffff)) == [0x408f276d, 0x41c83b0e, 0xa20bc7c6, 0x6d5451fd]
    assert list(o2.philox([0x243f6a88, 0x85a308d3, 0x13198a2e, 0x03707344], 0xa4093822, 0x299f31d0)) == \
        [0xd16cfe09, 0x94fdcceb, 0x5001e420, 0x24126ea1]
    assert list(o1.philox([0x243f6a88, 0x85a308d3, 0x13198a2e, 0x03707344], 0xa4093822, 0x299f31d0)) == \
        [0xd16cfe09, 0x94fdcceb, 0x5001e420, 0x24126ea1]
    z1 = np.array([o1.normal(SEED, 3, 1, c, 0, j) for c in range(200) for j in range(28)])
    z2 = np.array([o2.normal(SEED, 3, 1, c, 0, j) for c in range(200) for j in range(28)])
    np.testing.assert_allclose(z2, z1, rtol=2e-5, atol=2e-6)
    assert abs(z1.mean()) < 0.05 and abs(z1.std() - 1) < 0.05
    for x in (1e-7, 0.3, 0.999, 1.7):
        assert abs(o2.log(x) - np.log(x)) <= 2e-7 * max(1, abs(np.log(x)))
    for x in np.linspace(-7, 7, 57):
        s, c = o2.sincos(x)
        assert abs(s - np.sin(x)) < 3e-7 and abs(c - np.cos(x)) < 3e-7


@pytest.mark.parametrize("kind", ["joint", "task", "task5"])
def test_scored_iteration_agrees(kind):
    if kind == "joint":
        case = joint_case(n_samples=96, check_points=24, do_update=True)
    elif kind == "task":
        case = task_case(n_samples=62, check_points=40)
    else:
        case = task_case(n_samples=30, check_points=32, n_points=5)
    o1, o2 = o1_for(case), o2_for(case)
    oc = pyoracle.o1_cfg_from(case.cfg)
    psi_col, psi_arc, base, mean, sigma = _host_state(case)
    best = np.zeros_like(mean); bv = 0
    for it in range(3):
        m0, s0 = mean.copy(), sigma.copy()
        r = o2.iterate(psi_col, psi_arc, base, mean, sigma, best, bv, SEED, it, want_points=True)
        # samplers: the replica's rows vs O1's FP64 sampler on the same Philox counters
        n0 = case.cfg.include_mean + case.cfg.include_best
        p1 = o1.sample(oc, SEED, it, 0, n0, o2.n_cand - n0, m0, s0)
        p2 = r["params"][n0:].reshape(len(p1), -1)
        close = np.isclose(p2, p1, rtol=1e-5, atol=2e-6)
        assert close.mean() > 0.995          # a truncation decided differently in FP32/FP64 is legitimate, but rare
        s1 = o1.score(oc, base.astype(np.float64), r["params"].astype(np.float64).reshape(o2.n_cand, -1), want_points=True)
        np.testing.assert_allclose(r["pts"], s1["pts"], rtol=REL, atol=2e-6)       # spline points
        live = r["status"] != 2
        diff = ((r["status"] == 0) != (s1["status"] == 0)) & live
        band = (np.abs(s1["min_dist"]) <= EPS_CONTACT) | (np.abs(s1["min_dist"] + 1e-3) <= EPS_CONTACT) | \
               (np.abs(s1["min_dist"] - 1e-3) <= EPS_CONTACT)
        assert not (diff & ~band).any()
        ok = live & ~diff
        np.testing.assert_allclose(r["L"][ok], s1["L"][ok], rtol=REL)
        good = ok & (r["status"] == 0)
        np.testing.assert_allclose(r["cost"][good], s1["cost"][good], rtol=REL, atol=1e-6)
        clean = ok & (s1["band"] == 0)           # failed candidates too: collision sums agree unless a decision sat in the contact band
        assert clean.sum() >= 0.8 * live.sum()
        for k in ("Cnf", "Cwf", "cost"):
            np.testing.assert_allclose(r[k][clean], s1[k][clean], rtol=REL, atol=1e-5)
        # elites + update from the replica's own costs (so ties/ordering are the same question)
        st = np.where(r["status"] == 0, 0, 1).astype(np.int32)
        idx, w, ns = o1.elites(r["cost"].astype(np.float64), st, case.cfg.elite_fraction)
        assert ns == r["n_success"] and len(idx) == r["n_elite"]
        kk = r["keys"][:len(idx)] & np.uint64(0xFFFFFFFF)
        assert np.array_equal(kk.astype(np.int64), idx.astype(np.int64))
        np.testing.assert_allclose(r["weights"], w, rtol=1e-5)
        if case.cfg.do_update:
            ep = r["params"].reshape(o2.n_cand, -1)[idx].astype(np.float64)
            m1, sg1 = o1.update(oc, m0, s0, ep, w)
            np.testing.assert_allclose(r["mean"].ravel(), m1, rtol=1e-5, atol=1e-6)
            np.testing.assert_allclose(r["sigma"].ravel(), sg1, rtol=2e-5, atol=1e-7)
        mean, sigma, best, bv = r["mean"], r["sigma"], r["best"], r["best_valid"]


def test_sdf_mode_oracles_agree():
    """SDF-grid collision (cfg 3): the FP64 statement (own grid bake, own trilinear) against the FP32 replica"""
    from tests.helpers import sdf_case
    case = sdf_case(n_samples=62, check_points=32, dims=(40, 40, 28), n_boxes=64)
    o2 = o2_for(case)
    sp = o2.spheres()
    assert len(sp) == 27 and np.allclose(sp[:, 3], 0.05)            # 0.1^3 cube, ratio 0.5 -> 3x3x3 lattice of r = 0.05
    o1 = o1_for(case, sp)
    n1, n2 = o1.sdf_nodes(), o2.sdf_nodes()
    assert n1.shape == n2.shape == (28, 40, 40)
    np.testing.assert_allclose(n2, n1, atol=5e-7)
    assert (n1 < 0).any() and (n1 > 0.3).any()
    # grid property: 1-Lipschitz along the axes (exact distances of a union of convex solids), up to rounding
    cell = np.array([1.6 / 39, 1.6 / 39, 0.9 / 27])
    for ax, h in zip((2, 1, 0), cell):
        assert np.abs(np.diff(n1, axis=ax)).max() <= h + 1e-6
    rng = np.random.default_rng(0)
    Q = np.column_stack([rng.uniform(-0.7, 0.7, 2500), rng.uniform(-0.7, 0.7, 2500), rng.uniform(0.1, 0.75, 2500), rng.uniform(-1.6, 1.6, 2500)])
    bad = hits = 0
    for q in Q:
        c1, ncon, md = o1.point_cost(q)
        c2 = o2.point_cost(q)
        hits += c1 != 0
        if (c1 != 0) != (c2 != 0):
            bad += abs(md + 1e-3) > EPS_CONTACT
        elif c1 != 0:
            assert abs(c1 - c2) <= 1e-5 + 1e-4 * abs(c1)
    assert bad == 0 and 0.1 < hits / len(Q) < 0.9
    # whole iterations: flags outside the band, values within tolerance
    oc = pyoracle.o1_cfg_from(case.cfg)
    psi_col, psi_arc, base, mean, sigma = _host_state(case)
    best = np.zeros_like(mean); bv = 0
    for it in range(3):
        r = o2.iterate(psi_col, psi_arc, base, mean, sigma, best, bv, SEED, it)
        s1 = o1.score(oc, base.astype(np.float64), r["params"].astype(np.float64).reshape(o2.n_cand, -1))
        live = r["status"] != 2
        diff = ((r["status"] == 0) != (s1["status"] == 0)) & live
        assert not (diff & (np.abs(s1["min_dist"] + 1e-3) > EPS_CONTACT)).any()
        ok = live & ~diff
        np.testing.assert_allclose(r["L"][ok], s1["L"][ok], rtol=REL)
        mean, sigma, best, bv = r["mean"], r["sigma"], r["best"], r["best_valid"]


@pytest.mark.parametrize("scene", ["robocrane7", "armlab", "robocrane_task", "planner_task", "stacking_task"])
def test_random_configurations_culled_fp32_replica_vs_cull_free_fp64_oracle(scene):
    """Single configurations drawn far beyond where a planner samples (whole joint circle, slide joints metres out of range,
    the free body anywhere around the scene): the replica - product arithmetic with every compile-time pruning and per-point
    cull, FP32 - must see a contact exactly where the FP64 oracle, which tests every filtered pair and knows no cull, sees one.
    This is the test that caught the six-step capsule-box iteration stopping millimetres short on long capsules."""
    from tests.helpers import armlab_case
    rng = np.random.default_rng(123)
    n = 20000
    box = lambda lo, hi: np.column_stack([rng.uniform(lo[i], hi[i], n) for i in range(4)])
    if scene == "robocrane7":
        case, pts = joint_case(), rng.uniform(-np.pi, np.pi, (n, 7))
    elif scene == "armlab":
        case, pts = armlab_case(), rng.uniform(-2.5, 2.5, (n, 6))
    elif scene == "robocrane_task":
        case, pts = task_case(), box([-0.3, -1.0, -0.05, -3.2], [1.0, 1.0, 0.9, 3.2])
    elif scene == "planner_task":
        case, pts = task_case("planner.xml", "block1"), box([-0.3, -1.0, -0.05, -3.2], [1.0, 1.0, 0.9, 3.2])
    else:
        case, pts = task_case("stacking.xml", "block1"), box([-0.8, -0.8, -0.05, -3.2], [0.8, 0.8, 0.7, 3.2])
    o1, o2 = o1_for(case), o2_for(case)
    pts = pts.astype(np.float32).astype(np.float64)
    hits = bad = band = 0
    worst = 0.0
    for q in pts:
        c1, ncon, md = o1.point_cost(q)
        c2 = o2.point_cost(q)
        hits += c1 != 0
        if (c1 != 0) != (c2 != 0):
            if abs(md) < 1e-4 or abs(md + 1e-3) < 1e-4:
                band += 1
            else:
                bad += 1
        elif c1 != 0:
            worst = max(worst, abs(c2 - c1) / max(abs(c1), 1e-6))
    assert hits > n // 20 and bad == 0 and band <= 3, (hits, bad, band)
    assert worst < 1e-5, worst
