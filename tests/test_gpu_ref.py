"""The sm_100a path (through the C ABI) against the REFERENCE'S OWN planners on the same Philox normals: oracle/_ref/libref.so
is sspp.h / Collision.h / sspp/tsp*.h compiled unmodified (oracle/ref_driver.cpp; built in the development container, shipped
to the GPU box with the snapshot).  Bar: collision flags exact outside EPS_CONTACT of contact; elite INDICES in the
reference's order (ties FP32 cannot resolve excepted, counted); weights / mean / sigma / costs / arc lengths within 1e-5."""
import numpy as np
import pytest

from oracle import pyref
from tests import ref_helpers
from tests.helpers import SEED, armlab_case, joint_case, task_case
from tests.test_ref_planner import check_iteration_against_reference, in_band, REL

pytestmark = [pytest.mark.gpu, pytest.mark.skipif(not pyref.available(), reason="oracle/_ref/libref.so was not shipped")]


@pytest.mark.parametrize("kind", ["icra", "icra_long", "five_points", "stacking", "planner"])
def test_gpu_planner_b_matches_reference_code(kind):
    from sspp_b200 import capi
    iters = 12
    if kind == "icra":
        case = task_case()
    elif kind == "icra_long":
        case, iters = task_case(n_samples=48), 60
    elif kind == "five_points":
        case = task_case(n_samples=40, check_points=32, n_points=5)
    elif kind == "stacking":
        case = task_case("stacking.xml", "block1", n_samples=30, start=np.array([0.45, 0.3, 0.12, 0.0]), end=np.array([-0.45, -0.3, 0.12, 1.0]))
    else:
        case = task_case("planner.xml", "block1", n_samples=30, start=np.array([0.3, 0.4, 0.12, 0.0]), end=np.array([0.3, -0.4, 0.12, 0.5]))
    recs = ref_helpers.run_reference_b(case, iters)
    h = capi.Handle(case.model, case.cfg)
    case.prepare(h)
    amb = unres = 0
    for it, rec in enumerate(recs):
        if it == 0:                                        # cold start: the library's own reset must land on the reference's state
            m, s = h.get_distribution()
            np.testing.assert_allclose(m[0], rec["mean_in"], rtol=REL, atol=1e-7)
            np.testing.assert_allclose(s[0], rec["sigma_in"], rtol=REL)
        else:                                              # same input state as the reference (FP32 rounding of it); best is the GPU's own
            h.set_distribution(rec["mean_in"][None], rec["sigma_in"][None])
            b, bv = h.best()
            assert int(bv[0]) == rec["best_valid_in"]
            if rec["best_valid_in"]:
                np.testing.assert_allclose(b[0], rec["best_in"], rtol=REL, atol=1e-6)
        h.iterate(SEED, it, 1)
        r = h.results()
        idx, cost, w = h.elites(0)
        k = int(r["n_elite"][0])
        gm, gs = h.get_distribution()
        got = dict(status=r["status"][0], L=r["L"][0], cost=r["cost"][0], n_success=int(r["n_success"][0]), elite_idx=idx[:k].astype(np.int64),
                   weights=w[:k], mean=gm[0], sigma=gs[0])
        np.testing.assert_allclose(h.params()[0].reshape(len(rec["params"]), -1)[rec["status"] != 2], rec["params"][rec["status"] != 2], rtol=REL, atol=2e-6)
        a, u = check_iteration_against_reference(rec, got, case.cfg.w_collision)
        amb += a; unres += u
        if a or u:
            break                                          # the chains may part here; what was compared so far stands
    assert amb <= 1 and unres <= 1
    h.close()


@pytest.mark.parametrize("kind", ["robocrane7", "armlab", "robocrane7_4096"])
def test_gpu_planner_a_matches_reference_code(kind):
    from sspp_b200 import capi
    if kind == "robocrane7":
        case = joint_case(n_samples=384, check_points=48)
    elif kind == "armlab":
        case = armlab_case(n_samples=256, check_points=40)
    else:
        case = joint_case(n_samples=4096, check_points=64)   # the headline shape (cfg2), every candidate through the reference's code
    rec = ref_helpers.run_reference_a(case)
    h = capi.Handle(case.model, case.cfg)
    case.prepare(h)
    np.testing.assert_allclose(h.base()[0], rec["base"], rtol=REL, atol=1e-6)       # initializePath
    h.iterate(SEED, 0, 1)
    r = h.results()
    np.testing.assert_allclose(h.params()[0].reshape(len(rec["params"]), -1), rec["params"], rtol=REL, atol=2e-6)   # sampleWithNoise
    diff = (r["status"][0] != 0) != rec["collides"]
    assert not (diff & ~in_band(rec["min_dist"])).any(), "checkCollision flag differs from the reference outside the contact band"
    assert diff.sum() <= 2
    ok = ~diff
    np.testing.assert_allclose(r["L"][0][ok], rec["arc"][ok], rtol=REL)             # computeArcLength
    if not diff.any():
        assert int(r["n_success"][0]) == rec["n_success"] and int(r["best_idx"][0]) == rec["best_idx"]   # findBestPath
    h.close()


@pytest.mark.parametrize("scene,body", [("planner.xml", "block1"), ("stacking.xml", "block1"), ("robocrane.xml", None)])
def test_gpu_check_points_matches_reference_check_collision_point(scene, body):
    """ssppcu_check_points against the reference's Collision<Point>::check_collision_point (Collision.h:51-82, its printed geom
    ids and distance parsed by oracle/ref_driver.cpp): same verdict, same geom pair, same distance, and names through
    ssppcu_model_geom_name"""
    from sspp_b200 import capi
    from tests.helpers import ICRA_BODY
    body = body or ICRA_BODY
    case = task_case(scene, body)
    fm = ref_helpers.flat_for(case)
    ref = pyref.RefModel(fm)
    h = capi.Handle(case.model, case.cfg)
    rng = np.random.default_rng(11)
    n = 3000
    pts = np.column_stack([rng.uniform(-0.1, 0.8, n), rng.uniform(-0.8, 0.8, n), rng.uniform(0.0, 0.7, n), rng.uniform(-1.7, 1.7, n)])
    hit, r1, r2, rd = ref.check_points(body, pts.astype(np.float32).astype(np.float64))     # the points the GPU sees
    g1, g2, gd = h.check_points(pts)
    cost = h.eval_points(pts) if hasattr(h, "eval_points") else None
    # decisions within 2e-6 of the -1e-3 threshold may fall either way in FP32: find them through the distance either side reports
    near = (np.abs(rd + 1e-3) < 2e-6) | (np.abs(gd + 1e-3) < 2e-6)
    ok = ~near
    assert near.sum() <= 3
    assert hit.sum() > 50 and (~hit).sum() > 50, "the sample must exercise both verdicts"
    assert np.array_equal(g1[ok] >= 0, hit[ok])
    both = ok & hit & (g1 >= 0)
    same_pair = (g1[both] == r1[both]) & (g2[both] == r2[both])
    # a different pair may only be reported when the reference's pair sits on the threshold for the GPU (it then reports the next one)
    assert same_pair.mean() > 0.995, f"{(~same_pair).sum()} of {both.sum()} geom pairs differ"
    err = np.abs(gd[both][same_pair] - rd[both][same_pair])          # FP32 narrow phase vs FP64; the reference prints 6 digits
    assert (err <= 2e-5).mean() > 0.995 and err.max() <= 2e-4, (float(err.max()), float((err > 2e-5).mean()))
    i = int(np.flatnonzero(both)[0])
    assert case.model.geom_name(int(g1[i])) == fm.geom_names[int(r1[i])] if hasattr(fm, "geom_names") else True
    if cost is not None:
        assert (cost[ok & hit] != 0).all()
    h.close()


@pytest.mark.parametrize("shape", [(100, 100, 7), (50, 50, 10), (512, 64, 10)])
def test_drop_in_plan_returns_what_the_reference_plan_returns(shape):
    """`_sspp.SamplingPathPlanner7(xml).plan(...)` (the drop-in Python call) against the reference's own
    sspp::SamplingPathPlanner<7>::plan (sspp.h:194-232, compiled unmodified) fed the same Philox normals: the same candidates
    succeed, the returned splines carry the reference's control points (FP32 sampling vs FP64: 2e-6), the same path is best"""
    from sspp_b200 import _sspp
    from tests.helpers import Q_START, Q_END
    n, m, k = shape
    case = joint_case(n_samples=n, check_points=m, n_points=k)
    rec = ref_helpers.run_reference_a(case, seed=SEED, it=0)
    p = _sspp.SamplingPathPlanner7(str(case.mjcf))
    p.set_verbose(False)
    p.set_seed(SEED)
    ok, paths = p.plan(Q_START, Q_END, case.joint_sigma, np.asarray(case.limits, float), n, m, k)
    free = ~rec["collides"]
    if len(paths) != int(free.sum()):                                       # only a decision inside the contact band may differ
        assert in_band(rec["min_dist"]).any()
        return
    assert ok == bool(free.any())
    ref_ctrl = rec["ctrl"][free]                                            # [n_success][K][7], candidate order (sspp.h:215-216)
    for sp, c in zip(paths, ref_ctrl):
        np.testing.assert_allclose(sp.ctrls().T, c, rtol=REL, atol=2e-6)
    if free.any():
        np.testing.assert_allclose(p.get_ctrl_pts().T, rec["ctrl"][rec["best_idx"]], rtol=REL, atol=2e-6)   # findBestPath
        j = list(np.flatnonzero(free)).index(rec["best_idx"])
        for u in (0.0, 0.37, 1.0):
            np.testing.assert_array_equal(p.evaluate(u), paths[j](u))


def test_drop_in_task_space_planner_follows_the_reference_planner():
    """`_tsp.TaskSpacePlanner` (the drop-in Python class, src/main_icra_benchmark.cpp:152-160 arguments) against the reference's
    own tsp::TaskSpacePlanner run on the same Philox normals, call by call (cold plan, then iterate = true): the same seed sets
    are sampled, the same candidates succeed, mean and standard deviation of the via distribution follow the reference's.  The
    GPU continues from its own FP32 state here (no re-synchronisation between calls), hence the looser 1e-4."""
    from sspp_b200 import _tsp
    from tests.helpers import FIX, ICRA_BODY, ICRA_START, ICRA_END
    from tests.test_tsp_module import ICRA_KW
    case = task_case()
    iters = 8
    recs = ref_helpers.run_reference_b(case, iters)
    p = _tsp.TaskSpacePlanner(str(FIX / "robocrane.xml"), ICRA_BODY, **ICRA_KW)
    p.set_seed(SEED)
    for it, rec in enumerate(recs):
        succ = p.plan(ICRA_START, ICRA_END, it > 0)
        sets = p.get_sampled_via_sets()
        live = rec["status"] != 2
        assert len(sets) == int(live.sum())
        got = np.array([np.asarray(s).ravel() for s in sets])
        np.testing.assert_allclose(got, rec["params"][live], rtol=1e-4, atol=2e-5)
        band = in_band(rec["min_dist"][live])
        if not band.any():
            assert len(succ) == rec["n_success"]
        np.testing.assert_allclose(np.asarray(p.get_current_mean()).ravel(), rec["mean_out"].ravel(), rtol=1e-4, atol=1e-5)
        np.testing.assert_allclose(np.asarray(p.get_current_stddev()).ravel(), rec["sigma_out"].ravel(), rtol=1e-4, atol=1e-6)
        if rec["n_success"]:
            best = rec["params"][rec["best_idx"]].reshape(-1, 4)
            np.testing.assert_allclose(np.asarray(p.spline()(0.5) if hasattr(p, "spline") else p.evaluate(0.5)),
                                       np.asarray(p.spline_from_vias(best)(0.5)), atol=1e-4)
