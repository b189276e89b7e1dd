"""Whole planners of the REFERENCE (sspp.h's SamplingPathPlanner, tsp.h's TaskSpacePlanner — compiled unmodified into
oracle/_ref/libref.so, see oracle/ref_driver.cpp) against the oracles on the same Philox normals.

  reference code == O1 to 1e-12 (asserted inside tests/ref_helpers.py while the records are produced)
  FP32 replica O2 (== the GPU bit for bit, tests/test_gpu_parity.py) vs the reference: collision flags exact outside the
  contact band EPS_CONTACT, elite order exact unless two costs are closer than FP32 resolves, values within 1e-5."""
import numpy as np
import pytest

from oracle import pyref
from sspp_b200 import capi
from tests import ref_helpers
from tests.helpers import SEED, armlab_case, joint_case, o2_for, task_case

pytestmark = pytest.mark.skipif(not pyref.available(), reason="oracle/_ref/libref.so is not built (needs /root/reference)")

EPS_CONTACT = 1e-4
REL = 1e-5


def in_band(md):
    return (np.abs(md) <= EPS_CONTACT) | (np.abs(md + 1e-3) <= EPS_CONTACT) | (np.abs(md - 1e-3) <= EPS_CONTACT)


def check_iteration_against_reference(rec, got, w_collision=1.0):
    """got: status / L / cost / elite candidate indices / weights / mean / sigma of an FP32 implementation that started the
    iteration from rec's input state.  Returns (ambiguous flags, elite positions whose order FP32 cannot resolve)."""
    live = rec["status"] != 2
    assert np.array_equal(got["status"] == 2, ~live)
    diff = ((got["status"] == 0) != (rec["status"] == 0)) & live
    assert not (diff & ~in_band(rec["min_dist"])).any(), "collision flag differs from the reference outside the contact band"
    ok = live & ~diff
    np.testing.assert_allclose(got["L"][ok], rec["L"][ok], rtol=REL)
    good = ok & (rec["status"] == 0)
    np.testing.assert_allclose(got["cost"][good], rec["cost"][good], rtol=REL, atol=1e-6)
    unresolved = 0
    if not diff.any():
        assert got["n_success"] == rec["n_success"]
        k = len(rec["elite_idx"])
        assert len(got["elite_idx"]) == k
        for j in np.where(got["elite_idx"] != rec["elite_idx"])[0]:
            a, b = rec["cost"][got["elite_idx"][j]], rec["cost"][rec["elite_idx"][j]]
            assert abs(a - b) <= 2e-6 * max(abs(a), abs(b)), "elite order differs from the reference beyond FP32 resolution"
            unresolved += 1
        if unresolved == 0:
            np.testing.assert_allclose(got["weights"], rec["weights"], rtol=REL)
            np.testing.assert_allclose(got["mean"], rec["mean_out"], rtol=REL, atol=1e-6)
            np.testing.assert_allclose(got["sigma"], rec["sigma_out"], rtol=2e-5, atol=1e-7)
    return int(diff.sum()), unresolved


@pytest.mark.parametrize("kind", ["icra", "icra_long", "five_points", "stacking", "planner"])
def test_planner_b_reference_vs_oracles(kind):
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
    recs = ref_helpers.run_reference_b(case, iters)           # reference code == O1 asserted inside
    assert any(r["n_success"] > 0 for r in recs) and any((r["status"] == 1).any() for r in recs)
    o2 = o2_for(case)
    psi_col, _ = capi.planner_grids(case.cfg)
    base = capi.base_points(case.cfg, case.start[0], case.end[0])
    amb = unres = 0
    for it, rec in enumerate(recs):
        r = o2.iterate(psi_col, None, base, rec["mean_in"], rec["sigma_in"], rec["best_in"], rec["best_valid_in"], SEED, it)
        k = r["n_elite"]
        got = dict(status=r["status"], L=r["L"], cost=r["cost"], n_success=r["n_success"],
                   elite_idx=(r["keys"][:k] & np.uint64(0xFFFFFFFF)).astype(np.int64), weights=r["weights"], mean=r["mean"], sigma=r["sigma"])
        a, u = check_iteration_against_reference(rec, got, case.cfg.w_collision)
        amb += a; unres += u
    assert amb <= max(2, iters // 10) and unres <= 2


@pytest.mark.parametrize("kind", ["robocrane7", "armlab"])
def test_planner_a_reference_vs_oracles(kind):
    case = joint_case(n_samples=384, check_points=48) if kind == "robocrane7" else armlab_case(n_samples=256, check_points=40)
    rec = ref_helpers.run_reference_a(case)                   # reference code == O1 asserted inside
    assert 0.1 < rec["collides"].mean() < 0.9
    o2 = o2_for(case)
    psi_col, psi_arc = capi.planner_grids(case.cfg)
    base = capi.base_points(case.cfg, case.start[0], case.end[0])
    np.testing.assert_allclose(base, rec["base"], rtol=REL, atol=1e-6)            # product's host spline set-up vs initializePath
    nf = case.cfg.n_points - 2 * case.cfg.degree
    mean = base[case.cfg.degree:case.cfg.degree + nf]
    sigma = np.tile((case.joint_sigma * case.limits).astype(np.float32), (nf, 1))
    r = o2.iterate(psi_col, psi_arc, base, mean, sigma, np.zeros_like(mean), 0, SEED, 0)
    np.testing.assert_allclose(r["params"].reshape(len(rec["params"]), -1), rec["params"], rtol=REL, atol=2e-6)
    diff = (r["status"] != 0) != rec["collides"]
    assert not (diff & ~in_band(rec["min_dist"])).any(), "checkCollision flag differs from the reference outside the contact band"
    ok = ~diff
    np.testing.assert_allclose(r["L"][ok], rec["arc"][ok], rtol=REL)               # computeArcLength
    if not diff.any():
        assert r["n_success"] == rec["n_success"] and r["best_idx"] == rec["best_idx"]   # findBestPath
