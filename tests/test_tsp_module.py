"""Planner B host mirror (tsp::TaskSpacePlanner, reference include/sspp/tsp.h:7-109), its `_tsp` Python
module and the re-hosted C++ benchmark drivers (reference src/main_icra_benchmark.cpp, src/main_tsp_benchmark.cpp)."""
import subprocess

import numpy as np
import pytest

from tests.helpers import FIX, ROOT, ICRA_BODY, ICRA_START, ICRA_END, SEED, task_case, o1_for

XML = str(FIX / "robocrane.xml")
ICRA_KW = dict(stddev_initial=0.2, stddev_min=1e-4, stddev_max=0.5, stddev_increase_factor=1.5, stddev_decay_factor=0.9,
               elite_fraction=0.3, sample_count=15, check_points=40, init_points=3, collision_weight=1.0, z_min=0.1,
               limits_min=[0.0, -0.7, 0.1, -1.6], limits_max=[0.7, 0.7, 0.6, 1.6], sigma_floor=0.005, var_ema_beta=0.2, mean_lr=0.5,
               floor_margin=0.01, floor_penalty_scale=10.0)


def test_surface_and_loud_failure_without_gpu():
    import torch
    from sspp_b200 import _tsp
    for name in ("plan", "get_succesful_path_candidates", "get_failed_path_candidates", "get_sampled_via_sets", "get_sampled_via_pts",
                 "get_via_pts", "get_current_mean", "get_current_stddev", "get_limits_min", "get_limits_max", "evaluate", "get_path_pts",
                 "get_ctrl_pts", "get_knot_vector", "set_verbose", "spline_from_via", "spline_from_vias", "reset"):
        assert hasattr(_tsp.TaskSpacePlanner, name), name            # every public method of tsp.h:58-102
    assert [s for s in ("Converged", "Failed", "BelowFloor", "MaxIter", "Unknown") if hasattr(_tsp.SolverStatus, s)].__len__() == 5
    with pytest.raises(RuntimeError, match="Failed to load MuJoCo model"):
        _tsp.TaskSpacePlanner("/nonexistent.xml", ICRA_BODY)
    if not torch.cuda.is_available():
        with pytest.raises(RuntimeError, match="no CPU path"):
            _tsp.TaskSpacePlanner(XML, ICRA_BODY)
        r = subprocess.run([str(ROOT / "drivers" / "icra_benchmark"), XML, ICRA_BODY, "2"], capture_output=True, text=True)
        assert r.returncode == 4 and "no CPU path" in r.stderr


@pytest.mark.gpu
def test_planner_b_semantics():
    from sspp_b200 import _tsp
    p = _tsp.TaskSpacePlanner(XML, ICRA_BODY, **ICRA_KW)
    p.set_seed(SEED)
    case = task_case()
    o1 = o1_for(case)
    succ = p.plan(ICRA_START, ICRA_END, False)
    fail = p.get_failed_path_candidates()
    sets = p.get_sampled_via_sets()
    assert len(succ) + len(fail) == len(sets) == 16                  # mean + 15 samples; no forwarded best on a cold call
    np.testing.assert_allclose(p.get_via_pts(), np.linspace(ICRA_START, ICRA_END, 3), atol=1e-12)
    np.testing.assert_allclose(sets[0][0, :2], (ICRA_START[:2] + ICRA_END[:2]) / 2, atol=1e-6)   # the mean set comes first ...
    assert abs(sets[0][0, 2] - 0.2) < 1e-6                            # ... with the adapter quirk: dist z_min = stddev_initial
    for c in succ:
        assert c.status == _tsp.SolverStatus.Converged and c.C_nf == 0.0 and c.via.shape == (1, 4)
        s = p.spline_from_vias(c.via)
        L = sum(np.linalg.norm(s((i + 1) / 40) - s(i / 40)) for i in range(40))     # tsp_evaluator.h:24-25: all 4 dims
        assert abs(L - c.L) < 1e-5 * max(L, 1)
        for i in range(1, 41):
            cost, ncon, md = o1.point_cost(s(i / 40))
            assert cost == 0.0 or abs(md + 1e-3) < 1e-4
    for c in fail:
        assert c.status == _tsp.SolverStatus.Failed and c.C_nf < 0.0  # penetration costs are negative (Collision.h:96)
    # warm iterations keep the distribution, add the forwarded best and shrink sigma on success
    s0 = p.get_current_stddev()
    found = bool(succ)
    for _ in range(12):
        found |= bool(p.plan(ICRA_START, ICRA_END, True))
    assert found and len(p.get_sampled_via_sets()) == 17
    assert (p.get_current_stddev() < s0).all()
    np.testing.assert_allclose(p.evaluate(0.0), ICRA_START, atol=1e-6)
    np.testing.assert_allclose(p.evaluate(1.0), ICRA_END, atol=1e-6)
    assert p.get_ctrl_pts().shape == (4, 3) and len(p.get_knot_vector()) == 6 and p.get_path_pts(7).shape == (7, 4)
    # a cold call resets: sigma back to the (quirk) initial 0.3 before its own adaptation step
    p.plan(ICRA_START, ICRA_END, False)
    assert (p.get_current_stddev() > 0.2).all()
    # fused: 60 iterations in one launch give the same state as 60 plan() calls
    a = _tsp.TaskSpacePlanner(XML, ICRA_BODY, **ICRA_KW); a.set_seed(7)
    b = _tsp.TaskSpacePlanner(XML, ICRA_BODY, **ICRA_KW); b.set_seed(7)
    a.plan_iterations(ICRA_START, ICRA_END, 60, False)
    b.plan(ICRA_START, ICRA_END, False)
    for _ in range(59):
        b.plan(ICRA_START, ICRA_END, True)
    assert np.array_equal(a.get_current_mean(), b.get_current_mean()) and np.array_equal(a.get_current_stddev(), b.get_current_stddev())
    assert np.array_equal(a.get_ctrl_pts(), b.get_ctrl_pts()) and a.any_success()


@pytest.mark.gpu
def test_benchmark_drivers_run():
    r = subprocess.run([str(ROOT / "drivers" / "icra_benchmark"), XML, ICRA_BODY, "5", "--budgets_ms=5,10", "--converged"],
                       capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stderr
    assert "=== Converged (max_iter=60, N=5) ===" in r.stdout and "=== Anytime (budgets ms: 5,10, N=5) ===" in r.stdout
    assert "[B=10] Cold: succ 5/5" in r.stdout and "Warm: succ 5/5" in r.stdout
    r = subprocess.run([str(ROOT / "drivers" / "tsp_benchmark"), XML, ICRA_BODY, "20"], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stderr
    assert "=== Benchmark Results (N=20) ===" in r.stdout and "[Warm start] iterate=true" in r.stdout
