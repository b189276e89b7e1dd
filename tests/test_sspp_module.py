"""The drop-in Python surface `_sspp` (reference: src/sspp_bindings.cpp:14-68): names, signatures,
error behaviour on the CPU; plan() semantics against the oracles on the GPU."""
import numpy as np
import pytest

from tests.helpers import FIX, ROOT, Q_START, Q_END, SEED, joint_case, o1_for

XML = str(FIX / "robocrane7.xml")


@pytest.fixture(scope="module")
def sp():
    from sspp_b200 import _sspp
    return _sspp


def test_surface_matches_reference_names(sp):
    for d in (3, 6, 7, 9):
        assert hasattr(sp, f"SamplingPathPlanner{d}") and hasattr(sp, f"Spline{d}")
    P = sp.SamplingPathPlanner7
    for name in ("initializePath", "evaluate", "sampleWithNoise", "checkCollision", "computeArcLength", "findBestPath",
                 "get_ctrl_pts", "plan"):
        assert hasattr(P, name), name
    assert hasattr(sp.Spline7, "ctrls")
    assert sp.create_model_capsule(1234) is not None and sp.create_data_capsule(99) is not None
    doc = P.plan.__doc__
    import re
    order = [doc.index(k) for k in ("start:", "end:", "sigma:", "limits:", "sample_count:", "check_points:", "init_points:")]
    assert order == sorted(order)                                     # positional order of src/sspp_bindings.cpp:50
    for kw, dflt in (("sample_count", 50), ("check_points", 50), ("init_points", 10)):
        assert re.search(rf"{kw}: [^,)]*= {dflt}\b", doc), (kw, doc)


def test_errors_like_the_reference(sp):
    with pytest.raises(RuntimeError, match="Failed to load MuJoCo model from XML"):
        sp.SamplingPathPlanner7("/nonexistent/scene.xml")
    p = sp.SamplingPathPlanner7(XML)
    s = sp.Spline7()
    with pytest.raises(TypeError):
        p.sampleWithNoise(s, 0.1, np.ones(7), object())
    with pytest.raises(TypeError):
        p.checkCollision(s, 10, object())
    with pytest.raises(TypeError):
        p.initializePath(np.zeros(6), np.zeros(7), s)


def test_initialize_evaluate_arclength_host_side(sp):
    from oracle import bsplines_np  # noqa: F401  (oracle only as the checker)
    case = joint_case()
    o1 = o1_for(case)
    p = sp.SamplingPathPlanner7(XML)
    s = sp.Spline7()
    assert p.initializePath(Q_START, Q_END, s, 10) is True
    c = s.ctrls()
    assert c.shape == (7, 10) and c.dtype == np.float64
    V = np.linspace(0, 1, 10)[:, None] * (Q_END - Q_START)[None, :] + Q_START[None, :]
    np.testing.assert_allclose(c.T, o1.interpolate(V, 3), atol=1e-12)
    np.testing.assert_allclose(s.knots(), o1.knots(10, 3), atol=1e-15)
    for u in (0.0, 0.123, 0.5, 0.999, 1.0):
        np.testing.assert_allclose(p.evaluate(s, u), o1.spline_eval(c.T.copy(), 3, u), atol=1e-12)
        np.testing.assert_allclose(s(u), (1 - u) * Q_START + u * Q_END, atol=1e-12)     # a straight line stays straight
    assert abs(p.computeArcLength(s, 50) - np.linalg.norm(Q_END - Q_START)) < 1e-12
    best = sp.Spline7()
    assert p.findBestPath([s], best, 10) is True
    np.testing.assert_array_equal(best.ctrls(), c)
    assert p.findBestPath([], best) is False
    for K in (4, 7, 16):
        assert p.initializePath(Q_START, Q_END, s, K) and s.ctrls().shape == (7, K)


@pytest.mark.gpu
def test_plan_semantics_on_gpu(sp, capfd):
    case = joint_case(n_samples=100, check_points=100, n_points=7)
    o1 = o1_for(case)
    p = sp.SamplingPathPlanner7(XML)
    p.set_seed(SEED)
    lim = np.full(7, np.pi)
    ok, paths = p.plan(Q_START, Q_END, 0.08, lim, 100, 100, 7)          # scripts/main.py:168-173 parameters
    out = capfd.readouterr().out
    assert f"Sampled 100 splines. Successful paths found: {len(paths)}" in out
    assert ok and 0 < len(paths) < 100
    # every returned spline is collision free at the reference's check grid, per the FP64 oracle
    lengths = []
    for s in paths:
        c = s.ctrls()
        assert c.shape == (7, 7)
        np.testing.assert_allclose(c[:, :3].T, paths[0].ctrls()[:, :3].T, atol=0)      # boundary control points fixed (sspp.h:121-127)
        for i in range(0, 101, 5):
            cost, ncon, md = o1.point_cost(p.evaluate(s, i / 100))
            assert ncon == 0 or abs(md) < 1e-4
        lengths.append(p.computeArcLength(s, 100))
    # the stored best path is the shortest success (findBestPath, sspp.h:171-192)
    best = p.get_ctrl_pts()
    j = int(np.argmin(lengths))
    np.testing.assert_allclose(best, paths[j].ctrls(), atol=0)
    np.testing.assert_allclose(p.evaluate(0.0), Q_START, atol=1e-6)
    np.testing.assert_allclose(p.evaluate(1.0), Q_END, atol=1e-6)
    # an empty population: the reference's loop does not run, plan() prints its line and returns false (sspp.h:207-224)
    ok0, paths0 = p.plan(Q_START, Q_END, 0.08, lim, 0, 100, 7)
    assert ok0 is False and paths0 == [] and "Sampled 0 splines. Successful paths found: 0" in capfd.readouterr().out
    for bad in (dict(check_points=1), dict(check_points=5000), dict(init_points=3), dict(init_points=60)):     # limits of this build, stated loudly
        with pytest.raises(RuntimeError):
            p.plan(Q_START, Q_END, 0.08, lim, **dict(dict(sample_count=10, check_points=20, init_points=10), **bad))
    # defaults + repeated calls draw new samples; same seed reproduces
    ok2, paths2 = p.plan(Q_START, Q_END, 0.08, lim)
    assert len(paths2) <= 50
    p.set_seed(SEED)
    ok3, paths3 = p.plan(Q_START, Q_END, 0.08, lim, 100, 100, 7)
    assert len(paths3) == len(paths) and all(np.array_equal(a.ctrls(), b.ctrls()) for a, b in zip(paths, paths3))


@pytest.mark.gpu
def test_plan_result_objects_are_recycled_only_when_nobody_holds_them(sp):
    """plan() hands the Spline objects of an earlier result list out again only if nothing else refers to them: a path, or a
    numpy view of a path's control points, that the caller kept never changes under its holder"""
    import gc
    p = sp.SamplingPathPlanner7(XML)
    p.set_verbose(False)
    p.set_seed(SEED)
    lim = np.full(7, np.pi)
    ok, paths = p.plan(Q_START, Q_END, 0.08, lim, 200, 64, 10)
    assert ok and len(paths) > 20
    keep = paths[3]; keep_c = keep.ctrls().copy()
    view = paths[5].ctrls(); view_c = view.copy()            # keeps paths[5] alive through its base
    lazy = paths[7]                                           # never touched: its control points are still unbuilt
    ids_before = {id(x) for x in paths}
    first = [x.ctrls().copy() for x in paths]
    del paths
    gc.collect()
    ok2, paths2 = p.plan(Q_START, Q_END, 0.08, lim, 200, 64, 10)       # the next draw of the same planner: different samples
    assert ok2 and not any(y is keep or y is lazy for y in paths2)
    np.testing.assert_array_equal(keep.ctrls(), keep_c)
    np.testing.assert_array_equal(view, view_c)
    np.testing.assert_array_equal(lazy.ctrls(), first[7])
    assert len({id(y) for y in paths2}) == len(paths2)
    assert ids_before & {id(y) for y in paths2}, "released objects are expected to be handed out again"
    assert not any(np.array_equal(y.ctrls(), first[0]) for y in paths2[:5])
    p.set_seed(SEED)
    ok3, paths3 = p.plan(Q_START, Q_END, 0.08, lim, 200, 64, 10)       # paths2 is still held: fresh objects, same values as the first call
    assert len(paths3) == len(first) and all(np.array_equal(a.ctrls(), b) for a, b in zip(paths3, first))
    assert not ({id(y) for y in paths2} & {id(y) for y in paths3})
    # the loop a caller writes: the previous list is still bound while the next call runs; two object sets alternate
    seen = []
    for it in range(6):
        ok, paths = p.plan(Q_START, Q_END, 0.08, lim, 200, 64, 10)
        seen.append({id(y) for y in paths})
    assert seen[5] & seen[3] and not (seen[5] & seen[4])


@pytest.mark.gpu
def test_plan_from_several_python_threads(sp):
    """plan() releases the GIL around launch + wait; calls on ONE planner from several threads take turns (a mutex in the binding)
    and every caller gets a consistent result list"""
    import threading
    p = sp.SamplingPathPlanner7(XML)
    p.set_verbose(False)
    lim = np.full(7, np.pi)
    out, errs = [], []

    def work():
        try:
            for _ in range(15):
                ok, paths = p.plan(Q_START, Q_END, 0.08, lim, 256, 32, 10)
                d = p.plan_device(Q_START, Q_END, 0.08, lim, 256, 32, 10)
                assert (d["status"] == 0).sum() == d["n_success"]
                out.append((ok, len(paths), all(np.isfinite(x.ctrls()).all() and x.ctrls().shape == (7, 10) for x in paths[:5])))
        except Exception as e:                                   # surfaces in the main thread below
            errs.append(e)
    ts = [threading.Thread(target=work) for _ in range(4)]
    [t.start() for t in ts]; [t.join() for t in ts]
    assert not errs, errs
    assert len(out) == 60 and all(ok and n > 0 and fin for ok, n, fin in out)


@pytest.mark.gpu
def test_plan_device_buffers(sp):
    import torch
    p = sp.SamplingPathPlanner7(XML)
    p.set_verbose(False)
    d = p.plan_device(Q_START, Q_END, 0.08, np.full(7, np.pi), 512, 64, 10, fetch_ctrl=True)
    assert d["n_candidates"] == 512 and d["ctrl"].shape == (512, 4, 7)
    assert (d["status"] == 0).sum() == d["n_success"]
    ok = d["status"] == 0
    # the collision-free candidates' rows travel with the results, compacted in candidate order; all rows only on request
    assert np.array_equal(d["success_idx"], np.flatnonzero(ok)) and np.array_equal(d["ctrl_success"], d["ctrl"][ok])
    assert d["best_idx"] == int(np.flatnonzero(ok)[np.argmin(d["arc_length"][ok])])

    class Dev:
        def __init__(self, ptr, shape, typestr):
            self.__cuda_array_interface__ = dict(shape=shape, typestr=typestr, data=(ptr, False), version=2)
    t = torch.as_tensor(Dev(d["device_ptrs"]["arc_length"], (512,), "<f4"), device="cuda")
    np.testing.assert_array_equal(t.cpu().numpy(), d["arc_length"])          # zero-copy view of the device results
    assert "ctrl" not in p.plan_device(Q_START, Q_END, 0.08, np.full(7, np.pi), 512, 64, 10)


def test_importable_under_the_references_name(tmp_path):
    """the only line a maintainer adds to the reference's `sspp/__init__.py` makes
    `from sspp import _sspp` (scripts/main.py:1) resolve to this module, with the reference's class names"""
    import subprocess
    import sys
    pkg = tmp_path / "sspp"
    pkg.mkdir()
    (pkg / "__init__.py").write_text("from sspp_b200 import _sspp, _tsp\n")
    code = ("from sspp import _sspp, _tsp\n"
            "assert all(hasattr(_sspp, n) for n in ('SamplingPathPlanner3', 'SamplingPathPlanner6', 'SamplingPathPlanner7', 'SamplingPathPlanner9', "
            "'Spline3', 'Spline6', 'Spline7', 'Spline9'))\n"
            "assert hasattr(_tsp, 'TaskSpacePlanner')\n"
            "print('ok')\n")
    env = dict(__import__("os").environ, PYTHONPATH=f"{tmp_path}:{ROOT}")
    out = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, env=env, timeout=300)
    assert out.returncode == 0 and out.stdout.strip().endswith("ok"), out.stderr[-500:]
