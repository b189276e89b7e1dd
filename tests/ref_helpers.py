"""Reference-side runs for the parity tests: the reference's own planners (oracle/_ref/libref.so = its headers compiled
unmodified, oracle/ref_driver.cpp) driven with the Philox normals the GPU uses, laid out the way the GPU lays candidates out.

While producing them every quantity is cross-checked against oracle O1 (FP64, same normals) to 1e-12, so a test that uses
these records compares the GPU / the FP32 replica with the REFERENCE'S code, and O1 with it on the way."""
from __future__ import annotations

import numpy as np

from oracle import pyoracle, pyref
from oracle.flatmodel import FlatModel
from tests.helpers import SEED

TOL = 1e-12


def flat_for(case) -> FlatModel:
    stem = case.mjcf.stem
    try:
        return FlatModel.load(stem)
    except FileNotFoundError:
        return FlatModel.from_capi(case.model)


def tsp_args_from(case) -> pyref.TspArgs:
    """constructor arguments of tsp::TaskSpacePlanner that give the reference the effective configuration of a task_case
    (tests/helpers.py) — through the adapter's own quirks (tsp.h:53: stddev_initial lands in dist_.z_min, sigma0 stays 0.3,
    the evaluator keeps its defaults)"""
    c = case.cfg
    assert c.mode == 1 and case.sigma0 == 0.3 and c.eval_zmin == 0.0 and c.floor_margin == 0.01 and c.floor_scale == 10.0, \
        "not reachable through the reference's adapter"
    a = pyref.TspArgs(c.dist_zmin, c.stddev_min, c.stddev_max, c.inc, c.dec, c.elite_fraction, c.n_samples, c.check_points, c.n_points,
                      c.w_collision, c.sample_zmin)
    for d in range(4):
        a.lo[d], a.hi[d] = c.lo[d], c.hi[d]
    a.sigma_floor, a.var_ema_beta, a.mean_lr, a.floor_margin, a.floor_penalty_scale = c.sigma_floor, c.var_beta, c.mean_lr, c.floor_margin, c.floor_scale
    return a


def run_reference_b(case, iters: int, seed: int = SEED, query: int = 0):
    """`iters` calls of tsp::TaskSpacePlanner::plan (cold start, then iterate = true) on case's query.  Returns one record per
    iteration in the GPU's candidate layout (slot 0 mean, slot 1 forwarded best or ABSENT, slots 2.. samples):
      mean_in / sigma_in / best_in / best_valid_in   state the iteration sampled from  [Kv][4]
      params [n_cand][Kv*4], L, Cnf, Cwf, cost, status (0 converged, 1 failed, 2 absent), min_dist (O1, for the contact band)
      elite_idx (candidate slots, best first), weights, n_success, best_idx
      mean_out / sigma_out                           state after update + adapt"""
    c = case.cfg
    fm = flat_for(case)
    ref = pyref.RefTsp(pyref.RefModel(fm), case.body, tsp_args_from(case))
    o1 = pyoracle.O1(fm, 1, 4, case.body)
    oc = pyoracle.o1_cfg_from(c)
    Kv, K, N = c.n_points - 2, c.n_points, c.n_samples + 2
    start, end = np.asarray(case.start[query], float), np.asarray(case.end[query], float)
    base = np.array([(1 - i / (K - 1)) * start + (i / (K - 1)) * end for i in range(K)])
    lo, hi = np.array([c.lo[d] for d in range(4)]), np.array([c.hi[d] for d in range(4)])
    mean = base[1:1 + Kv].copy()
    mean[:, 2] = np.maximum(np.maximum(mean[:, 2], c.sample_zmin), c.dist_zmin)
    mean = np.clip(mean, lo, hi)
    sigma = np.full((Kv, 4), max(min(max(case.sigma0, c.stddev_min), c.stddev_max), c.sigma_floor))
    best, best_valid = np.zeros((Kv, 4)), 0
    spr = 3 if c.lo[3] == c.hi[3] else 4
    out = []
    for it in range(iters):
        pyref.arm_tape(pyref.MODE_TASK, seed, it, query, 2, spr, max(Kv, 1))
        r = ref.plan(start, end, it > 0)
        # --- O1 on the same normals, GPU layout ---
        params = np.zeros((N, Kv * 4))
        m0 = mean.copy(); m0[:, 2] = np.maximum(m0[:, 2], c.sample_zmin)               # tsp_planner.h:83-84
        params[0] = m0.ravel()
        params[1] = best.ravel() if best_valid else mean.ravel()
        params[2:] = o1.sample(oc, seed, it, query, 2, c.n_samples, mean, sigma)
        s = o1.score(oc, base, params)
        status = s["status"].astype(np.int32).copy()
        if not best_valid:
            status[1] = 2
        slots = np.array([0] + ([1] if best_valid else []) + list(range(2, N)))          # reference seed order -> GPU slots
        assert r["n"] == len(slots)
        np.testing.assert_allclose(r["vias"].reshape(len(slots), -1), params[slots], rtol=0, atol=TOL)
        assert np.array_equal(r["status"], status[slots]), "collision flags: reference code vs O1"
        for k in ("L", "Cnf", "Cwf"):
            np.testing.assert_allclose(r[k], s[k][slots], rtol=TOL, atol=TOL)
        cost = s["L"] + c.w_collision * s["Cwf"]
        idx, w, ns = o1.elites(cost, status, c.elite_fraction)
        rec = dict(mean_in=mean.copy(), sigma_in=sigma.copy(), best_in=best.copy(), best_valid_in=best_valid, params=params, L=s["L"],
                   Cnf=s["Cnf"], Cwf=s["Cwf"], cost=cost, status=status, min_dist=s["min_dist"], elite_idx=idx.copy(), weights=w.copy(),
                   n_success=ns, best_idx=int(idx[0]) if ns else -1)
        # the reference's own elites / weights over ITS successes_, mapped to slots
        rs = np.where(r["status"] == 0)[0]
        if len(rs):
            rcost = r["L"][rs] + c.w_collision * r["Cwf"][rs]
            ri, rw = pyref.elites(rcost, c.elite_fraction)
            assert np.array_equal(slots[rs[ri]], idx), "elite indices: reference code vs O1"
            np.testing.assert_allclose(rw, w, rtol=0, atol=1e-15)
        m1, s1 = o1.update(oc, mean, sigma, params[idx], w)
        mean, sigma = m1.reshape(Kv, 4), s1.reshape(Kv, 4)
        if ns:
            best, best_valid = params[idx[0]].reshape(Kv, 4).copy(), 1
        if Kv >= 1:                                                                       # all the adapter exposes: the first via
            np.testing.assert_allclose(r["mean0"], mean[0], rtol=0, atol=TOL)
            np.testing.assert_allclose(r["sigma0"], sigma[0], rtol=0, atol=TOL)
        rec.update(mean_out=mean.copy(), sigma_out=sigma.copy(), path=r["path"])
        out.append(rec)
    return out


def run_reference_a(case, seed: int = SEED, it: int = 0, query: int = 0):
    """One sspp::SamplingPathPlanner<dof>::plan-shaped pass (sspp.h:194-232) candidate by candidate: sampleWithNoise,
    checkCollision, computeArcLength, then plan() itself for the success list and the best path.  Cross-checked against O1."""
    c = case.cfg
    fm = flat_for(case)
    ref = pyref.RefSspp(pyref.RefModel(fm), c.dof)
    o1 = pyoracle.O1(fm, 0, c.dof)
    oc = pyoracle.o1_cfg_from(c)
    K, M, n, p = c.n_points, c.check_points, c.n_samples, c.degree
    start, end = np.asarray(case.start[query], float), np.asarray(case.end[query], float)
    pyref.arm_tape(pyref.MODE_JOINT, seed, it, query, 0)
    r = ref.candidates(start, end, case.joint_sigma, case.limits, n, M, K)
    V = np.array([(1 - i / (K - 1)) * start + (i / (K - 1)) * end for i in range(K)])
    base = o1.interpolate(V, p)
    ctrl0, knots = ref.init_path(start, end, K)
    np.testing.assert_allclose(ctrl0, base, rtol=0, atol=TOL)                     # initializePath vs O1's interpolation
    np.testing.assert_allclose(knots, o1.knots(K, p), rtol=0, atol=1e-15)
    nf = K - 2 * p
    mean = base[p:p + nf]; sigma = np.tile(case.joint_sigma * np.asarray(case.limits, float), (nf, 1))
    params = o1.sample(oc, seed, it, query, 0, n, mean, sigma)
    np.testing.assert_allclose(r["ctrl"][:, p:p + nf].reshape(n, -1), params, rtol=0, atol=TOL)
    fixed = np.delete(r["ctrl"], np.s_[p:p + nf], axis=1)
    assert (fixed == np.delete(ctrl0, np.s_[p:p + nf], axis=0)[None]).all()          # only columns p .. K-p-1 are perturbed (sspp.h:123)
    s = o1.score(oc, base, params)
    assert np.array_equal(r["collides"], s["status"] != 0), "collision flags: reference code vs O1"
    np.testing.assert_allclose(r["arc"], s["L"], rtol=TOL, atol=TOL)
    pyref.arm_tape(pyref.MODE_JOINT, seed, it, query, 0)
    pl = ref.plan(start, end, case.joint_sigma, case.limits, n, M, K)
    ok = np.where(~r["collides"])[0]
    assert pl["n_success"] == len(ok) and pl["found"] == (len(ok) > 0)
    np.testing.assert_allclose(pl["succ_ctrl"], r["ctrl"][ok], rtol=0, atol=0)        # successes in candidate order (one thread)
    best_idx = int(ok[np.argmin(r["arc"][ok])]) if len(ok) else -1
    if len(ok):
        np.testing.assert_allclose(pl["best_ctrl"], r["ctrl"][best_idx], rtol=0, atol=0)
    return dict(base=base, params=params, collides=r["collides"], arc=r["arc"], min_dist=s["min_dist"], best_idx=best_idx, n_success=len(ok),
                ctrl=r["ctrl"])
