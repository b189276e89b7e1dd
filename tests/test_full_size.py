"""BASELINE.json's configurations at their full sizes.

cfg 3: stacking scene + 64 clutter boxes, 65536 candidates x 64 collocation points, SDF grid 128^3, 5 control points:
       bit-exact against the replica oracle (the SDF read path is cheap enough on the CPU to replay the whole
       population), plus the properties that do not depend on size (idempotence, a ranking that is a permutation,
       normalised monotone weights, the update a convex combination of the elites) and the population-sharded form.
cfg 4: 1024 independent ICRA-style queries in one launch: seeded subsample of queries against the replica.
(cfg 2 at full size: tests/test_gpu_parity.py::test_joint_iterations_match_oracles[4096-64-True].)"""
import numpy as np
import pytest

from tests.helpers import SEED, compare_iteration, sdf_case, task_case

pytestmark = pytest.mark.gpu
EXACT = ("status_mismatch", "elite_mismatch", "key_mismatch", "value_bits_mismatch", "param_bits_mismatch", "dist_bits_mismatch")


def test_cfg3_full_size_matches_replica_and_shards():
    from sspp_b200 import capi
    from sspp_b200.sharding import PopulationExchange
    N = 65536
    case = sdf_case(n_samples=N - 2, check_points=64, n_points=5, dims=(128, 128, 128))
    h = capi.Handle(case.model, case.cfg); case.prepare(h)
    rep = compare_iteration(h, case, iters=2)            # O1 leg on 512 seeded candidates of the 65536 per iteration
    for k in EXACT + ("o1_flag_mismatch",):
        assert rep[k] == 0, rep
    assert rep["o1_max_rel"] <= 1e-5 and rep["o1_checked"] >= 1000, rep
    assert 0 < min(rep["n_success"]) and max(rep["n_success"]) < N
    h.close()

    h = capi.Handle(case.model, case.cfg); case.prepare(h)
    twin = capi.Handle(case.model, case.cfg); case.prepare(twin)
    shard_case = sdf_case(n_samples=N, check_points=64, n_points=5, dims=(128, 128, 128))
    shard_case.cfg.shard_offset, shard_case.cfg.shard_global = 0, N
    hs = capi.Handle(shard_case.model, shard_case.cfg); shard_case.prepare(hs)
    px = PopulationExchange(hs)
    for it in range(3):
        mean0, _ = h.get_distribution()
        h.iterate(SEED, it, 1); twin.iterate(SEED, it, 1); px.iterate(SEED, it); px.check()
        r, rt = h.results(), twin.results()
        for k in ("L", "Cnf", "Cwf", "cost"):                      # idempotence: same counters, same bits on a second handle
            assert np.array_equal(r[k].view(np.uint32), rt[k].view(np.uint32))
        assert np.array_equal(r["status"], rt["status"])
        status, cost = r["status"][0], r["cost"][0]
        n_succ = int((status == 0).sum())
        assert r["n_success"][0] == n_succ
        k = int(r["n_elite"][0])
        idx, ecost, w = h.elites(0)
        idx64 = idx.astype(np.int64)
        assert np.array_equal(np.sort(idx64[:N]), np.arange(N))     # every candidate ranked exactly once
        top = idx64[:n_succ]
        assert (status[top] == 0).all() and (np.diff(cost[top]) >= 0).all()
        ties = np.diff(cost[top]) == 0
        assert (np.diff(top)[ties] > 0).all()                       # ties break by candidate index
        assert np.array_equal(ecost[:n_succ].view(np.uint32), cost[top].view(np.uint32))
        assert abs(float(w[:k].astype(np.float64).sum()) - 1.0) < 1e-4 and (np.diff(w[:k]) <= 0).all() and (w[:k] > 0).all()
        assert r["best_idx"][0] == idx[0] and r["best_cost"][0] == cost[idx[0]]
        P = h.params()[0].reshape(N, -1)
        mean1, sigma1 = h.get_distribution()
        lr = case.cfg.mean_lr
        target = (mean1[0].astype(np.float64).ravel() - (1 - lr) * mean0[0].astype(np.float64).ravel()) / lr
        El = P[idx64[:k]].astype(np.float64)
        assert np.allclose(target, (w[:k, None].astype(np.float64) * El).sum(0), rtol=5e-4, atol=5e-5)
        assert (sigma1 >= case.cfg.stddev_min * (1 - 1e-6)).all() and (sigma1 <= case.cfg.stddev_max * (1 + 1e-6)).all()
        mi, mc, mw = px.h.merged(N)                                 # population shard, exchange fused into the kernel (one rank)
        assert np.array_equal(mi, idx) and np.array_equal(mw[:k].view(np.uint32), w[:k].view(np.uint32))
        ms, ss = px.h.get_distribution()
        assert np.array_equal(ms, mean1) and np.array_equal(ss, sigma1)
    psi_col, _ = h.collocation()
    pts = h.base()[0].copy(); pts[1:-1] = P[idx64[0]].reshape(-1, 4)
    assert np.count_nonzero(h.eval_points((psi_col.astype(np.float64) @ pts.astype(np.float64)).astype(np.float32))) <= 1
    for x in (h, twin, px.h):
        x.close()


def test_cfg4_full_batch_queries_match_replica():
    from sspp_b200 import capi
    rng = np.random.default_rng(2)
    Q = 1024
    s = np.column_stack([rng.uniform(0.2, 0.6, Q), rng.uniform(-0.4, 0.4, Q), rng.uniform(0.14, 0.4, Q), rng.uniform(-1.5, 1.5, Q)])
    e = np.column_stack([rng.uniform(0.2, 0.6, Q), rng.uniform(-0.4, 0.4, Q), rng.uniform(0.14, 0.4, Q), rng.uniform(-1.5, 1.5, Q)])
    case = task_case(n_queries=Q, start=s, end=e)
    for q in rng.choice(Q, 3, replace=False):
        h = capi.Handle(case.model, case.cfg); case.prepare(h)
        rep = compare_iteration(h, case, iters=12, query=int(q))
        for k in EXACT + ("o1_flag_mismatch",):
            assert rep[k] == 0, (int(q), rep)
        assert rep["o1_max_rel"] <= 1e-5, (int(q), rep)
        h.close()
    h = capi.Handle(case.model, case.cfg); case.prepare(h)
    h.iterate(SEED, 0, 60)                                           # a whole 60-iteration plan of every query in one launch
    r = h.results()
    h2 = capi.Handle(case.model, case.cfg); case.prepare(h2)
    for it in range(60):
        h2.iterate(SEED, it, 1)
    r2 = h2.results()
    for k in ("cost", "status", "n_success", "best_idx"):
        assert np.array_equal(r[k], r2[k]), k
    m, sg = h.get_distribution(); m2, s2 = h2.get_distribution()
    assert np.array_equal(m, m2) and np.array_equal(sg, s2)
    assert (r["n_success"] > 0).mean() > 0.5
    h.close(); h2.close()
