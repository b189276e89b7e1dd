"""Random shapes through the additive C-ABI entries: ssppcu_plan_step == set_queries + noise/reset + iterate (bit for bit),
success rows == the converged candidates' sampled rows, candidate_ctrl == init spline with the row written in,
check_points consistent with eval_points.  usage: python scripts/fuzz_gpu_api.py [n_cases] [seed]"""
import sys, json
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import numpy as np
from sspp_b200 import capi
from tests.helpers import joint_case, task_case, armlab_case, Q_START, Q_END, ARMLAB_START, ARMLAB_END

n_cases = int(sys.argv[1]) if len(sys.argv) > 1 else 60
rng = np.random.default_rng(int(sys.argv[2]) if len(sys.argv) > 2 else 5)
bad = 0
for i in range(n_cases):
    kind = str(rng.choice(["joint", "armlab", "task", "task_q"]))
    ns = int(rng.choice([5, 30, 100, 500, 2000])); cp = int(rng.choice([8, 16, 40, 64])); upd = bool(rng.integers(0, 2)); nq = 1
    if kind == "joint":
        nq = int(rng.choice([1, 1, 2, 3]))
        case = joint_case(n_samples=ns, check_points=cp, n_points=int(rng.choice([7, 10, 14])), do_update=upd, n_queries=nq)
        case.start = Q_START + rng.normal(0, 0.2, (nq, 7)); case.end = Q_END + rng.normal(0, 0.2, (nq, 7))
    elif kind == "armlab":
        case = armlab_case(n_samples=ns, check_points=cp, n_points=int(rng.choice([7, 10])), do_update=upd)
        case.start = (ARMLAB_START + rng.normal(0, 0.3, 6))[None]; case.end = (ARMLAB_END + rng.normal(0, 0.3, 6))[None]
    else:
        nq = 1 if kind == "task" else int(rng.choice([2, 7, 40, 300]))
        s = np.column_stack([rng.uniform(0.2, 0.6, nq), rng.uniform(-0.4, 0.4, nq), rng.uniform(0.14, 0.4, nq), rng.uniform(-1.5, 1.5, nq)])
        e = np.column_stack([rng.uniform(0.2, 0.6, nq), rng.uniform(-0.4, 0.4, nq), rng.uniform(0.14, 0.4, nq), rng.uniform(-1.5, 1.5, nq)])
        case = task_case(n_samples=min(ns, 500) if nq > 1 else ns, check_points=cp, n_points=int(rng.choice([3, 5])), n_queries=nq, start=s, end=e)
        case.cfg.do_update = int(upd)
    seed = int(rng.integers(1, 2**60)); n_it = int(rng.integers(1, 4))
    try:
        joint = case.cfg.mode == 0
        a = capi.Handle(case.model, case.cfg); b = capi.Handle(case.model, case.cfg); b.set_success_rows(True)
        for rep in range(2):
            case.prepare(a); a.iterate(seed, 7 * rep, n_it)
            b.plan_step(case.start, case.end, case.joint_sigma if joint else case.sigma0, case.limits if joint else None, seed, 7 * rep, n_it)
            ra, rb = a.results(), b.results()
            for k in ("L", "Cnf", "Cwf", "cost"):
                assert np.array_equal(ra[k].view(np.uint32), rb[k].view(np.uint32)), k
            for k in ("status", "n_success", "n_elite", "best_idx"):
                assert np.array_equal(ra[k], rb[k]), k
            assert np.array_equal(a.params(), b.params()) and np.array_equal(a.base(), b.base())
            for x, y in zip(a.get_distribution() + tuple(a.best()), b.get_distribution() + tuple(b.best())):
                assert np.array_equal(x, y)
            P = b.params()
            for q in range(min(nq, 3)):
                rows, idx = b.success_rows(q)
                ok = np.flatnonzero(rb["status"][q] == 0)
                assert np.array_equal(idx, ok) and np.array_equal(rows, P[q].reshape(len(P[q]), -1)[ok])
                if len(ok):
                    c = b.candidate_ctrl(int(ok[len(ok) // 2]), q)
                    fb = 1 if not joint else case.cfg.degree
                    assert np.array_equal(c[fb:fb + P.shape[2]], P[q, ok[len(ok) // 2]])
        # check_points vs eval_points on random configurations (primitives mode)
        D = case.cfg.dof
        pts = (rng.uniform(-2, 2, (500, D)) if joint else np.column_stack([rng.uniform(0, 0.7, 500), rng.uniform(-0.7, 0.7, 500), rng.uniform(0, 0.6, 500), rng.uniform(-1.6, 1.6, 500)])).astype(np.float32)
        g1, g2, dist = b.check_points(pts); cost = b.eval_points(pts)
        assert ((g1 >= 0) <= (cost != 0)).all() and (dist[g1 >= 0] < -1e-3).all() and (g1[g1 >= 0] < g2[g1 >= 0]).all()
        a.close(); b.close(); ok_case = True; err = None
    except Exception as ex:
        ok_case = False; err = repr(ex)[:300]; bad += 1
    print(json.dumps(dict(case=i, kind=kind, n_samples=ns, check_points=cp, queries=nq, update=upd, iters=n_it, ok=ok_case, error=err)), flush=True)
print("FAILED CASES:", bad)
