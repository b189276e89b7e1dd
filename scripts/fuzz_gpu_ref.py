"""Random Planner A / Planner B cases: the GPU (through the C ABI) against the REFERENCE'S OWN planner code (oracle/_ref, the
reference's headers compiled unmodified) on the same Philox normals - the randomised version of tests/test_gpu_ref.py.
usage: python scripts/fuzz_gpu_ref.py [n_cases] [seed]"""
import sys, json
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import numpy as np
from sspp_b200 import capi
from tests import ref_helpers
from tests.helpers import joint_case, task_case, armlab_case, Q_START, Q_END, ARMLAB_START, ARMLAB_END
from tests.test_ref_planner import check_iteration_against_reference, in_band, REL

def planner_a(case, seed):
    rec = ref_helpers.run_reference_a(case, seed=seed)
    h = capi.Handle(case.model, case.cfg); case.prepare(h)
    np.testing.assert_allclose(h.base()[0], rec["base"], rtol=REL, atol=1e-6)
    h.iterate(seed, 0, 1)
    r = h.results()
    np.testing.assert_allclose(h.params()[0].reshape(len(rec["params"]), -1), rec["params"], rtol=REL, atol=2e-6)
    diff = (r["status"][0] != 0) != rec["collides"]
    assert not (diff & ~in_band(rec["min_dist"])).any(), "collision flag differs outside the contact band"
    ok = ~diff
    np.testing.assert_allclose(r["L"][0][ok], rec["arc"][ok], rtol=REL)
    if not diff.any():
        assert int(r["n_success"][0]) == rec["n_success"] and int(r["best_idx"][0]) == rec["best_idx"]
    h.close()
    return dict(n_success=rec["n_success"], in_band=int(diff.sum()))

def planner_b(case, iters, seed):
    recs = ref_helpers.run_reference_b(case, iters, seed=seed)
    h = capi.Handle(case.model, case.cfg); case.prepare(h)
    amb = unres = 0
    for it, rec in enumerate(recs):
        if it > 0:
            h.set_distribution(rec["mean_in"][None], rec["sigma_in"][None])
        h.iterate(seed, it, 1)
        r = h.results(); idx, cost, w = h.elites(0); k = int(r["n_elite"][0]); gm, gs = h.get_distribution()
        got = dict(status=r["status"][0], L=r["L"][0], cost=r["cost"][0], n_success=int(r["n_success"][0]), elite_idx=idx[:k].astype(np.int64),
                   weights=w[:k], mean=gm[0], sigma=gs[0])
        a, u = check_iteration_against_reference(rec, got, case.cfg.w_collision)
        amb += a; unres += u
        if a or u: break
    h.close()
    return dict(n_success=[r["n_success"] for r in recs], in_band=amb, unresolved_order=unres)

n_cases = int(sys.argv[1]) if len(sys.argv) > 1 else 40
rng = np.random.default_rng(int(sys.argv[2]) if len(sys.argv) > 2 else 99)
bad = 0
for i in range(n_cases):
    kind = str(rng.choice(["A7", "A6", "B", "Bplanner", "Bstacking"]))
    ns = int(rng.choice([5, 17, 48, 100, 300, 1000])); cp = int(rng.choice([8, 16, 33, 40, 64]))
    seed = int(rng.integers(1, 2**60))
    try:
        if kind == "A7":
            case = joint_case(n_samples=ns, check_points=cp, n_points=int(rng.choice([7, 8, 10, 14])))
            case.start = (Q_START + rng.normal(0, 0.3, 7))[None]; case.end = (Q_END + rng.normal(0, 0.3, 7))[None]; case.joint_sigma = float(rng.choice([0.02, 0.08, 0.3]))
            out = planner_a(case, seed)
        elif kind == "A6":
            case = armlab_case(n_samples=ns, check_points=cp, n_points=int(rng.choice([7, 10, 12])))
            case.start = (ARMLAB_START + rng.normal(0, 0.4, 6))[None]; case.end = (ARMLAB_END + rng.normal(0, 0.4, 6))[None]
            out = planner_a(case, seed)
        else:
            scene, body = {"B": ("robocrane.xml", None), "Bplanner": ("planner.xml", "block1"), "Bstacking": ("stacking.xml", "block1")}[kind]
            s = np.array([rng.uniform(0.1, 0.6), rng.uniform(-0.5, 0.5), rng.uniform(0.12, 0.5), rng.uniform(-1.5, 1.5)])
            e = np.array([rng.uniform(0.1, 0.6), rng.uniform(-0.5, 0.5), rng.uniform(0.12, 0.5), rng.uniform(-1.5, 1.5)])
            kw = dict(n_samples=min(ns, 300), check_points=cp, n_points=int(rng.choice([3, 4, 5, 7])), start=s, end=e)
            case = task_case(scene, body, **kw) if body else task_case(**kw)
            out = planner_b(case, int(rng.integers(1, 8)), seed)
        ok = True
    except Exception as ex:
        ok = False; out = dict(error=repr(ex)[:300]); bad += 1
    print(json.dumps(dict(case=i, kind=kind, n_samples=ns, check_points=cp, n_points=int(case.cfg.n_points), ok=ok, **out)), flush=True)
print("FAILED CASES:", bad)
