"""Random planner shapes, scenes, start / goal pairs and seeds through the fused iteration on the GPU: every output bit for bit
against the replica, collision flags and costs against the FP64 oracle (tests.helpers.compare_iteration).
usage: python scripts/fuzz_gpu.py [n_cases] [seed]"""
import sys, json
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import numpy as np
from sspp_b200 import capi
from tests.helpers import joint_case, task_case, armlab_case, sdf_case, compare_iteration, Q_START, Q_END, ARMLAB_START, ARMLAB_END

n_cases = int(sys.argv[1]) if len(sys.argv) > 1 else 40
rng = np.random.default_rng(int(sys.argv[2]) if len(sys.argv) > 2 else 2026)
bad = 0
for i in range(n_cases):
    kind = rng.choice(["joint", "armlab", "task", "task_planner", "task_stacking", "sdf"], p=[0.3, 0.2, 0.15, 0.1, 0.1, 0.15])
    ns = int(rng.choice([5, 17, 62, 100, 254, 600, 1500, 4094]))
    cp = int(rng.choice([8, 16, 33, 40, 64, 100]))
    upd = bool(rng.integers(0, 2))
    iters = int(rng.integers(1, 4))
    if kind == "joint":
        K = int(rng.choice([7, 8, 10, 14]))
        case = joint_case(n_samples=ns, check_points=cp, n_points=K, do_update=upd, w_collision=float(rng.choice([0.0, 1.0])), limit_scale=float(rng.choice([0.0, 2.0])))
        case.start = (Q_START + rng.normal(0, 0.3, 7))[None]; case.end = (Q_END + rng.normal(0, 0.3, 7))[None]
        case.joint_sigma = float(rng.choice([0.02, 0.08, 0.3]))
    elif kind == "armlab":
        K = int(rng.choice([7, 10, 12]))
        case = armlab_case(n_samples=ns, check_points=cp, n_points=K, do_update=upd, w_collision=float(rng.choice([0.0, 1.0])))
        case.start = (ARMLAB_START + rng.normal(0, 0.4, 6))[None]; case.end = (ARMLAB_END + rng.normal(0, 0.4, 6))[None]
    elif kind == "sdf":
        case = sdf_case(n_samples=min(ns, 1500), check_points=cp, n_points=int(rng.choice([3, 5, 6])), dims=(48, 48, 48))
    else:
        scene, body = {"task": ("robocrane.xml", None), "task_planner": ("planner.xml", "block1"), "task_stacking": ("stacking.xml", "block1")}[kind]
        s = np.array([rng.uniform(0.1, 0.6), rng.uniform(-0.5, 0.5), rng.uniform(0.12, 0.5), rng.uniform(-1.5, 1.5)])
        e = np.array([rng.uniform(0.1, 0.6), rng.uniform(-0.5, 0.5), rng.uniform(0.12, 0.5), rng.uniform(-1.5, 1.5)])
        kw = dict(n_samples=ns, check_points=cp, n_points=int(rng.choice([3, 4, 5, 7])), start=s, end=e)
        case = task_case(scene, body, **kw) if body else task_case(**kw)
        case.cfg.do_update = int(upd)
    seed = int(rng.integers(1, 2**62))
    try:
        h = capi.Handle(case.model, case.cfg); case.prepare(h)
        rep = compare_iteration(h, case, seed=seed, iters=iters, o1_subsample=256)
        h.close()
    except Exception as ex:
        rep = dict(error=str(ex)[:200])
    fails = {k: v for k, v in rep.items() if k.endswith("_mismatch") and v} if "error" not in rep else rep
    ok = not fails and rep.get("o1_max_rel", 0) < 2e-4
    bad += not ok
    print(json.dumps(dict(case=i, kind=str(kind), n_samples=ns, check_points=cp, n_points=int(case.cfg.n_points), update=upd, iters=iters, ok=ok, fails=fails,
                          o1_max_rel=rep.get("o1_max_rel"), ambiguous=rep.get("o1_flag_ambiguous"), n_success=rep.get("n_success"))), flush=True)
print("FAILED CASES:", bad)
