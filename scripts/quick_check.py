"""Experiment loop helper: parity smoke against the replica oracle + cfg2 / cfg4 kernel times for the library in SSPP_CUDA_LIB."""
import sys, json, os
from pathlib import Path
ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
import numpy as np, torch
from sspp_b200 import capi
from tests.helpers import joint_case, task_case, compare_iteration, SEED

def t_us(h, n_inner, reps):
    st = torch.cuda.Stream(); h.set_stream(st.cuda_stream); h.save_state()
    for i in range(3): h.restore_state(); h.iterate(SEED, i * n_inner, n_inner)
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record(st)
    for r in range(reps): h.iterate(SEED, (3 + r) * n_inner, n_inner, restored=True)
    b.record(st); torch.cuda.synchronize()
    return a.elapsed_time(b) * 1e3 / reps

out = dict(lib=os.environ.get("SSPP_CUDA_LIB", "default"))
case = joint_case(n_samples=64, check_points=16, do_update=True)
h = capi.Handle(case.model, case.cfg); case.prepare(h)
rep = compare_iteration(h, case, seed=SEED, iters=2); h.close()
out["parity_joint"] = (rep["status_mismatch"], rep["elite_mismatch"])
case = joint_case(n_samples=3000, check_points=16, do_update=True)
h = capi.Handle(case.model, case.cfg); case.prepare(h)
rep = compare_iteration(h, case, seed=SEED, iters=2, check_o1=False); h.close()
out["parity_joint_3000"] = (rep["status_mismatch"], rep["elite_mismatch"], rep["dist_bits_mismatch"], rep["n_success"])
case = task_case(n_queries=4); h = capi.Handle(case.model, case.cfg); case.prepare(h)
rep = compare_iteration(h, case, seed=SEED, iters=2); h.close()
out["parity_task"] = (rep["status_mismatch"], rep["elite_mismatch"])
hint = int(os.environ.get("SSPP_TILE_HINT", "0"), 0)
case = joint_case(n_samples=4096, check_points=64, do_update=True); case.cfg.tile_hint = hint
h = capi.Handle(case.model, case.cfg); case.prepare(h)
out["cfg2_us"] = round(t_us(h, 1, 200), 1); out["cfg2_gpts"] = round(4096 * 64 / out["cfg2_us"] * 1e-3, 3)
r = h.results(); out["cfg2_succ"] = int(r["n_success"][0])
import time
for i in range(5): case.prepare(h); h.iterate(SEED, i, 1); h.results(); h.best()
t0 = time.perf_counter()
for i in range(200): case.prepare(h); h.iterate(SEED, 10 + i, 1); r = h.results(); b = h.best()
out["cfg2_e2e_us"] = round((time.perf_counter() - t0) / 200 * 1e6, 1); out["cfg2_e2e_succ"] = int(r["n_success"][0])
h.close()
Q = 1024; rng = np.random.default_rng(1000)
s = np.column_stack([rng.uniform(0.2, 0.6, Q), rng.uniform(-0.4, 0.4, Q), rng.uniform(0.14, 0.4, Q), rng.uniform(-1.5, 1.5, Q)])
e = np.column_stack([rng.uniform(0.2, 0.6, Q), rng.uniform(-0.4, 0.4, Q), rng.uniform(0.14, 0.4, Q), rng.uniform(-1.5, 1.5, Q)])
case = task_case(n_queries=Q, start=s, end=e); h = capi.Handle(case.model, case.cfg); case.prepare(h)
out["cfg4_ms"] = round(t_us(h, 60, 10) * 1e-3, 3); out["cfg4_gpts"] = round(Q * 17 * 40 * 60 / out["cfg4_ms"] * 1e-6, 3); h.close()
print(json.dumps(out))
