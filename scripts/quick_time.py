"""First timing probe: config 2 (robocrane7, 4096 x 64 x 10) and the ICRA batch, CUDA events on the launch stream."""
import sys, json
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import numpy as np, torch
from sspp_b200 import capi
from tests.helpers import joint_case, task_case, SEED

def timeit(h, n, reps=5):
    st = torch.cuda.Stream()
    h.set_stream(st.cuda_stream)
    h.iterate(SEED, 0, 3); torch.cuda.synchronize()
    best = 1e9
    for r in range(reps):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(st); h.iterate(SEED, 3 + r * n, n); b.record(st); torch.cuda.synchronize()
        best = min(best, a.elapsed_time(b) / n)
    return best

out = {}
for ns, M in ((4096, 64), (16384, 64), (65536, 64)):
    case = joint_case(n_samples=ns, check_points=M, do_update=True)
    h = capi.Handle(case.model, case.cfg); case.prepare(h)
    ms = timeit(h, 50)
    d = h.dims
    out[f"joint_{ns}x{M}"] = dict(ms=ms, pts_per_s=ns * M / ms * 1e3, threads=d.threads, cpc=d.cand_per_cta, ctas=d.ctas_per_query, smem=d.smem_bytes, succ=int(h.results()["n_success"][0]))
    h.close()
for Q in (1024, 8192):
    rng = np.random.default_rng(1)
    s = np.column_stack([rng.uniform(0.2, 0.6, Q), rng.uniform(-0.4, 0.4, Q), rng.uniform(0.14, 0.4, Q), rng.uniform(-1.5, 1.5, Q)])
    e = np.column_stack([rng.uniform(0.2, 0.6, Q), rng.uniform(-0.4, 0.4, Q), rng.uniform(0.14, 0.4, Q), rng.uniform(-1.5, 1.5, Q)])
    case = task_case(n_queries=Q, start=s, end=e)
    h = capi.Handle(case.model, case.cfg); case.prepare(h)
    ms = timeit(h, 60, reps=3)
    d = h.dims
    out[f"task_Q{Q}"] = dict(ms_per_iter=ms, queries_per_s=Q / (ms * 60) * 1e3, pts_per_s=Q * 17 * 40 / ms * 1e3, threads=d.threads, cpc=d.cand_per_cta, ctas=d.ctas_per_query, smem=d.smem_bytes)
    h.close()
print(json.dumps(out, indent=1))
