"""cfg4 (one CTA per query, 17 candidates x 41 points = 697 items per iteration): thread counts that change the number of item
rounds per iteration (256 threads: 3 rounds, the last 72 % full; 352: 2 rounds, 99 % full).  usage: python scripts/sweep_cfg4_threads.py"""
import sys, json
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import numpy as np, torch
from sspp_b200 import capi
from tests.helpers import task_case, SEED
VARIANTS = [(256, 2), (256, 3), (320, 2), (320, 3), (448, 2), (384, 2), (128, 7), (192, 5)]
def hint(variant, cpc, threads): return (variant + 1) | (cpc << 8) | (threads << 20)
Q = 1024
rng = np.random.default_rng(1000)
s = np.column_stack([rng.uniform(0.2, 0.6, Q), rng.uniform(-0.4, 0.4, Q), rng.uniform(0.14, 0.4, Q), rng.uniform(-1.5, 1.5, Q)])
e = np.column_stack([rng.uniform(0.2, 0.6, Q), rng.uniform(-0.4, 0.4, Q), rng.uniform(0.14, 0.4, Q), rng.uniform(-1.5, 1.5, Q)])
ref = None
for v1, th in [(0, 0), (5, 352), (6, 352), (5, 384), (6, 384), (3, 320), (5, 256), (6, 256), (5, 704 // 2), (5, 448)]:
    case = task_case(n_queries=Q, start=s, end=e)
    case.cfg.tile_hint = 0 if v1 == 0 else hint(v1 - 1, 0, th)
    try:
        h = capi.Handle(case.model, case.cfg)
    except Exception as ex:
        print(json.dumps(dict(variant=v1 - 1, threads=th, error=str(ex)[:100]))); continue
    st = torch.cuda.Stream(); h.set_stream(st.cuda_stream)
    case.prepare(h); h.save_state()
    ts = []
    for r in range(8):
        h.restore_state()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(st); h.iterate(SEED, 0, 60); b.record(st); torch.cuda.synchronize()
        if r >= 2: ts.append(a.elapsed_time(b))
    res = h.results(); d = h.dims
    sig = (int(res["n_success"].sum()), float(np.nansum(np.where(np.isfinite(res["best_cost"]), res["best_cost"], 0))))
    ref = ref or sig
    print(json.dumps(dict(variant=d.variant, lb=VARIANTS[d.variant], threads=d.threads, smem=d.smem_bytes, ms_med=round(float(np.median(ts)), 3),
                          ms_min=round(min(ts), 3), same_result=(sig == ref))), flush=True)
    h.close()
