"""Tile / launch-bound sweep of the fused iteration on cfg2 (and optionally cfg4): kernel time per iteration by CUDA events.
usage: python scripts/sweep_tiles.py [cfg2|cfg4] """
import sys, json, itertools
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import numpy as np, torch
from sspp_b200 import capi
from tests.helpers import joint_case, task_case, SEED

VARIANTS = [(256, 2), (256, 3), (320, 2), (320, 3), (448, 2), (384, 2), (128, 7), (192, 5)]
which = sys.argv[1] if len(sys.argv) > 1 else "cfg2"

def hint(variant, cpc, threads): return (variant + 1) | (cpc << 8) | (threads << 20)

def timeit(h, n_inner, reps=30):
    st = torch.cuda.Stream(); h.set_stream(st.cuda_stream)
    h.save_state()
    for i in range(3): h.restore_state(); h.iterate(SEED, i * n_inner, n_inner)
    torch.cuda.synchronize()
    ts = []
    for r in range(reps):
        h.restore_state()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(st); h.iterate(SEED, (3 + r) * n_inner, n_inner); b.record(st); torch.cuda.synchronize()
        ts.append(a.elapsed_time(b))
    return float(np.median(ts)), float(np.min(ts))

rows = []
if which == "cfg2":
    combos = [(0, 0, 0)]
    for v, (mt, mb) in enumerate(VARIANTS):
        for th in sorted({mt, min(mt, 256), min(mt, 320)}):
            slots = mb * 148
            base = -(-4096 // slots)
            for cpc in sorted({base, base + 1, base + 2, -(-4096 // (slots // 2)) if slots >= 2 else base}):
                combos.append((v + 1, cpc, th))
    for v1, cpc, th in combos:
        case = joint_case(n_samples=4096, check_points=64, do_update=True)
        case.cfg.tile_hint = 0 if v1 == 0 else hint(v1 - 1, cpc, th)
        try:
            h = capi.Handle(case.model, case.cfg)
        except Exception as e:
            rows.append(dict(variant=v1 - 1, cpc=cpc, threads=th, error=str(e)[:80])); continue
        case.prepare(h)
        med, mn = timeit(h, 1)
        d = h.dims; r = h.results()
        rows.append(dict(variant=d.variant, lb=VARIANTS[d.variant], threads=d.threads, cpc=d.cand_per_cta, ctas=d.ctas_per_query, grid=d.grid,
                         smem=d.smem_bytes, us_med=round(med * 1e3, 1), us_min=round(mn * 1e3, 1), gpts=round(4096 * 64 / med * 1e-6, 3),
                         succ=int(r["n_success"][0]), best=float(r["best_cost"][0])))
        print(json.dumps(rows[-1]), flush=True)
        h.close()
else:
    Q = 1024
    rng = np.random.default_rng(1000)
    s = np.column_stack([rng.uniform(0.2, 0.6, Q), rng.uniform(-0.4, 0.4, Q), rng.uniform(0.14, 0.4, Q), rng.uniform(-1.5, 1.5, Q)])
    e = np.column_stack([rng.uniform(0.2, 0.6, Q), rng.uniform(-0.4, 0.4, Q), rng.uniform(0.14, 0.4, Q), rng.uniform(-1.5, 1.5, Q)])
    for v1, th in [(0, 0)] + [(v + 1, th) for v, (mt, mb) in enumerate(VARIANTS) for th in (64, 96, 128, 192, 256) if th <= mt]:
        case = task_case(n_queries=Q, start=s, end=e)
        case.cfg.tile_hint = 0 if v1 == 0 else hint(v1 - 1, 0, th)
        try:
            h = capi.Handle(case.model, case.cfg)
        except Exception as ex:
            print(json.dumps(dict(variant=v1 - 1, threads=th, error=str(ex)[:80]))); continue
        case.prepare(h)
        med, mn = timeit(h, 60, reps=5)
        d = h.dims; r = h.results()
        print(json.dumps(dict(variant=d.variant, lb=VARIANTS[d.variant], threads=d.threads, smem=d.smem_bytes, ms_med=round(med, 3), ms_min=round(mn, 3),
                              gpts=round(Q * 17 * 40 * 60 / med * 1e-6, 3), succ=int(r["n_success"].sum()), best=float(r["best_cost"][:8].sum()))), flush=True)
        h.close()
