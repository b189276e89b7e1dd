"""cfg2 with tiles smaller than one per CTA (tiles beyond the first wave are handed out dynamically):
kernel time per iteration, back to back on one stream (the bench's measure).  usage: python scripts/sweep_small_tiles.py"""
import sys, json
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import numpy as np, torch
from sspp_b200 import capi
from tests.helpers import joint_case, SEED

def hint(variant, cpc, threads): return (variant + 1) | (cpc << 8) | (threads << 20)

for cpc in (0, 14, 12, 10, 9, 8, 7, 6, 5, 4, 3):
    case = joint_case(n_samples=4096, check_points=64, do_update=True)
    case.cfg.tile_hint = hint(2, cpc, 320) if cpc else 0
    h = capi.Handle(case.model, case.cfg); case.prepare(h); h.save_state()
    st = torch.cuda.Stream(); h.set_stream(st.cuda_stream)
    for i in range(5): h.iterate(SEED, i, 1, restored=True)
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record(st)
    for i in range(200): h.iterate(SEED, 5 + i, 1, restored=True)
    b.record(st); torch.cuda.synchronize()
    d = h.dims; r = h.results()
    print(json.dumps(dict(cpc=d.cand_per_cta, tiles=d.ctas_per_query, grid=d.grid, smem=d.smem_bytes, us=round(a.elapsed_time(b) / 200 * 1e3, 2),
                          succ=int(r["n_success"][0]))), flush=True)
    h.close()
