"""Task-scene population of 8192 candidates x 64 points (one shard of the sharded cfg-3-sized population, primitives mode):
candidates per tile vs time per iteration; tiles beyond the first wave are handed out dynamically."""
import sys, json
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import torch
from sspp_b200 import capi
from tests.helpers import task_case, SEED

def hint(variant, cpc, threads): return (variant + 1) | (cpc << 8) | (threads << 20)

for cpc in (0, 28, 20, 14, 10, 7, 5):
    case = task_case(n_samples=8190, check_points=64, n_points=5)
    case.cfg.tile_hint = hint(2, cpc, 320) if cpc else 0
    h = capi.Handle(case.model, case.cfg); case.prepare(h); h.save_state()
    st = torch.cuda.Stream(); h.set_stream(st.cuda_stream)
    for i in range(5): h.iterate(SEED, i, 1, restored=True)
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record(st)
    for i in range(100): h.iterate(SEED, 5 + i, 1, restored=True)
    b.record(st); torch.cuda.synchronize()
    d = h.dims
    print(json.dumps(dict(cpc=d.cand_per_cta, tiles=d.ctas_per_query, grid=d.grid, us=round(a.elapsed_time(b) / 100 * 1e3, 2),
                          succ=int(h.results()["n_success"][0]))), flush=True)
    h.close()
