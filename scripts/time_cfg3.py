"""cfg 3 timing on one GPU: 65536 candidates x 64 points, SDF 128^3, K=5 (3 vias), plus the 8192-candidate shard shape."""
import sys, json, os
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import numpy as np, torch
from sspp_b200 import capi
from tests.helpers import sdf_case, SEED
out = {}
for n in (8190, 65534):
    case = sdf_case(n_samples=n, check_points=64, n_points=5, dims=(128, 128, 128))
    case.cfg.tile_hint = int(os.environ.get('TILE_HINT', '0'), 0)
    h = capi.Handle(case.model, case.cfg); case.prepare(h); h.save_state()
    st = torch.cuda.Stream(); h.set_stream(st.cuda_stream)
    for i in range(3): h.restore_state(); h.iterate(SEED, i, 1)
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record(st)
    for i in range(20): h.restore_state(); h.iterate(SEED, 3 + i, 1)
    b.record(st); torch.cuda.synchronize()
    ms = a.elapsed_time(b) / 20
    d = h.dims
    out[f"sdf_{n + 2}x64"] = dict(ms=ms, pts_per_s=(n + 2) * 64 / ms * 1e3, sphere_samples_per_s=(n + 2) * 64 * 27 / ms * 1e3,
                                  cand_per_cta=d.cand_per_cta, variant=d.variant, grid=d.grid, ctas=d.ctas_per_query, smem=d.smem_bytes, n_success=int(h.results()["n_success"][0]))
    h.close()
print(json.dumps(out, indent=1))
