"""Launch overhead vs per-iteration time: one launch running n_inner iterations (cfg2)."""
import sys, json
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import numpy as np, torch
from sspp_b200 import capi
from tests.helpers import joint_case, SEED
case = joint_case(n_samples=4096, check_points=64, do_update=True)
if len(sys.argv) > 1: case.cfg.tile_hint = int(sys.argv[1], 0)
h = capi.Handle(case.model, case.cfg); case.prepare(h); h.save_state()
st = torch.cuda.Stream(); h.set_stream(st.cuda_stream)
for n in (1, 2, 4, 8, 16):
    ts = []
    for r in range(25):
        h.restore_state()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(st); h.iterate(SEED, 100 * r, n); b.record(st); torch.cuda.synchronize()
        ts.append(a.elapsed_time(b) * 1e3)
    # back-to-back launches, no sync in between
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record(st)
    for r in range(50): h.iterate(SEED, 100 * r, n)
    b.record(st); torch.cuda.synchronize()
    print(json.dumps(dict(n_inner=n, us_single_med=round(float(np.median(ts)), 1), us_single_min=round(float(np.min(ts)), 1),
                          us_b2b=round(a.elapsed_time(b) * 1e3 / 50, 1))), flush=True)
