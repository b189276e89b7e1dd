"""Phase timeline of the cooperative iteration kernel (cfg2) from the per-CTA %globaltimer probe."""
import sys, json, os
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import numpy as np, torch
from sspp_b200 import capi
from tests.helpers import joint_case, SEED
ns = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
case = joint_case(n_samples=ns, check_points=64, do_update=True)
case.cfg.tile_hint = int(os.environ.get("SSPP_TILE_HINT", "0"), 0)
h = capi.Handle(case.model, case.cfg); case.prepare(h); h.save_state()
h.phase_probe(True)
acc = []
for r in range(20):
    h.restore_state(); h.iterate(SEED, r, 1)
    st = h.phase_probe(True, read=True).astype(np.int64)
    if r >= 3: acc.append(st)
A = np.stack(acc)                      # [rep][cta][6]
t0 = A[:, :, 0].min(axis=1)[:, None, None]
R = (A - t0) * 1e-3                    # us since the first CTA started
names = ["start", "scored", "past_bar1", "ranked", "past_bar2", "updated", "rank_staged", "rank_counted"]
out = {n: dict(min=round(float(R[:, :, i].min(axis=1).mean()), 1), med=round(float(np.median(R[:, :, i], axis=1).mean()), 1),
               max=round(float(R[:, :, i].max(axis=1).mean()), 1)) for i, n in enumerate(names)}
print(json.dumps(dict(grid=int(A.shape[1]), phases_us=out)))
