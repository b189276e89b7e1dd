"""Per-CTA durations of the update phase (barrier 2 -> updated) of the cooperative kernel, cfg2; mirror off / on."""
import sys, json, os
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import numpy as np
from sspp_b200 import capi
from tests.helpers import joint_case, SEED
case = joint_case(n_samples=4096, check_points=64, do_update=True)
h = capi.Handle(case.model, case.cfg); case.prepare(h); h.save_state()
for mirror in (False, True):
    h.set_result_mirror(mirror)
    h.phase_probe(True)
    acc = []
    for r in range(20):
        h.iterate(SEED, r, 1, restored=True)
        st = h.phase_probe(True, read=True).astype(np.int64)
        if r >= 3: acc.append(st)
    A = np.stack(acc)
    upd = (A[:, :, 5] - A[:, :, 4]).mean(axis=0) * 1e-3
    rank = (A[:, :, 3] - A[:, :, 2]).mean(axis=0) * 1e-3
    t0 = A[:, :, 0].min(axis=1)[:, None]
    end = ((A[:, :, 5] - t0) * 1e-3).max(axis=1).mean()
    sub = {n: [round(float(x), 2) for x in ((A[:, :4, i] - A[:, :4, 4]).mean(axis=0) * 1e-3)] for i, n in ((8, "n_succ_read"), (9, "weights"), (10, "gathered"), (11, "summed"), (5, "done"))}
    print(json.dumps(dict(mirror=mirror, update_subphases_us_since_barrier2_cta0_3=sub, kernel_end_us=round(float(end), 1), update_us_cta0_31=[round(float(x), 1) for x in upd[:32]],
                          update_us_last4=[round(float(x), 1) for x in upd[-4:]], rank_us_min_med_max=[round(float(x), 1) for x in (rank.min(), np.median(rank), rank.max())],
                          rank_slowest_ctas=[int(i) for i in np.argsort(-rank)[:6]])))
