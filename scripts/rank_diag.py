import sys, json, os
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import numpy as np, torch
from sspp_b200 import capi
from tests.helpers import joint_case, SEED
case = joint_case(n_samples=4096, check_points=64, do_update=True)
h = capi.Handle(case.model, case.cfg); case.prepare(h); h.save_state(); h.phase_probe(True)
for r in range(5): h.iterate(SEED, r, 1, restored=True)
st = h.phase_probe(True, read=True).astype(np.int64)
res = h.results()
ok = (res["status"][0] == 0)
cpc = h.dims.cand_per_cta
nsucc = np.array([ok[i * cpc:(i + 1) * cpc].sum() for i in range(st.shape[0])])
dur = (st[:, 7] - st[:, 6]) * 1e-3
stage = (st[:, 6] - st[:, 2]) * 1e-3
print("count dur percentiles", np.percentile(dur, [0, 10, 50, 90, 100]).round(2))
print("stage dur percentiles", np.percentile(stage, [0, 10, 50, 90, 100]).round(2))
for k in range(0, cpc + 1):
    m = nsucc == k
    if m.any(): print(k, int(m.sum()), dur[m].mean().round(2), dur[m].max().round(2))
order = np.argsort(-dur)[:10]
print("slowest CTAs", order.tolist(), dur[order].round(2).tolist(), nsucc[order].tolist())
