"""Experiment probe (one library per run, SSPP_CUDA_LIB): cfg2 device time per restored iteration, phase timeline, and a CRC
of everything an iteration leaves behind (sorted keys, elite weights, distribution, per-candidate results) after a few
iterations - equal CRCs across libraries = the change kept every bit."""
import sys, json, zlib
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import numpy as np, torch
from sspp_b200 import capi
from tests.helpers import joint_case, task_case, SEED
case = joint_case(n_samples=4096, check_points=64, do_update=True)
h = capi.Handle(case.model, case.cfg); st = torch.cuda.Stream(); h.set_stream(st.cuda_stream)
case.prepare(h); h.save_state(); h.set_result_mirror(False)
crc = 0
for it in range(4):
    h.iterate(SEED, it, 1)
    r = h.results()
    for k in sorted(r):
        crc = zlib.crc32(np.ascontiguousarray(r[k]).tobytes(), crc)
    m, s = h.get_distribution()[:2]
    crc = zlib.crc32(np.ascontiguousarray(m).tobytes(), crc); crc = zlib.crc32(np.ascontiguousarray(s).tobytes(), crc)
    e = h.elites()
    for x in (e if isinstance(e, (tuple, list)) else [e]):
        crc = zlib.crc32(np.ascontiguousarray(x).tobytes(), crc)
def dev_us(fn, n=300):
    for i in range(10): fn(i)
    torch.cuda.synchronize()
    best = 1e9
    for rep in range(5):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(st)
        for i in range(n): fn(10 + i)
        b.record(st); torch.cuda.synchronize()
        best = min(best, a.elapsed_time(b) * 1e3 / n)
    return round(best, 2)
us = dev_us(lambda i: h.iterate(SEED, i, 1, restored=True))
h.phase_probe(True)
acc = []
for r in range(20):
    h.iterate(SEED, r, 1, restored=True)
    stp = h.phase_probe(True, read=True).astype(np.int64)
    if r >= 3: acc.append(stp)
A = np.stack(acc); t0 = A[:, :, 0].min(axis=1)[:, None, None]; R = (A - t0) * 1e-3
names = ["start", "scored", "past_bar1", "ranked", "past_bar2", "updated", "rank_staged", "rank_counted"]
ph = {n: [round(float(R[:, :, i].min(axis=1).mean()), 1), round(float(np.median(R[:, :, i], axis=1).mean()), 1), round(float(R[:, :, i].max(axis=1).mean()), 1)] for i, n in enumerate(names)}
print(json.dumps(dict(us_per_iter=us, crc=crc, phases_min_med_max=ph)))
