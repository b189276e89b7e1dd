"""Device time of one cfg2 step launched three ways: iterate_restored (snapshot in HBM), plan_step (inputs in the parameter
block), set_queries + set_joint_noise + iterate (staging ring).  CUDA events on the launch stream, mirror off."""
import sys, json
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import numpy as np, torch
from sspp_b200 import capi
from tests.helpers import joint_case, SEED
case = joint_case(n_samples=4096, check_points=64, do_update=True)
h = capi.Handle(case.model, case.cfg); st = torch.cuda.Stream(); h.set_stream(st.cuda_stream)
case.prepare(h); h.save_state(); h.set_result_mirror(False)
def dev_us(fn, n=200):
    for i in range(5): fn(i)
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record(st)
    for i in range(n): fn(10 + i)
    b.record(st); torch.cuda.synchronize()
    return round(a.elapsed_time(b) * 1e3 / n, 2)
out = dict(restored=dev_us(lambda i: h.iterate(SEED, i, 1, restored=True)),
           plan_step=dev_us(lambda i: h.plan_step(case.start, case.end, case.joint_sigma, case.limits, SEED, i, 1)),
           three_calls=dev_us(lambda i: (case.prepare(h), h.iterate(SEED, i, 1))))
print(json.dumps(out))
