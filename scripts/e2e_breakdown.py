#!/usr/bin/env python
"""Where an end-to-end step of bench.py (host buffers in, every candidate's result out) spends its time.
Each stage is timed on the host clock over REPS steps; 'alone' stages are followed by a stream synchronisation so that
the device part of the stage is included.  Usage (GPU box): python scripts/e2e_breakdown.py [cfg2|cfg4]"""
import json
import sys
import time
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import bench  # noqa: E402
from sspp_b200 import capi  # noqa: E402

import os
REPS = 300
name = sys.argv[1] if len(sys.argv) > 1 else "cfg2"
if len(sys.argv) > 2:
    os.environ["SSPPCU_RESULT_MIRROR"] = sys.argv[2]       # 0: results by D2H copies only (read when the handle is created)
case, _ = bench.workload_case(name, 0)
case.cfg.device = 0
h = capi.Handle(case.model, case.cfg)
case.prepare(h)
h.save_state()
inner = 60 if name == "cfg4" else 1
for i in range(5):
    h.iterate(bench.SEED, i, inner); h.sync()


def timed(fn):
    h.sync()
    t0 = time.perf_counter()
    for i in range(REPS):
        fn(i)
    h.sync()
    return (time.perf_counter() - t0) / REPS * 1e6


out = {}
out["prepare_enqueue_then_sync_each_us"] = timed(lambda i: (case.prepare(h), h.sync()))
out["prepare_enqueue_only_us"] = timed(lambda i: case.prepare(h))
out["iterate_sync_us"] = timed(lambda i: (h.iterate(bench.SEED, i, inner), h.sync()))
out["iterate_restored_sync_us"] = timed(lambda i: (h.iterate(bench.SEED, i, inner, restored=True), h.sync()))
out["results_alone_us"] = timed(lambda i: h.results())
out["best_after_results_us"] = timed(lambda i: (h.results(), h.best())) - out["results_alone_us"]
out["sync_alone_us"] = timed(lambda i: h.sync())


def full(i):
    case.prepare(h)
    h.iterate(bench.SEED, i, inner)
    h.results()
    h.best()


out["full_step_us"] = timed(full)
res, best = h.results(), h.best()


def full_reuse(i):
    case.prepare(h)
    h.iterate(bench.SEED, i, inner)
    h.results(res)
    h.best(best)


out["full_step_reused_buffers_us"] = timed(full_reuse)
out["iterate_results_us"] = timed(lambda i: (h.iterate(bench.SEED, i, inner), h.results(res)))
out["mirror"] = os.environ.get("SSPPCU_RESULT_MIRROR", "1")
if hasattr(h, "plan_step"):
    out["plan_step_us"] = timed(lambda i: h.plan_step(case.start, case.end, case.joint_sigma if case.cfg.mode == 0 else case.sigma0,
                                                      case.limits, bench.SEED, i, inner))
print(json.dumps(out, indent=1))
