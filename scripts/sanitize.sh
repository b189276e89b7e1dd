#!/bin/bash
# compute-sanitizer over the hot kernel's hand-rolled synchronisation (mbarriers, grid barriers, shared-memory atomics, the
# dynamic tile counter, the cross-GPU flag protocol at world = 1).  Usage (GPU box, repo root): bash scripts/sanitize.sh <tag>
TAG=${1:-rXX}; OUT=gpurun_out; mkdir -p $OUT
cat > /tmp/sanitize_target.py <<'PY'
import sys
sys.path.insert(0, ".")
import numpy as np, torch
import __graft_entry__ as ge
from sspp_b200 import capi
from sspp_b200.sharding import PopulationExchange
from tests.helpers import SEED, joint_case, task_case, sdf_case
ge.smoke()                                                       # cooperative kernel, joint mode, 2 iterations vs the oracles
c = task_case(n_queries=160)                                     # one-CTA-per-query kernel (cfg4 shape), fused iterations
h = capi.Handle(c.model, c.cfg); c.prepare(h); h.iterate(SEED, 0, 6); h.results(); h.close()
c = joint_case(n_samples=2600, check_points=32, do_update=True)  # more tiles than CTAs: dynamic tile hand-out
c.cfg.tile_hint = (4 << 8) | 3
h = capi.Handle(c.model, c.cfg); c.prepare(h); h.set_success_rows(True)
h.plan_step(c.start, c.end, c.joint_sigma, c.limits, SEED, 0, 2); h.results(); h.success_rows(0); h.work_counters(); h.close()
c = sdf_case(n_samples=1024, check_points=32, n_points=5, dims=(32, 32, 24))
c.cfg.shard_offset, c.cfg.shard_global = 0, 1024
h = capi.Handle(c.model, c.cfg); c.prepare(h)
px = PopulationExchange(h)                                       # world = 1: publishes into its own exchange block
for it in range(3):
    px.iterate(SEED, it)
px.check(); h.close()
print("sanitize target done")
PY
for TOOL in memcheck racecheck synccheck; do
  timeout 1500 compute-sanitizer --tool $TOOL --print-limit 20 python /tmp/sanitize_target.py > $OUT/${TAG}_sanitize_${TOOL}.log 2>&1
  echo "$TOOL rc=$? : $(grep -E 'ERROR SUMMARY|RACECHECK SUMMARY' $OUT/${TAG}_sanitize_${TOOL}.log | tail -1)"; tail -3 $OUT/${TAG}_sanitize_${TOOL}.log | cut -c1-300
done
