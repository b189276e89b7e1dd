#!/usr/bin/env python
"""Config 5 (BASELINE.json): pairwise path-cost matrix over 256 waypoints = 32 640 Planner-B plans (cold start + 59
warm iterations each, main_tsp_benchmark.cpp:73-93 parameters) on N GPUs of one box.

  python scripts/cost_matrix_bench.py                         # one GPU
  python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 scripts/cost_matrix_bench.py

Every rank plans its block of 1020-pair batches (one launch per batch: 1020 queries x 60 fused iterations), then ONE
all-gather of the per-pair costs.  Timed region: waypoints on the host -> whole matrix on the host of every rank,
barrier + synchronize on both sides, max over ranks; repeated REPS times after one warm-up build.  Rank 0 prints one
JSON line; `matrix_crc` must be the same number for every N (the matrix does not depend on the rank count)."""
import json
import os
import sys
import time
import zlib
from pathlib import Path

import numpy as np

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

from sspp_b200 import cost_matrix as cm  # noqa: E402
from tests.helpers import task_case  # noqa: E402

W, BATCH, ITERS, REPS = int(os.environ.get("CM_WAYPOINTS", 256)), int(os.environ.get("CM_BATCH", 1020)), 60, 3


def main():
    sys.stdout.flush()
    json_fd = os.dup(1)                                  # stdout = the JSON line only (NCCL prints its version there)
    os.dup2(2, 1)
    rank, world = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
    local = int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    case = task_case(n_queries=BATCH)
    case.cfg.device = local
    planner = cm.HandlePlanner(case.model, case.cfg, case.sigma0, ITERS)
    wp = cm.sample_waypoints(W)
    dev = torch.device("cuda", local)

    def build():
        return cm.cost_matrix(planner, wp, BATCH, rank, world, device=dev if world > 1 else None)

    def fence():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()

    m = build()                                          # warm-up (first launch, NCCL set-up)
    times = []
    for _ in range(REPS):
        fence()
        t0 = time.perf_counter()
        m = build()
        fence()
        t = torch.tensor([time.perf_counter() - t0], device=dev)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        times.append(float(t.item()))
    n_pairs = W * (W - 1) // 2
    s = float(np.median(times))
    d = planner.h.dims
    if rank == 0:
        off = m[np.triu_indices(W, 1)]
        os.write(json_fd, (json.dumps(dict(
            workload=f"cfg5: {W} waypoints, {n_pairs} Planner-B plans x {ITERS} iterations x {d.n_cand} candidates x "
                     f"{case.cfg.check_points} collocation points, robocrane.xml gripper, batches of {BATCH} pairs",
            n_gpus=world, seconds=s, plans_per_s=n_pairs / s,
            points_per_s=n_pairs * ITERS * d.n_cand * case.cfg.check_points / s,
            launches_per_rank=planner.launches // (REPS + 1), reps=REPS, times=times,
            finite_entries=int(np.isfinite(off).sum()), mean_finite_cost=float(off[np.isfinite(off)].mean()),
            symmetric=bool(np.array_equal(m, m.T)), matrix_crc=zlib.crc32(m.tobytes()))) + "\n").encode())
    planner.close()
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
