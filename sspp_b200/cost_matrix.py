"""Pairwise path-cost matrix over waypoints (BASELINE.json config 5): every unordered pair i<j of W waypoints
is one independent Planner-B query - `plan(start, end, iterate=false)` followed by warm iterations on the same
distribution, the parameters of /root/reference/src/main_tsp_benchmark.cpp:73-93 - and the matrix entry is the best
`L + w*C_wf` the plan found (`tsp_planner.h:123,131-138`), +inf where no candidate was collision free.  The reference
plans such pairs one after another on the host (`scripts/main_tsp.py` builds its tour costs that way); here the pair
list is cut into fixed batches, a batch is ONE handle launch (all queries of the batch x all iterations fused), and
the batches are dealt to the ranks in contiguous blocks (`sharding.block_range`): no collective while planning, one
all-gather of the per-pair costs (4 bytes each) at the end.

A pair's random numbers are a function of (seed + batch index, iteration, slot in the batch, candidate) only, and the
batch size is part of the problem definition, not of the machine: the matrix is bit-identical for every world size.
"""
from __future__ import annotations

from typing import Callable

import numpy as np

from .sharding import block_range, gather_rows

SEED = 0x53535050

# plans one batch: (start [B,4], end [B,4], seed) -> best cost [B] float32, +inf where the plan found no free path
BatchPlanner = Callable[[np.ndarray, np.ndarray, int], np.ndarray]


def sample_waypoints(n: int, seed: int = 2024) -> np.ndarray:
    """waypoints drawn like config 4's end points (x, y, z, yaw inside the limits of main_icra_benchmark.cpp:160,
    away from the workspace rim)"""
    rng = np.random.default_rng(seed)
    return np.column_stack([rng.uniform(0.2, 0.6, n), rng.uniform(-0.4, 0.4, n), rng.uniform(0.14, 0.4, n),
                            rng.uniform(-1.5, 1.5, n)])


def pair_list(n: int) -> tuple[np.ndarray, np.ndarray]:
    """all unordered pairs i<j in row-major order: n*(n-1)/2 of them (256 waypoints -> 32 640)"""
    return np.triu_indices(n, 1)


def n_batches(n_pairs: int, batch: int) -> int:
    return (n_pairs + batch - 1) // batch


def plan_pairs(planner: BatchPlanner, waypoints: np.ndarray, batch: int, rank: int = 0, world: int = 1,
               seed: int = SEED) -> np.ndarray:
    """this rank's share of the pair costs: the costs of batches [lo, hi) of the pair list, flattened, the last batch
    of the list cut to its real length"""
    wp = np.asarray(waypoints, np.float64)
    if wp.ndim != 2 or wp.shape[1] != 4:
        raise ValueError("waypoints must be [W, 4] (x, y, z, yaw)")
    if batch < 1:
        raise ValueError("batch must be positive")
    ii, jj = pair_list(len(wp))
    n_pairs = len(ii)
    lo, hi = block_range(n_batches(n_pairs, batch), rank, world)
    out = []
    for b in range(lo, hi):
        p0, p1 = b * batch, min((b + 1) * batch, n_pairs)
        idx = np.arange(p0, p0 + batch)
        idx[idx >= n_pairs] = n_pairs - 1              # the last batch is padded with its last pair
        cost = np.asarray(planner(wp[ii[idx]], wp[jj[idx]], seed + b), np.float32)
        if cost.shape != (batch,):
            raise ValueError(f"batch planner returned shape {cost.shape}, expected ({batch},)")
        out.append(cost[: p1 - p0])
    return np.concatenate(out) if out else np.zeros(0, np.float32)


def assemble(n: int, pair_costs: np.ndarray) -> np.ndarray:
    """symmetric [n, n] float32 matrix with a zero diagonal from the costs in pair_list order"""
    ii, jj = pair_list(n)
    if len(pair_costs) != len(ii):
        raise ValueError(f"{len(pair_costs)} costs for {len(ii)} pairs")
    m = np.zeros((n, n), np.float32)
    m[ii, jj] = pair_costs
    m[jj, ii] = pair_costs
    return m


def cost_matrix(planner: BatchPlanner, waypoints: np.ndarray, batch: int, rank: int = 0, world: int = 1,
                seed: int = SEED, device=None, group=None) -> np.ndarray:
    """the whole matrix on every rank.  Ranks own whole batches, so the per-rank pieces are gathered batch-wise:
    every rank pads its piece to the same number of batches."""
    n = len(waypoints)
    n_pairs = n * (n - 1) // 2
    local = plan_pairs(planner, waypoints, batch, rank, world, seed)
    if world == 1:
        return assemble(n, local)
    nb = n_batches(n_pairs, batch)
    lo, hi = block_range(nb, rank, world)
    rows = np.full((hi - lo, batch), np.inf, np.float32)           # one row per batch (the list's last one is short)
    rows.reshape(-1)[: len(local)] = local
    full = gather_rows(rows, nb, device=device, group=group)
    return assemble(n, full.reshape(-1)[:n_pairs])


class HandlePlanner:
    """BatchPlanner on one GPU: a task-space handle with `batch` queries; a batch is set_queries + reset_distribution +
    ONE launch of `iters` fused iterations + one read-back of the per-query scalars"""

    def __init__(self, model, cfg, sigma0: float, iters: int = 60):
        from . import capi
        if cfg.mode != capi.MODE_TASK:
            raise ValueError("the cost matrix is a Planner-B (task space) workload")
        self.h = capi.Handle(model, cfg)
        self.batch = self.h.dims.n_queries
        self.sigma0, self.iters = float(sigma0), int(iters)
        self.launches = 0

    def __call__(self, start: np.ndarray, end: np.ndarray, seed: int) -> np.ndarray:
        h = self.h
        h.set_queries(start, end)
        h.reset_distribution(self.sigma0)                          # Planner::reset: cold start (tsp_planner.h:54-69)
        h.iterate(seed, 0, self.iters)
        self.launches += 1 if h.dims.ctas_per_query == 1 else self.iters
        r = h.results()
        # the best candidate seen over the plan: last_best_ is forwarded into every later iteration (include_best), so
        # the last iteration's best is the plan's best whenever any iteration succeeded
        return np.where(r["any_success"] > 0, r["best_cost"], np.float32(np.inf)).astype(np.float32)

    def close(self):
        self.h.close()
