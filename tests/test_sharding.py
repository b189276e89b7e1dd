"""Multi-rank host logic on the CPU: world_size-2 gloo process groups.

The collective code of sspp_b200/sharding.py is exercised for real (all_gather over gloo); the device
compute behind it is stood in for by the replica oracle O2 (tests only), so the check is: a population
sharded over 2 ranks reaches, bit for bit, the state of the unsharded population."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from tests.helpers import SEED, task_case, joint_case, o2_for
from sspp_b200 import capi
from sspp_b200.sharding import PopulationSharded, block_range, gather_rows


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


class ReplicaShard:
    """checker backend: O2 restating what a DeviceShard's handle does"""

    def __init__(self, case, offset, n_local, n_global):
        cfg = case.cfg
        cfg.n_samples, cfg.shard_offset, cfg.shard_global = n_local, offset, n_global
        self.o2 = o2_for(case)
        self.offset, self.n_local = offset, n_local
        self.psi_col, self.psi_arc = capi.planner_grids(cfg)
        self.base = capi.base_points(cfg, case.start[0], case.end[0])
        from tests.test_oracle_agreement import _host_state
        _, _, _, self.mean, self.sigma = _host_state(case)
        self.best = np.zeros_like(self.mean); self.bv = 0
        self.last = None

    def score(self, seed, it):
        r = self.o2.iterate(self.psi_col, self.psi_arc, self.base, self.mean, self.sigma, self.best, self.bv, seed, it,
                            cand_offset=self.offset, n_local=self.n_local, do_elites=False)
        return torch.from_numpy(r["keys"].view(np.int64).copy())

    def merge(self, gathered, n_shards, seed, it):
        m = self.o2.merge_update(gathered.numpy().view(np.uint64), seed, it, self.mean, self.sigma, self.best, self.bv)
        self.mean, self.sigma, self.best, self.bv, self.last = m["mean"], m["sigma"], m["best"], m["best_valid"], m


def _make_case(kind):
    if kind == "task":
        return task_case(n_samples=62, check_points=24)          # 64 candidates with the mean / best slots
    return joint_case(n_samples=64, check_points=16, do_update=True)


def _worker(rank, world, port, kind, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        case = _make_case(kind)
        lo, hi = block_range(64, rank, world)
        ps = PopulationSharded(ReplicaShard(case, lo, hi - lo, 64))
        states = []
        for it in range(4):
            ps.iterate(SEED, it)
            m = ps.b.last
            states.append((m["keys"].copy(), m["weights"].copy(), m["mean"].copy(), m["sigma"].copy(), m["best"].copy(), m["n_success"], m["best_idx"]))
        # query sharding helpers
        rows = np.arange(rank * 100, rank * 100 + (block_range(7, rank, world)[1] - block_range(7, rank, world)[0]) * 3, dtype=np.float32).reshape(-1, 3)
        g = gather_rows(rows, 7)
        out[rank] = (states, g)
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("kind", ["task", "joint"])
def test_population_sharded_equals_unsharded(kind):
    world = 2
    mgr = mp.Manager()
    out = mgr.dict()
    mp.spawn(_worker, args=(world, _free_port(), kind, out), nprocs=world, join=True)
    s0, g0 = out[0]
    s1, g1 = out[1]
    # the unsharded population on one "rank"
    case = _make_case(kind)
    o2 = o2_for(case)
    psi_col, psi_arc = capi.planner_grids(case.cfg)
    base = capi.base_points(case.cfg, case.start[0], case.end[0])
    from tests.test_oracle_agreement import _host_state
    _, _, _, mean, sigma = _host_state(case)
    best = np.zeros_like(mean); bv = 0
    assert o2.n_cand == 64
    for it in range(4):
        r = o2.iterate(psi_col, psi_arc, base, mean, sigma, best, bv, SEED, it)
        mean, sigma, best, bv = r["mean"], r["sigma"], r["best"], r["best_valid"]
        for st in (s0[it], s1[it]):            # both ranks agree with each other and with the unsharded run, bit for bit
            keys, w, m, s, b, ns, bi = st
            assert np.array_equal(keys, r["keys"])
            assert ns == r["n_success"] and (ns == 0 or bi == r["best_idx"])
            assert np.array_equal(w.view(np.uint32), r["weights"].view(np.uint32))
            assert np.array_equal(m.view(np.uint32), mean.view(np.uint32))
            assert np.array_equal(s.view(np.uint32), sigma.view(np.uint32))
            if ns:
                assert np.array_equal(b.view(np.uint32), best.view(np.uint32))
    assert np.array_equal(g0, g1) and g0.shape == (7, 3)
    assert np.array_equal(g0[:4].ravel(), np.arange(0, 12)) and np.array_equal(g0[4:].ravel(), np.arange(100, 109))


def test_block_range_covers_everything():
    for n in (0, 1, 7, 1024, 32640):
        for world in (1, 2, 3, 8):
            r = [block_range(n, k, world) for k in range(world)]
            assert r[0][0] == 0 and r[-1][1] == n and all(r[i][1] == r[i + 1][0] for i in range(world - 1))
            sizes = [b - a for a, b in r]
            assert max(sizes) - min(sizes) <= 1
