"""Config 5 (pairwise path-cost matrix): the host logic on the CPU with a stand-in batch planner (pair
list, fixed batches dealt to ranks, gloo world-2 gather, symmetric assembly), and on the GPU the real handle planner:
the matrix must not depend on how many ranks planned it."""
import os
import socket

import numpy as np
import pytest
import torch.distributed as dist
import torch.multiprocessing as mp

from sspp_b200 import cost_matrix as cm


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def stub_planner(start, end, seed):
    """stand-in: joint-space distance, +inf for 'blocked' pairs, and a term that exposes the batch seed"""
    d = np.linalg.norm(np.asarray(end) - np.asarray(start), axis=1).astype(np.float32)
    d = d + np.float32((seed % 7) * 1e-3)
    d[np.asarray(start)[:, 0] + np.asarray(end)[:, 0] > 0.9] = np.inf
    return d


def test_pair_list_and_assembly():
    ii, jj = cm.pair_list(256)
    assert len(ii) == 32640 and (ii < jj).all()
    c = np.arange(6, dtype=np.float32) + 1
    m = cm.assemble(4, c)
    assert np.array_equal(m, m.T) and (np.diag(m) == 0).all()
    assert m[0, 1] == 1 and m[0, 3] == 3 and m[1, 2] == 4 and m[2, 3] == 6
    with pytest.raises(ValueError):
        cm.assemble(4, c[:5])


@pytest.mark.parametrize("n,batch", [(9, 8), (12, 66), (12, 5), (2, 4)])
def test_ranks_partition_the_batches_and_agree_with_one_rank(n, batch):
    wp = cm.sample_waypoints(n, seed=3)
    one = cm.plan_pairs(stub_planner, wp, batch)
    assert len(one) == n * (n - 1) // 2
    for world in (2, 3, 8):
        parts = [cm.plan_pairs(stub_planner, wp, batch, r, world) for r in range(world)]
        assert np.array_equal(np.concatenate(parts), one)
    # every batch sees its own seed, the last one is padded to full length
    seen = []
    cm.plan_pairs(lambda s, e, seed: (seen.append((len(s), seed)), np.zeros(len(s), np.float32))[1], wp, batch)
    assert all(b == batch for b, _ in seen) and [s for _, s in seen] == [cm.SEED + i for i in range(len(seen))]


def test_bad_inputs_are_rejected():
    with pytest.raises(ValueError):
        cm.plan_pairs(stub_planner, np.zeros((4, 3)), 4)
    with pytest.raises(ValueError):
        cm.plan_pairs(stub_planner, cm.sample_waypoints(4), 0)
    with pytest.raises(ValueError):
        cm.plan_pairs(lambda s, e, seed: np.zeros(3, np.float32), cm.sample_waypoints(4), 4)


def _worker(rank, world, port, n, batch, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        out[rank] = cm.cost_matrix(stub_planner, cm.sample_waypoints(n, seed=3), batch, rank, world)
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("n,batch", [(14, 8), (6, 32)])
def test_two_gloo_ranks_build_the_matrix_of_one(n, batch):
    world = 2
    out = mp.Manager().dict()
    mp.spawn(_worker, args=(world, _free_port(), n, batch, out), nprocs=world, join=True)
    one = cm.cost_matrix(stub_planner, cm.sample_waypoints(n, seed=3), batch)
    assert np.array_equal(out[0], one) and np.array_equal(out[1], one)
    assert np.array_equal(one, one.T) and np.isinf(one).any() and np.isfinite(one).any()


@pytest.mark.gpu
def test_handle_planner_matrix_is_independent_of_the_rank_count():
    from tests.helpers import task_case
    n, batch, iters = 10, 16, 6
    wp = cm.sample_waypoints(n, seed=11)
    case = task_case(n_queries=batch)

    def matrix(world):
        parts = []
        for r in range(world):                       # every "rank" is a fresh handle, as a process would hold
            p = cm.HandlePlanner(case.model, case.cfg, case.sigma0, iters)
            parts.append(cm.plan_pairs(p, wp, batch, r, world))
            p.close()
        return cm.assemble(n, np.concatenate(parts))

    m1, m3 = matrix(1), matrix(3)
    assert np.array_equal(m1.view(np.uint32), m3.view(np.uint32))
    assert np.array_equal(m1, m1.T) and (np.diag(m1) == 0).all()
    off = m1[np.triu_indices(n, 1)]
    assert np.isfinite(off).sum() > len(off) // 2
    # a finite entry is a path length plus penalties: never shorter than the straight line between the waypoints
    ii, jj = np.triu_indices(n, 1)
    straight = np.linalg.norm(wp[ii] - wp[jj], axis=1)
    ok = np.isfinite(off)
    assert (off[ok] >= straight[ok] * (1 - 1e-5)).all()
    # one batch planned directly through the handle gives the same numbers as the matrix holds
    p = cm.HandlePlanner(case.model, case.cfg, case.sigma0, iters)
    direct = p(wp[ii[:batch]], wp[jj[:batch]], cm.SEED)
    p.close()
    assert np.array_equal(direct.view(np.uint32), off[:batch].view(np.uint32))
