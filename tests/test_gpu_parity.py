"""Parity of the sm_100a path (through the C ABI) against the checkers, on a B200.

Bar (BASELINE.json north_star): statuses (collision flags), sorted elite indices, weights and the
updated distribution bit-exact against the replica O2 on the same Philox counters; against the
independent FP64 oracle O1, flags exact outside EPS_CONTACT of contact and path values within
REL = 1e-5 relative."""
import numpy as np
import pytest

from tests.helpers import SEED, armlab_case, compare_iteration, joint_case, task_case, o1_for, o2_for

pytestmark = pytest.mark.gpu
REL = 1e-5


def _handle(case):
    from sspp_b200 import capi
    h = capi.Handle(case.model, case.cfg)
    case.prepare(h)
    return h


def _assert_exact(rep):
    for k in ("status_mismatch", "elite_mismatch", "key_mismatch", "value_bits_mismatch", "param_bits_mismatch",
              "dist_bits_mismatch", "o1_flag_mismatch"):
        assert rep[k] == 0, rep
    assert rep["o1_max_rel"] <= REL, rep


@pytest.mark.parametrize("n_samples,check_points,update", [(64, 16, True), (50, 50, False), (100, 100, False),
                                                           (1000, 64, True), (4096, 64, True)])
def test_joint_iterations_match_oracles(n_samples, check_points, update):
    case = joint_case(n_samples=n_samples, check_points=check_points, do_update=update)
    h = _handle(case)
    rep = compare_iteration(h, case, iters=3 if n_samples <= 1000 else 2)   # O1 leg: every candidate, or 512 seeded ones of the 4096
    _assert_exact(rep)
    assert max(rep["n_success"]) > 0
    h.close()


@pytest.mark.parametrize("n_points,n_samples,check_points", [(3, 15, 40), (5, 15, 40), (3, 254, 64), (4, 1022, 32)])
def test_task_iterations_match_oracles(n_points, n_samples, check_points):
    case = task_case(n_points=n_points, n_samples=n_samples, check_points=check_points)
    h = _handle(case)
    rep = compare_iteration(h, case, iters=6)
    _assert_exact(rep)
    h.close()


def test_task_stacking_scene():
    case = task_case("stacking.xml", "block1", n_samples=126, start=np.array([0.205, 0.0, 0.12, 0.0]),
                     end=np.array([-0.3, 0.3, 0.12, 1.0]))
    h = _handle(case)
    _assert_exact(compare_iteration(h, case, iters=4))
    h.close()


@pytest.mark.parametrize("n_samples,check_points,update", [(64, 16, True), (1000, 64, True)])
def test_armlab_scene_iterations_match_oracles(n_samples, check_points, update):
    """fixtures/armlab.xml: frames, a two-joint body (a chain of one-joint links in the kernel tables), a
    slide joint, capsule / sphere / cylinder / box partners with per-class margins"""
    case = armlab_case(n_samples=n_samples, check_points=check_points, do_update=update, w_collision=0.5)
    h = _handle(case)
    rep = compare_iteration(h, case, iters=3, eps_contact=2e-4)
    _assert_exact(rep)
    assert 0 < max(rep["n_success"]) < n_samples          # the start-goal pair has both outcomes
    h.close()


def test_eval_points_matches_replica_bitwise():
    case = joint_case()
    h = _handle(case)
    o2 = o2_for(case)
    rng = np.random.default_rng(2)
    Q = rng.uniform(-2, 2, (4000, 7)).astype(np.float32)
    g = h.eval_points(Q)
    r = np.array([o2.point_cost(q) for q in Q], np.float32)
    assert np.array_equal(g.view(np.uint32), r.view(np.uint32))
    assert 0.05 < (g > 0).mean() < 0.95
    h.close()
    case = task_case()
    h = _handle(case)
    o2 = o2_for(case)
    Q = np.column_stack([rng.uniform(0, 0.7, 4000), rng.uniform(-0.7, 0.7, 4000), rng.uniform(0.0, 0.6, 4000),
                         rng.uniform(-1.6, 1.6, 4000)]).astype(np.float32)
    g = h.eval_points(Q)
    r = np.array([o2.point_cost(q) for q in Q], np.float32)
    assert np.array_equal(g.view(np.uint32), r.view(np.uint32))
    h.close()


def test_many_queries_one_launch_equals_single_queries():
    """query batching (cfg 4): Q queries in one handle == each query alone, bit for bit; inner
    iterations inside one launch == one launch per iteration."""
    Q = 6
    rng = np.random.default_rng(9)
    starts = np.column_stack([rng.uniform(0.2, 0.6, Q), rng.uniform(-0.4, 0.4, Q), rng.uniform(0.14, 0.4, Q), rng.uniform(-1.5, 1.5, Q)])
    ends = np.column_stack([rng.uniform(0.2, 0.6, Q), rng.uniform(-0.4, 0.4, Q), rng.uniform(0.14, 0.4, Q), rng.uniform(-1.5, 1.5, Q)])
    case = task_case(n_queries=Q, start=starts, end=ends)
    h = _handle(case)
    h.iterate(SEED, 0, 5)
    rb = h.results(); mb, sb = h.get_distribution()
    h.close()
    h = _handle(case)
    for it in range(5):
        h.iterate(SEED, it, 1)
    r1 = h.results(); m1, s1 = h.get_distribution()
    h.close()
    for k in ("L", "Cnf", "Cwf", "cost", "status", "n_success", "best_idx"):
        assert np.array_equal(rb[k], r1[k]), k
    assert np.array_equal(mb, m1) and np.array_equal(sb, s1)
    # each query alone (query index is a Philox counter word, so pass it through a Q-query handle is
    # the definition; the replica with query=q is the check)
    h = _handle(case)
    for q in (0, Q - 1):
        hq = _handle(case)
        rep = compare_iteration(hq, case, iters=3, query=q)
        _assert_exact(rep)
        hq.close()
    h.close()


def test_no_device_memory_growth_and_errors():
    from sspp_b200 import capi
    case = joint_case()
    bad = joint_case()
    bad.cfg.n_points = 4        # fewer than 2*degree+... leaves no free rows is allowed; below degree+1 is not
    bad.cfg.n_points = 3
    with pytest.raises(capi.SsppError):
        capi.Handle(bad.model, bad.cfg)
    bad = joint_case(); bad.cfg.dof = 9
    with pytest.raises(capi.SsppError):
        capi.Handle(bad.model, bad.cfg)
    bad = task_case(); bad.cfg.body_name = b"nope"
    with pytest.raises(capi.SsppError, match="not found"):
        capi.Handle(bad.model, bad.cfg)
    for _ in range(20):
        h = _handle(case); h.iterate(SEED, 0, 1); h.sync(); h.close()


@pytest.mark.parametrize("kind,n_global", [("task", 1024), ("joint", 4096)])
def test_population_shards_match_unsharded(kind, n_global):
    """population sharding (cfg 3 shape) emulated on one GPU: two shard handles + the merge kernels reach the
    state of the unsharded handle bit for bit (multi-rank transport is covered by tests/test_sharding.py)."""
    import torch
    from sspp_b200 import capi
    from sspp_b200.sharding import DeviceShard

    def mk(n_samples, offset=0, glob=0):
        case = task_case(n_samples=n_samples, check_points=32) if kind == "task" else \
            joint_case(n_samples=n_samples, check_points=32, do_update=True)
        case.cfg.shard_offset, case.cfg.shard_global = offset, glob
        h = capi.Handle(case.model, case.cfg)
        case.prepare(h)
        return h
    slots = 2 if kind == "task" else 0
    full = mk(n_global - slots)
    half = n_global // 2
    shards = [DeviceShard(mk(half, 0, n_global)), DeviceShard(mk(half, half, n_global))]
    for it in range(4):
        full.iterate(SEED, it, 1)
        r = full.results()
        fi, fc, fw = full.elites(0)
        for s in shards:
            s.score(SEED, it)
        torch.cuda.synchronize()                     # the shards run on their own streams
        keys = torch.cat([s.keys for s in shards])
        torch.cuda.synchronize()
        for s in shards:
            s.merge(keys, 2, SEED, it)
            mi, mc, mw = s.h.merged(n_global)
            assert np.array_equal(mi, fi) and np.array_equal(mc.view(np.uint32), fc.view(np.uint32))
            k = int(r["n_elite"][0])
            assert np.array_equal(mw[:k].view(np.uint32), fw[:k].view(np.uint32))
            rs = s.h.results()
            assert rs["n_success"][0] == r["n_success"][0] and rs["best_idx"][0] == r["best_idx"][0]
            m0, s0 = full.get_distribution(); m1, s1 = s.h.get_distribution()
            assert np.array_equal(m0, m1) and np.array_equal(s0, s1)
            b0, v0 = full.best(); b1, v1 = s.h.best()
            assert np.array_equal(b0, b1) and np.array_equal(v0, v1)
    for s in shards:
        s.h.close()
    full.close()


@pytest.mark.parametrize("kind,n_global", [("task", 2048), ("joint", 4096)])
def test_fused_exchange_single_rank_matches_unsharded(kind, n_global):
    """the exchange fused into the iteration kernel (ssppcu_iterate_exchange) with world = 1: publish into the
    rank's own block, flag, merge and update inside one launch reach the unsharded state bit for bit.  world > 1
    needs one GPU per rank: scripts/check_popshard.py under torchrun (profiles/README.md)."""
    import torch
    from sspp_b200 import capi
    from sspp_b200.sharding import PopulationExchange

    def mk(n_samples, offset=0, glob=0):
        case = task_case(n_samples=n_samples, check_points=32) if kind == "task" else \
            joint_case(n_samples=n_samples, check_points=32, do_update=True)
        case.cfg.shard_offset, case.cfg.shard_global = offset, glob
        h = capi.Handle(case.model, case.cfg)
        case.prepare(h)
        return h
    slots = 2 if kind == "task" else 0
    full = mk(n_global - slots)
    px = PopulationExchange(mk(n_global, 0, n_global))
    with pytest.raises(capi.SsppError):
        px.h.exchange_export(1, 0)                    # one block per handle
    for it in range(4):
        full.iterate(SEED, it, 1)
        px.iterate(SEED, it)
        px.check()
        r, rs = full.results(), px.h.results()
        fi, fc, fw = full.elites(0)
        mi, mc, mw = px.h.merged(n_global)
        assert np.array_equal(mi, fi) and np.array_equal(mc.view(np.uint32), fc.view(np.uint32))
        k = int(r["n_elite"][0])
        assert k > 0 and np.array_equal(mw[:k].view(np.uint32), fw[:k].view(np.uint32))
        assert rs["n_success"][0] == r["n_success"][0] and rs["best_idx"][0] == r["best_idx"][0]
        m0, s0 = full.get_distribution(); m1, s1 = px.h.get_distribution()
        assert np.array_equal(m0, m1) and np.array_equal(s0, s1)
        b0, v0 = full.best(); b1, v1 = px.h.best()
        assert np.array_equal(b0, b1) and np.array_equal(v0, v1)
    px.h.close(); full.close()


def test_sdf_mode_matches_oracles():
    """cfg 3 back end: grid baked on the device == the replica's grid bit for bit; scored iterations bit-exact
    against the replica and within tolerance of the FP64 statement."""
    from tests.helpers import sdf_case
    case = sdf_case(n_samples=510, check_points=64, n_points=5, dims=(64, 64, 48))
    h = _handle(case)
    o2 = o2_for(case)
    nodes, lo, cell = h.sdf()
    assert np.array_equal(nodes.view(np.uint32), o2.sdf_nodes().view(np.uint32))
    assert np.array_equal(h.spheres(), o2.spheres())
    rng = np.random.default_rng(4)
    Q = np.column_stack([rng.uniform(-0.9, 0.9, 3000), rng.uniform(-0.9, 0.9, 3000), rng.uniform(-0.1, 1.0, 3000), rng.uniform(-1.6, 1.6, 3000)]).astype(np.float32)
    g = h.eval_points(Q)                                   # includes points outside the grid (clamped reads)
    r = np.array([o2.point_cost(q) for q in Q], np.float32)
    assert np.array_equal(g.view(np.uint32), r.view(np.uint32))
    rep = compare_iteration(h, case, iters=6, eps_contact=1e-4)
    for k in ("status_mismatch", "elite_mismatch", "key_mismatch", "value_bits_mismatch", "param_bits_mismatch", "dist_bits_mismatch",
              "o1_flag_mismatch"):
        assert rep[k] == 0, rep
    assert rep["o1_max_rel"] <= REL, rep
    assert max(rep["n_success"]) > 0
    h.close()
    bad = task_case()                                      # robocrane gripper: geoms with different collision masks
    bad.cfg.collision_mode = 1
    for a in range(3):
        bad.cfg.sdf_dims[a], bad.cfg.sdf_lo[a], bad.cfg.sdf_hi[a] = 16, -1.0, 1.0
    from sspp_b200 import capi
    with pytest.raises(capi.SsppError, match="uniform collision mask"):
        capi.Handle(bad.model, bad.cfg)


@pytest.mark.parametrize("kind", ["joint", "task"])
def test_iterate_restored_equals_restore_then_iterate(kind):
    """ssppcu_iterate_restored = ssppcu_restore_state + ssppcu_iterate, bit for bit, in both kernel shapes"""
    from sspp_b200 import capi
    if kind == "joint":
        case = joint_case(n_samples=1000, check_points=32, do_update=True)
    else:
        rng = np.random.default_rng(3)
        Q = 160
        s = np.column_stack([rng.uniform(0.2, 0.6, Q), rng.uniform(-0.4, 0.4, Q), rng.uniform(0.14, 0.4, Q), rng.uniform(-1.5, 1.5, Q)])
        e = np.column_stack([rng.uniform(0.2, 0.6, Q), rng.uniform(-0.4, 0.4, Q), rng.uniform(0.14, 0.4, Q), rng.uniform(-1.5, 1.5, Q)])
        case = task_case(n_queries=Q, start=s, end=e)
    out = []
    for fused in (False, True):
        h = _handle(case)
        h.iterate(SEED, 0, 2)
        h.save_state()
        h.iterate(SEED, 2, 3)                       # move the live state away from the snapshot
        if fused:
            h.iterate(SEED, 7, 2, restored=True)
        else:
            h.restore_state(); h.iterate(SEED, 7, 2)
        r = h.results(); m, sg = h.get_distribution(); b, bv = h.best()
        out.append((r, m, sg, b, bv))
        h.close()
    (r0, m0, s0, b0, v0), (r1, m1, s1, b1, v1) = out
    for k in r0:
        assert np.array_equal(r0[k], r1[k]), k
    assert np.array_equal(m0, m1) and np.array_equal(s0, s1) and np.array_equal(b0, b1) and np.array_equal(v0, v1)


def test_best_read_after_results_needs_no_second_fetch_and_stays_current():
    """get_results brings best | best_valid along; get_best right after it reads the pinned mirror, any call in
    between (an iteration, a reset) makes it fetch again"""
    case = task_case(n_samples=126, check_points=40)
    h = _handle(case)
    h.iterate(SEED, 0, 1)
    b_direct, v_direct = h.best()                     # fetched
    h.results()
    b_mirror, v_mirror = h.best()                     # from the mirror get_results filled
    assert np.array_equal(b_direct, b_mirror) and np.array_equal(v_direct, v_mirror) and v_direct[0] == 1
    h.results()
    h.iterate(SEED, 1, 3)                             # state moves on after the results call
    b2, _ = h.best()
    h.results()
    b3, _ = h.best()
    assert np.array_equal(b2, b3)
    h.results()
    case.prepare(h)                                   # reset_distribution clears the forwarded best
    _, v4 = h.best()
    assert v4[0] == 0


@pytest.mark.gpu
@pytest.mark.parametrize("n_iters", [1, 3])
def test_results_mirrored_by_the_kernel_equal_the_copied_ones(n_iters):
    """the cooperative kernel copies the result block into the mapped pinned mirror itself (mirror_results): a read-back
    right after the launch does no D2H copy.  Any call in between (here: sync) makes get_results copy the device block
    instead - both must give the same bytes, for every array, the scalars and best, and re-used output buffers too"""
    case = joint_case(n_samples=2046, check_points=32, n_points=10, do_update=True)
    h = _handle(case)
    assert h.dims.ctas_per_query > 1                  # the cooperative shape
    buf = None
    for it in range(3):
        h.iterate(SEED, it * n_iters, n_iters)
        buf = h.results(buf); b_m, v_m = h.best()     # mirrored by the kernel
        r_m = {k: buf[k].copy() for k in h._RES_KEYS}
        h.sync()
        r_c = h.results(); b_c, v_c = h.best()        # copied from the device block
        for k in h._RES_KEYS:
            assert np.array_equal(r_m[k], r_c[k]), (it, k)
        assert np.array_equal(b_m, b_c) and np.array_equal(v_m, v_c)
        assert r_m["n_success"][0] > 0 and r_m["n_elite"][0] > 0 and v_m[0] == 1
    h.close()


def test_tile_shapes_and_kernel_variants_agree_bitwise():
    """the launch-bound variant, threads per CTA and candidates per CTA are pure tuning knobs"""
    ref = None
    for hint in (0, (0 + 1) | (14 << 8) | (256 << 20), (2 + 1) | (9 << 8) | (320 << 20), (1 + 1) | (31 << 8) | (256 << 20),
                 (5 + 1) | (20 << 8) | (384 << 20)):
        case = joint_case(n_samples=2000, check_points=24, do_update=True)
        case.cfg.tile_hint = hint
        h = _handle(case)
        h.iterate(SEED, 0, 2)
        r = h.results(); m, sg = h.get_distribution(); idx, cost, w = h.elites(0)
        cur = (r["cost"].copy(), r["status"].copy(), m, sg, idx, w)
        if ref is None:
            ref = cur
        else:
            for a, b in zip(ref, cur):
                assert np.array_equal(a, b), hex(hint)
        h.close()


def test_phase_probe_stamps_are_ordered():
    case = joint_case(n_samples=2048, check_points=32, do_update=True)
    h = _handle(case)
    h.phase_probe(True)
    h.iterate(SEED, 0, 1)
    st = h.phase_probe(True, read=True).astype(np.int64)
    assert st.shape[1] == 12 and (st[:, :6] > 0).all()
    assert (np.diff(st[:, :6], axis=1) >= 0).all()              # start <= scored <= barrier 1 <= ranked <= barrier 2 <= updated
    assert st[:, 2].min() >= st[:, 1].max() - 2000              # nobody passes barrier 1 before everybody has scored (timer granularity)
    h.close()


def test_set_distribution_round_trips_mean_sigma_and_either_alone():
    """ssppcu_set_distribution: both blocks in one staged copy, or either alone without touching the other"""
    case = task_case(n_samples=30, check_points=20, n_queries=3)
    h = _handle(case)
    d = h.dims
    rng = np.random.default_rng(5)
    m0, s0 = h.get_distribution()
    m1 = rng.uniform(-1, 1, m0.shape).astype(np.float32).astype(np.float64)
    s1 = rng.uniform(0.01, 0.5, s0.shape).astype(np.float32).astype(np.float64)
    h.set_distribution(mean=m1)
    m, s = h.get_distribution()
    assert np.array_equal(m, m1) and np.array_equal(s, s0)
    h.set_distribution(sigma=s1)
    m, s = h.get_distribution()
    assert np.array_equal(m, m1) and np.array_equal(s, s1)
    h.set_distribution(mean=m0, sigma=s0)
    m, s = h.get_distribution()
    assert np.array_equal(m, m0) and np.array_equal(s, s0)
    case.prepare(h)                                   # reset: best_valid | any_success cleared by one memset
    h.iterate(SEED, 0, 1)
    r = h.results()
    assert r["any_success"].shape == (d.n_queries,)


@pytest.mark.parametrize("kind", ["joint", "joint_multi", "task", "task_batch"])
def test_plan_step_equals_the_three_calls(kind):
    """ssppcu_plan_step (query + initial state in the kernel's parameter block, one launch) == ssppcu_set_queries +
    ssppcu_set_joint_noise / ssppcu_reset_distribution + ssppcu_iterate, bit for bit; success rows and candidate_ctrl read back
    what the kernel sampled"""
    from sspp_b200 import capi
    if kind == "joint":
        case = joint_case(n_samples=1000, check_points=32, do_update=True)
    elif kind == "joint_multi":
        case = joint_case(n_samples=300, check_points=24, do_update=True, n_queries=2)       # two queries still fit the parameter block
    elif kind == "task":
        case = task_case(n_samples=62, n_points=5)
    else:
        rng = np.random.default_rng(5)
        Q = 300                                                                               # too many for the block: the three-call path inside
        s = np.column_stack([rng.uniform(0.2, 0.6, Q), rng.uniform(-0.4, 0.4, Q), rng.uniform(0.14, 0.4, Q), rng.uniform(-1.5, 1.5, Q)])
        e = np.column_stack([rng.uniform(0.2, 0.6, Q), rng.uniform(-0.4, 0.4, Q), rng.uniform(0.14, 0.4, Q), rng.uniform(-1.5, 1.5, Q)])
        case = task_case(n_queries=Q, start=s, end=e)
    a = capi.Handle(case.model, case.cfg); b = capi.Handle(case.model, case.cfg)
    b.set_success_rows(True)
    joint = case.cfg.mode == 0
    for rep in range(2):                                       # the second round starts from a used handle (state must be re-armed)
        case.prepare(a)
        a.iterate(SEED, 10 * rep, 3)
        b.plan_step(case.start, case.end, case.joint_sigma if joint else case.sigma0, case.limits if joint else None, SEED, 10 * rep, 3)
        ra, rb = a.results(), b.results()
        for k in ("L", "Cnf", "Cwf", "cost"):
            assert np.array_equal(ra[k].view(np.uint32), rb[k].view(np.uint32)), k
        for k in ("status", "n_success", "n_elite", "best_idx"):
            assert np.array_equal(ra[k], rb[k]), k
        assert np.array_equal(a.params(), b.params()) and np.array_equal(a.base(), b.base())
        for x, y in zip(a.get_distribution() + a.best(), b.get_distribution() + b.best()):
            assert np.array_equal(x, y)
        P = b.params()
        for q in range(min(case.cfg.n_queries, 3)):
            rows, idx = b.success_rows(q)
            ok = np.flatnonzero(rb["status"][q] == 0)
            assert np.array_equal(idx, ok) and np.array_equal(rows, P[q].reshape(len(P[q]), -1)[ok])
            if len(ok):
                c = b.candidate_ctrl(int(ok[-1]), q)
                fb = 1 if not joint else case.cfg.degree
                assert np.array_equal(c[fb:fb + P.shape[2]], P[q, ok[-1]]) and np.array_equal(np.delete(c, np.s_[fb:fb + P.shape[2]], 0),
                                                                                             np.delete(b.base()[q], np.s_[fb:fb + P.shape[2]], 0))
    with pytest.raises(capi.SsppError):
        a.success_rows(0)                                      # never switched on for this handle
    a.close(); b.close()


def test_work_counters_describe_the_scored_iteration():
    from sspp_b200 import capi, roofline
    case = joint_case(n_samples=512, check_points=64, do_update=True)
    h = _handle(case)
    h.iterate(SEED, 0, 1)
    wc = h.work_counters()
    r = h.results()
    d = h.dims
    assert wc["points"] == 512 * 65 and wc["fk_hinge_axis"] == 7 * wc["points"]
    assert 0 < wc["broad"] <= d.n_pair * wc["points"] and wc["body_broad"] > 0      # the tool body's geoms are skipped while it is clear of everything
    assert wc["np_caps_box"] <= wc["np_caps_box_early"] + wc["np_caps_box"] <= wc["broad"]
    assert wc["contacts"] == int(round(float(r["Cnf"].sum())))              # planner A: every contact adds exactly 1 (sspp.h:144)
    # the same counters from the replica oracle's host build of the same header
    o2 = o2_for(case)
    import ctypes as C
    o2.lib.o2_work.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_int]
    psi_col, psi_arc = h.collocation()
    case.prepare(h)
    mean, sigma = h.get_distribution()
    o2.lib.o2_work(o2.h, 1, None, 0)
    o2.iterate(psi_col, psi_arc, h.base()[0], mean[0], sigma[0], np.zeros_like(mean[0]), 0, SEED, 0)
    out = np.zeros(len(capi.WORK_COUNTERS), np.uint64)
    o2.lib.o2_work(o2.h, 0, out.ctypes.data_as(C.c_void_p), len(out))
    assert {k: int(v) for k, v in zip(capi.WORK_COUNTERS, out)} == wc
    ex = roofline.executed_flop(wc, n_cand_total=512, check_points=64, dof=7, n_points=10, degree=3, mode=0)
    nom = roofline.work_model(case.model, 0, 7, 10)
    assert 500 < ex["flop_per_point_executed"] < nom["flop_per_point"]
    h.close()


def test_bound_plan_step_reads_the_callers_buffers_every_call():
    """Handle.bind_plan_step (addresses of the caller's host arrays taken once, like a C caller's pointers) == plan_step, and an
    in-place edit of a bound array changes the next call's result the way passing the edited array would"""
    from sspp_b200 import capi
    case = joint_case(n_samples=500, check_points=32, do_update=True)
    a = capi.Handle(case.model, case.cfg); b = capi.Handle(case.model, case.cfg)
    start = np.ascontiguousarray(case.start, np.float64).copy(); end = np.ascontiguousarray(case.end, np.float64).copy()
    lim = np.ascontiguousarray(case.limits, np.float64).copy()
    step = b.bind_plan_step(start, end, lim)
    best = None
    for rep in range(3):
        if rep == 2:
            start[...] += 0.01; end[...] -= 0.02                       # in place: the bound addresses stay valid
        a.plan_step(start.copy(), end.copy(), case.joint_sigma, lim.copy(), SEED, 5 * rep, 2)
        step(case.joint_sigma, SEED, 5 * rep, 2)
        ra, rb = a.results(), b.results()
        for k in ("L", "Cnf", "Cwf", "cost"):
            assert np.array_equal(ra[k].view(np.uint32), rb[k].view(np.uint32)), k
        assert np.array_equal(ra["status"], rb["status"]) and np.array_equal(a.base(), b.base())
        best = b.best(best)                                              # refilled in place through its cached pointers
        for x, y in zip(a.best(), best):
            assert np.array_equal(x, y)
    with pytest.raises(ValueError):
        b.bind_plan_step(np.zeros((1, 7), np.float32), end, lim)


@pytest.mark.parametrize("scene", ["robocrane7", "armlab", "robocrane_task", "planner_task", "stacking_task"])
def test_random_configurations_gpu_equals_replica_bitwise(scene):
    """the configurations of tests/test_oracle_agreement.py's random sweep (far outside any planner's sampling range) through the
    GPU: the same bits as the replica, which that sweep ties to the cull-free FP64 oracle"""
    rng = np.random.default_rng(123)
    n = 20000
    box = lambda lo, hi: np.column_stack([rng.uniform(lo[i], hi[i], n) for i in range(4)])
    if scene == "robocrane7":
        case, pts = joint_case(), rng.uniform(-np.pi, np.pi, (n, 7))
    elif scene == "armlab":
        case, pts = armlab_case(), rng.uniform(-2.5, 2.5, (n, 6))
    elif scene == "robocrane_task":
        case, pts = task_case(), box([-0.3, -1.0, -0.05, -3.2], [1.0, 1.0, 0.9, 3.2])
    elif scene == "planner_task":
        case, pts = task_case("planner.xml", "block1"), box([-0.3, -1.0, -0.05, -3.2], [1.0, 1.0, 0.9, 3.2])
    else:
        case, pts = task_case("stacking.xml", "block1"), box([-0.8, -0.8, -0.05, -3.2], [0.8, 0.8, 0.7, 3.2])
    h = _handle(case)
    o2 = o2_for(case)
    Q = pts.astype(np.float32)
    g = h.eval_points(Q)
    r = np.array([o2.point_cost(q) for q in Q], np.float32)
    assert np.array_equal(g.view(np.uint32), r.view(np.uint32)), int((g.view(np.uint32) != r.view(np.uint32)).sum())
    assert 0.02 < (g != 0).mean() < 0.98
    h.close()
