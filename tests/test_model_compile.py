"""The scene compiler behind ssppcu_create (sspp_b200/csrc/sspp_model.cpp): workspace-bound pruning of geom pairs
is exact — a pruned pair can never be in contact — checked against the FP64 oracle O1, which keeps every pair."""
import ctypes as C

import numpy as np

from tests.helpers import joint_case, task_case, o1_for, o2_for, ICRA_BODY


def _pair_counts(case):
    from sspp_b200 import capi
    a, b, c = C.c_int32(), C.c_int32(), C.c_int32()
    err = C.create_string_buffer(256)
    rc = capi.lib.ssppcu_model_pair_counts(case.model._h, case.cfg.mode, case.cfg.dof, (case.body or "").encode(),
                                           C.byref(a), C.byref(b), C.byref(c), err, 256)
    assert rc == 0, err.value
    return a.value, b.value, c.value


def test_pair_counts_after_pruning():
    # robocrane7: 16 arm/tool geoms x (floor + 3 boxes) = 64 filtered pairs; the arm hangs from z = 1.2 m, so the
    # proximal links can never reach the floor or the blocks: those pairs are dropped when the scene is compiled
    ms, mm, ss = _pair_counts(joint_case())
    assert mm == 0 and ss == 0
    assert 30 <= ms < 64
    # a free body is unbounded: nothing may be pruned (robocrane.xml gripper: 48 active pairs)
    ms, mm, ss = _pair_counts(task_case("robocrane.xml", ICRA_BODY))
    assert ms == 48 and ss == 12


def test_pruned_pairs_never_touch_at_the_workspace_boundary():
    """configurations that stretch the arm towards every obstacle, including the joint limits: the replica (pruned
    pair table) flags exactly the configurations the FP64 oracle (all pairs) flags"""
    case = joint_case()
    o1, o2 = o1_for(case), o2_for(case)
    rng = np.random.default_rng(5)
    lim = np.array([2.967, 2.094, 2.967, 2.094, 2.967, 2.094, 3.054])
    Q = [np.zeros(7)]
    for s in (-1.0, 1.0):
        for j in range(7):
            q = np.zeros(7); q[j] = s * lim[j]; Q.append(q)
    Q += list(rng.choice([-1.0, 0.0, 1.0], (400, 7)) * lim)            # corners of the joint box
    Q += list(rng.uniform(-1, 1, (1500, 7)) * lim)
    bad = 0
    for q in Q:
        c1, ncon, md = o1.point_cost(q)
        c2 = o2.point_cost(q)
        if (c1 != 0.0) != (c2 != 0.0) and abs(md) > 1e-4:
            bad += 1
    assert bad == 0


def test_coordinates_outside_jnt_range_keep_their_contacts():
    """Nothing on the path clamps the planned coordinates to jnt_range (neither does the reference: sspp.h:139-142 writes qpos
    as sampled), so pruning must not rely on the range: hinges far outside their range and a slide pushed well beyond its
    range still see every obstacle the FP64 oracle (all pairs) sees"""
    from tests.helpers import armlab_case
    rng = np.random.default_rng(8)
    for case, scale in ((joint_case(), np.full(7, 3.0 * np.pi)), (armlab_case(), np.array([6.0, 6.0, 6.0, 6.0, 6.0, 1.5]))):
        o1, o2 = o1_for(case), o2_for(case)
        Q = rng.uniform(-1, 1, (1500, case.cfg.dof)) * scale             # the armlab slide (range 0.04 m) travels +-1.5 m
        bad = hits = 0
        for q in Q:
            c1, ncon, md = o1.point_cost(q)
            c2 = o2.point_cost(q)
            hits += c1 != 0.0
            if (c1 != 0.0) != (c2 != 0.0) and abs(md) > 1e-4:
                bad += 1
        assert bad == 0 and hits > 20
