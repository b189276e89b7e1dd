"""Algorithmic work per scored collocation point (constants frozen here).

The scoring path moves ~16 B per CANDIDATE (L, Cnf, Cwf, cost, status, key) and reads a few KiB
of constants, so HBM is not the bound; FP32 issue is.  `flop_per_point` is the NOMINAL model the
roofline fraction is quoted on: every filtered geom pair is charged its full narrow phase, although
the kernel's bounding-sphere cull skips most of them (so nominal FLOP/s can exceed what the FP32
pipe physically executed; the ncu pipe utilisation in profiles/ is the physical measure).
FMA = 2 FLOP.
"""
from __future__ import annotations

import numpy as np

PLANE, SPHERE, CAPSULE, CYLINDER, BOX = 0, 2, 3, 5, 6

# narrow phase FLOP per pair, keyed by sorted (type_a, type_b)
PAIR_FLOP = {
    (PLANE, SPHERE): 8, (PLANE, CAPSULE): 17, (PLANE, CYLINDER): 22, (PLANE, BOX): 34,
    (SPHERE, SPHERE): 12, (SPHERE, CAPSULE): 30, (SPHERE, CYLINDER): 30, (SPHERE, BOX): 42,
    (CAPSULE, CAPSULE): 90, (CAPSULE, BOX): 126, (CYLINDER, BOX): 300, (BOX, BOX): 300,
}
PLACE_FLOP = {SPHERE: 18, CAPSULE: 36, CYLINDER: 72, BOX: 72}
FK_HINGE, FK_SLIDE, FK_FREE4 = 165, 90, 12
BYTES_PER_CANDIDATE = 4 * 4 + 1 + 8      # L, Cnf, Cwf, cost, status, sort key


def _pair_active(a, g1, g2) -> bool:
    b1, b2 = a["geom_body"][g1], a["geom_body"][g2]
    w1, w2 = a["body_weld"][b1], a["body_weld"][b2]
    if w1 == w2:
        return False
    if w1 != 0 and w2 != 0 and (a["body_weld"][a["body_parent"][w1]] == w2 or a["body_weld"][a["body_parent"][w2]] == w1):
        return False
    for e1, e2 in a["exclude"]:
        if (e1, e2) in ((b1, b2), (b2, b1)):
            return False
    if a["geom_type"][g1] == PLANE and a["geom_type"][g2] == PLANE:
        return False
    return bool((a["geom_contype"][g1] & a["geom_conaffinity"][g2]) or (a["geom_contype"][g2] & a["geom_conaffinity"][g1]))


def moving_bodies(a, mode: int, dof: int, body_id: int) -> np.ndarray:
    """per body: does it move with the planned coordinates"""
    nb = a["nbody"]
    moves = np.zeros(nb, bool)
    for b in range(1, nb):
        planned = (a["body_jtype"][b] in (2, 3) and 0 <= a["body_qadr"][b] < dof) if mode == 0 else (b == body_id)
        moves[b] = planned or moves[a["body_parent"][b]]
    return moves


def work_model(model, mode: int, dof: int, n_points: int, body: str = "") -> dict:
    """nominal FLOP per scored collocation point and its breakdown for a planner shape"""
    a = model.arrays()
    moves = moving_bodies(a, mode, dof, model.body_id(body) if mode == 1 else -1)
    ng = a["ngeom"]
    pair_hist: dict = {}
    touched = set()
    for g1 in range(ng):
        for g2 in range(g1 + 1, ng):
            m1, m2 = moves[a["geom_body"][g1]], moves[a["geom_body"][g2]]
            if not (m1 or m2) or not _pair_active(a, g1, g2):
                continue
            key = tuple(sorted((int(a["geom_type"][g1]), int(a["geom_type"][g2]))))
            pair_hist[key] = pair_hist.get(key, 0) + 1
            if m1: touched.add(g1)
            if m2: touched.add(g2)
    narrow = sum(PAIR_FLOP[k] * n for k, n in pair_hist.items())
    place = sum(PLACE_FLOP[int(a["geom_type"][g])] for g in touched)
    if mode == 0:
        n_h = sum(1 for b in range(a["nbody"]) if a["body_jtype"][b] == 3 and 0 <= a["body_qadr"][b] < dof)
        n_s = sum(1 for b in range(a["nbody"]) if a["body_jtype"][b] == 2 and 0 <= a["body_qadr"][b] < dof)
        fk = FK_HINGE * n_h + FK_SLIDE * n_s
        sfu = 2 * n_h
    else:
        fk, sfu = FK_FREE4, 2
    spline = 2 * n_points * dof
    cost = 3 * dof + 2 + (4 * dof if mode == 0 else 4)
    total = spline + fk + place + narrow + cost
    return dict(flop_per_point=float(total), spline=spline, fk=fk, placement=place, narrow_phase=narrow, cost_terms=cost,
                sfu_sincos=sfu, n_pairs=int(sum(pair_hist.values())), n_moving_geoms=len(touched),
                pair_hist={f"{k[0]}-{k[1]}": n for k, n in sorted(pair_hist.items())})
