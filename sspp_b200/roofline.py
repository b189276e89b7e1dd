"""Algorithmic work per scored collocation point (constants frozen here).

The scoring path moves ~16 B per CANDIDATE (L, Cnf, Cwf, cost, status, key) and reads a few KiB
of constants, so HBM is not the bound; FP32 issue is.  Two figures (SURVEY 8d asks for both):

  executed_flop()  F_executed: the floating-point operations the kernel's arithmetic performs for the work it
                   actually did, from the counters of ssppcu_work_counters (pairs that reached the narrow phase by
                   type, early exits, Newton steps, placements, broad-phase tests) times the operation count of the
                   corresponding code in sspp_b200/csrc/sspp_math.h.  `roofline.frac` is quoted on THIS figure.
  work_model()     F_nominal: every filtered geom pair charged its full narrow phase, before the compile-time pruning
                   and the per-point broad phase - the work the reference's semantics define (MuJoCo tests every
                   pair it keeps).  Reported as `frac_nominal`; it can exceed what the FP32 pipe executed.
FMA = 2 FLOP; add, mul, sqrt, div = 1; compare / min / max / abs / integer = 0.
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


# ---- F_executed: operation counts of the code paths of sspp_b200/csrc/sspp_math.h (FMA = 2) ----
SINCOS = 24                 # sincos_poly: range reduction 1 mul + 3 FMA, two minimax kernels
EXEC_FLOP = dict(
    fk_hinge_axis=18 + SINCOS + 1 + 18,      # frame origin (R p + p), sincos, sign, rot_axis (two columns mix)
    fk_hinge_gen=18 + SINCOS + 28 + 45,      # Rodrigues matrix + 3x3 product
    fk_slide=18 + 15 + 6,
    fk_free4=SINCOS,
    fk_r0=45,                                # rest orientation: one more 3x3 product
    fk_anchor=36,
    geom_pos=18 + 5,                         # centre + |p|^2 for the bounded partners' test
    geom_rot_box=45, geom_rot_axis=15,
    broad=6,                                 # one affine form: 3 FMA
    body_broad=6,
    np_plane=29,                             # box against a plane (the costliest of the plane forms)
    np_sphere_box=28,
    np_box_cull=28,                          # geom centre into the box frame + point-box distance + bound (before any orientation work)
    np_caps_box_early=29,                    # centre into the box frame + point-box distance + bound
    np_caps_box=29 + 25 + 2 * 19 + 12 + 5 + 15 + 1,   # ... + segment into the box frame, two bracket gradients, six kink parameters (sub + div), the secant step, end evaluation
    np_caps_newton=19,                       # one gradient evaluation at a kink inside the bracket (the counter keeps its first name)
    np_box_box_face=120, np_box_box=275,     # 6 face axes / all 15 axes + depth
    np_cyl_box_face=84, np_cyl_box=292,
    np_other=60,
    contacts=11,                             # centre-distance cost (Collision.h:96); the count rule adds 1.0
    mm_tests=138, mm_narrow=126,
    sdf_samples=49,                          # sphere centre, cell + fractions, trilinear blend, radius
)
SAMPLE_FLOP = 64            # one sampled element: Box-Muller (log, sqrt, sincos) + mean + sigma z


def executed_flop(counters: dict, *, n_cand_total: int, check_points: int, dof: int, n_points: int, degree: int, mode: int,
                  limit_penalty: bool = False) -> dict:
    """F_executed per scored collocation point (N * M points) from the work counters of one iteration.
    n_cand_total: candidates scored (all queries); returns the per-term breakdown too."""
    M = check_points
    pts = float(n_cand_total * M)
    task = mode == 1
    dp = (dof + 3) // 4 * 4
    W = degree + 1
    evals = (M + 1) + (M if task else M)             # collision grid + (TASK: the segment's start point again | JOINT: the arc grid)
    n_seg = M if task else M - 1
    spline = n_cand_total * (evals * 2 * W * dp + n_seg * (3 * dof + 1))
    n_free = n_points - 2 if task else n_points - 2 * degree
    sampling = n_cand_total * n_free * dof * SAMPLE_FLOP
    penalty = counters["points"] * (4 if task else (4 * dof if limit_penalty else 0))
    terms = {k: counters[k] * f for k, f in EXEC_FLOP.items()}
    fk = sum(terms[k] for k in terms if k.startswith("fk_"))
    place = terms["geom_pos"] + terms["geom_rot_box"] + terms["geom_rot_axis"]
    narrow = sum(terms[k] for k in terms if k.startswith("np_") or k.startswith("mm_")) + terms["contacts"]
    total = spline + sampling + penalty + fk + place + terms["broad"] + terms["body_broad"] + narrow + terms["sdf_samples"]
    return dict(flop_per_point_executed=total / pts, spline=spline / pts, sampling=sampling / pts, fk=fk / pts, placement=place / pts,
                broad_phase=(terms["broad"] + terms["body_broad"]) / pts, narrow_phase=narrow / pts, sdf=terms["sdf_samples"] / pts, penalty=penalty / pts,
                pairs_tested_per_point=counters["broad"] / pts,
                pairs_in_narrow_phase_per_point=sum(counters[k] for k in counters if k.startswith("np_") and k not in ("np_caps_newton", "np_box_cull")) / pts,
                contacts_per_point=counters["contacts"] / pts)
