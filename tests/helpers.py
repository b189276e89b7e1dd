"""Shared case builders and GPU-vs-oracle comparison used by tests/, smoke() and bench.py's checker leg.

Nothing here reads /root/reference: scenes come from fixtures/ (the reference's MJCF files are
mirrored there as inputs, see fixtures/README.md)."""
from __future__ import annotations

import ctypes as C
from dataclasses import dataclass, field
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent.parent
FIX = ROOT / "fixtures"
SEED = 0x53535050  # "SSPP"

# the only concrete 7-DoF configurations the reference holds (scripts/bspline_params.npy)
Q_START = np.array([1.1805, 0.7884, 1.6656, 1.4216, 0.8346, -1.6386, 1.3560])
Q_END = np.array([0.4323, 0.7750, 1.6877, 1.4162, 0.7718, -1.6073, 0.6237])
# the ICRA query: body points of block_green/ and block_orange/ lifted by 0.02 (src/main_icra_benchmark.cpp:176-178)
ICRA_START = np.array([0.5, 0.15, 0.136, np.pi / 2])
ICRA_END = np.array([0.5, -0.05, 0.136, np.pi / 2])
ICRA_BODY = "gripper_collision_with_block/"


@dataclass
class Case:
    mjcf: Path
    cfg: object                 # capi.PlannerConfig
    start: np.ndarray           # [Q][D]
    end: np.ndarray
    sigma0: float = 0.3
    limits: np.ndarray | None = None   # JOINT: per-dof noise scale
    joint_sigma: float = 0.08
    body: str = ""
    model: object = field(default=None, repr=False)

    def prepare(self, h):
        """put a fresh handle into the state the planners start from"""
        h.set_queries(self.start, self.end)
        if self.cfg.mode == 0:
            h.set_joint_noise(self.joint_sigma, self.limits)
        else:
            h.reset_distribution(self.sigma0)


def _capi():
    from sspp_b200 import capi
    return capi


def joint_case(n_samples=64, check_points=16, n_points=10, do_update=False, n_queries=1, mjcf=None,
               limit_scale=0.0, w_collision=0.0) -> Case:
    """Planner A shape (sspp.h:194-232) on the robocrane7 fixture; scripts/main.py:168-173 noise."""
    capi = _capi()
    cfg = capi.default_config(capi.MODE_JOINT)
    cfg.dof, cfg.degree, cfg.n_points = 7, 3, n_points
    cfg.check_points, cfg.n_samples, cfg.n_queries = check_points, n_samples, n_queries
    cfg.do_update = int(do_update)
    cfg.limit_scale, cfg.w_collision = limit_scale, w_collision
    cfg.stddev_min, cfg.stddev_max = 1e-4, 2.0
    path = Path(mjcf) if mjcf else FIX / "robocrane7.xml"
    rng = np.random.default_rng(7)
    start = np.tile(Q_START, (n_queries, 1)); end = np.tile(Q_END, (n_queries, 1))
    if n_queries > 1:
        start[1:] += rng.normal(0, 0.05, (n_queries - 1, 7)); end[1:] += rng.normal(0, 0.05, (n_queries - 1, 7))
    return Case(path, cfg, start, end, limits=np.full(7, np.pi), model=capi.Model(path=str(path)))


# collision-free end points whose connecting paths are about 85 % free at the default noise (found by a seeded search)
ARMLAB_START = np.array([1.704, -0.726, -0.67, 1.41, 0.307, 0.031])
ARMLAB_END = np.array([-1.691, 0.745, -0.453, -1.473, 0.578, 0.008])


def armlab_case(n_samples=64, check_points=16, n_points=10, do_update=False, n_queries=1, w_collision=0.0) -> Case:
    """Planner A shape on fixtures/armlab.xml: 6 planned coordinates, two of them joints of one body, one a
    slide joint inside a frame; capsule / sphere / cylinder / box partners."""
    capi = _capi()
    cfg = capi.default_config(capi.MODE_JOINT)
    cfg.dof, cfg.degree, cfg.n_points = 6, 3, n_points
    cfg.check_points, cfg.n_samples, cfg.n_queries = check_points, n_samples, n_queries
    cfg.do_update = int(do_update)
    cfg.w_collision = w_collision
    cfg.stddev_min, cfg.stddev_max = 1e-4, 2.0
    path = FIX / "armlab.xml"
    rng = np.random.default_rng(9)
    start = np.tile(ARMLAB_START, (n_queries, 1)); end = np.tile(ARMLAB_END, (n_queries, 1))
    if n_queries > 1:
        start[1:, :5] += rng.normal(0, 0.05, (n_queries - 1, 5)); end[1:, :5] += rng.normal(0, 0.05, (n_queries - 1, 5))
    limits = np.array([np.pi, np.pi, np.pi, np.pi, np.pi, 0.04])
    return Case(path, cfg, start, end, limits=limits, model=capi.Model(path=str(path)))


def task_case(scene="robocrane.xml", body=ICRA_BODY, n_samples=15, check_points=40, n_points=3, n_queries=1,
              start=None, end=None, faithful_quirks=True) -> Case:
    """Planner B shape with the parameters of src/main_icra_benchmark.cpp:152-160.  faithful_quirks
    reproduces the adapter's effective values: sigma0 = 0.3, dist z_min = stddev_initial,
    evaluator floor at z_min = 0."""
    capi = _capi()
    cfg = capi.default_config(capi.MODE_TASK)
    cfg.dof, cfg.degree, cfg.n_points = 4, 2, n_points
    cfg.check_points, cfg.n_samples, cfg.n_queries = check_points, n_samples, n_queries
    cfg.include_mean = cfg.include_best = cfg.do_update = 1
    cfg.body_name = body.encode()
    cfg.w_collision, cfg.elite_fraction = 1.0, 0.3
    cfg.inc, cfg.dec = 1.5, 0.9
    cfg.stddev_min, cfg.stddev_max, cfg.sigma_floor = 1e-4, 0.5, 0.005
    cfg.var_beta, cfg.mean_lr = 0.2, 0.5
    cfg.sample_zmin = 0.1
    cfg.dist_zmin = 0.2 if faithful_quirks else 0.1
    cfg.eval_zmin = 0.0 if faithful_quirks else 0.1
    cfg.floor_margin, cfg.floor_scale = 0.01, 10.0
    lo = [0.0, -0.7, 0.1, -1.6]; hi = [0.7, 0.7, 0.6, 1.6]   # main_icra_benchmark.cpp:160
    for d in range(4):
        cfg.lo[d], cfg.hi[d] = lo[d], hi[d]
    path = FIX / scene
    s = np.tile(ICRA_START if start is None else start, (n_queries, 1)) if np.ndim(start if start is not None else ICRA_START) == 1 else np.asarray(start)
    e = np.tile(ICRA_END if end is None else end, (n_queries, 1)) if np.ndim(end if end is not None else ICRA_END) == 1 else np.asarray(end)
    return Case(path, cfg, s, e, sigma0=0.3 if faithful_quirks else 0.2, body=body, model=capi.Model(path=str(path)))


def clutter_scene(n_boxes=64, seed=SEED, keep_clear=()):
    """cfg 3 scene: stacking.xml (floor + 3 boxes, block1 moving) plus n_boxes seeded static boxes,
    half extents U(0.02,0.08), centres U([-0.6,0.6]^2 x [0.05,0.5]); boxes touching the shipped blocks or a
    keep_clear point (start / goal, with the moving block's circumscribed radius) are rejected.  Returns MJCF text."""
    rng = np.random.default_rng(seed)
    base = (FIX / "stacking.xml").read_text()
    blocks = [np.array([0.205, 0, 0.1]), np.array([0.0, 0, 0.1]), np.array([-0.205, 0, 0.1])]
    geoms = []
    while len(geoms) < n_boxes:
        h = rng.uniform(0.02, 0.08, 3)
        c = np.array([rng.uniform(-0.6, 0.6), rng.uniform(-0.6, 0.6), rng.uniform(0.05, 0.5)])
        c[2] = max(c[2], h[2] + 0.002)                       # not sunk into the floor
        if any(np.all(np.abs(c - b) < h + 0.1 + 0.01) for b in blocks):
            continue
        if any(np.linalg.norm(np.maximum(np.abs(c - np.asarray(p)[:3]) - h, 0)) < 0.1 * np.sqrt(3) + 0.02 for p in keep_clear):
            continue
        geoms.append(f'    <geom name="clutter{len(geoms)}" type="box" size="{h[0]:.6f} {h[1]:.6f} {h[2]:.6f}" '
                     f'pos="{c[0]:.6f} {c[1]:.6f} {c[2]:.6f}" contype="1" conaffinity="1"/>')
    return base.replace("  <worldbody>\n", "  <worldbody>\n" + "\n".join(geoms) + "\n", 1)


CFG3_START = np.array([0.5, -0.5, 0.3, 0.0])
CFG3_END = np.array([-0.5, 0.5, 0.3, 1.0])


def sdf_case(n_samples=254, check_points=64, n_points=5, dims=(64, 64, 64), n_boxes=64, n_queries=1) -> Case:
    """cfg 3 shape: Planner B on the cluttered stacking scene, SDF-grid collision, via points sampled"""
    capi = _capi()
    text = clutter_scene(n_boxes, keep_clear=(CFG3_START, CFG3_END))
    path = ROOT / "fixtures" / f"_cfg3_clutter{n_boxes}.xml"
    if not path.exists() or path.read_text() != text:
        path.write_text(text)
    case = task_case(f"_cfg3_clutter{n_boxes}.xml", "block1", n_samples=n_samples, check_points=check_points, n_points=n_points,
                     n_queries=n_queries, start=CFG3_START, end=CFG3_END)
    cfg = case.cfg
    cfg.collision_mode = capi.COLLIDE_SDF
    cfg.use_center_dist = 0
    cfg.sdf_sphere_ratio = 0.5
    lo, hi = [-0.8, -0.8, 0.0], [0.8, 0.8, 0.9]
    for a in range(3):
        cfg.sdf_dims[a], cfg.sdf_lo[a], cfg.sdf_hi[a] = dims[a], lo[a], hi[a]
    for d, (l, h) in enumerate(zip([-0.7, -0.7, 0.1, -1.6], [0.7, 0.7, 0.78, 1.6])):   # paths may climb over the clutter
        cfg.lo[d], cfg.hi[d] = l, h
    return case


def o2_for(case: Case):
    from oracle import pyoracle
    return pyoracle.O2(str(case.mjcf), case.cfg)


def o1_for(case: Case, spheres=None):
    """FP64 oracle of the case; SDF-mode cases also need the product's sample spheres (an input, like the noise)"""
    from oracle import pyoracle
    o1 = pyoracle.O1(case.model, case.cfg.mode, case.cfg.dof, case.body)
    if case.cfg.collision_mode == 1:
        if spheres is None:
            raise ValueError("SDF-mode oracle needs the sample spheres")
        c = case.cfg
        o1.set_sdf(list(c.sdf_dims), list(c.sdf_lo), list(c.sdf_hi), spheres, 0.0)
    return o1


def compare_iteration(h, case: Case, seed=SEED, iters=1, query=0, check_o1=True, eps_contact=1e-4, o1_subsample=512):
    """Run `iters` fused iterations on the GPU handle and in the replica oracle O2 from the same
    state; compare every per-candidate output bit for bit, and (check_o1) the FP64 oracle O1 within
    tolerance on the GPU's own sampled rows - all of them, or a seeded subsample of `o1_subsample`
    candidates when there are more than 1000.  Returns a report dict of mismatch counts:
      o1_flag_mismatch   collision flag differs from O1 with no decision of the candidate inside the contact band
      o1_max_rel         L over candidates whose flag agrees; Cnf / Cwf / cost over candidates with no in-band decision"""
    o2 = o2_for(case)
    psi_col, psi_arc = h.collocation()
    base = h.base()[query]
    mean, sigma = h.get_distribution()
    mean, sigma = mean[query].astype(np.float32), sigma[query].astype(np.float32)
    best, bvalid = h.best()
    best, bvalid = best[query], int(bvalid[query])
    rep = dict(status_mismatch=0, elite_mismatch=0, key_mismatch=0, value_bits_mismatch=0, param_bits_mismatch=0,
               dist_bits_mismatch=0, o1_flag_mismatch=0, o1_flag_ambiguous=0, o1_max_rel=0.0, n_success=[])
    if check_o1:
        from oracle import pyoracle
        o1 = o1_for(case, h.spheres() if case.cfg.collision_mode == 1 else None)
        oc = pyoracle.o1_cfg_from(case.cfg)
    for it in range(iters):
        h.iterate(seed, it, 1)
        r = h.results()
        gp = h.params()[query]
        ref = o2.iterate(psi_col, psi_arc if case.cfg.mode == 0 else None, base, mean, sigma, best, bvalid, seed, it, query)
        mean, sigma, best, bvalid = ref["mean"], ref["sigma"], ref["best"], ref["best_valid"]
        rep["status_mismatch"] += int((r["status"][query] != ref["status"]).sum())
        rep["param_bits_mismatch"] += int((gp.view(np.uint32) != ref["params"].view(np.uint32)).sum())
        for k in ("L", "Cnf", "Cwf", "cost"):
            rep["value_bits_mismatch"] += int((r[k][query].view(np.uint32) != ref[k].view(np.uint32)).sum())
        idx, cost, w = h.elites(query)
        gkeys = idx.astype(np.uint64)
        rep["key_mismatch"] += int((gkeys != (ref["keys"] & np.uint64(0xFFFFFFFF))).sum())
        k = ref["n_elite"]
        rep["elite_mismatch"] += int(r["n_elite"][query] != k) + int(r["n_success"][query] != ref["n_success"])
        rep["elite_mismatch"] += int((w[:k].view(np.uint32) != ref["weights"].view(np.uint32)).sum())
        if ref["n_success"] > 0:
            rep["elite_mismatch"] += int(r["best_idx"][query] != ref["best_idx"])
        gm, gs = h.get_distribution()
        rep["dist_bits_mismatch"] += int((gm[query].astype(np.float32).view(np.uint32) != mean.view(np.uint32)).sum())
        rep["dist_bits_mismatch"] += int((gs[query].astype(np.float32).view(np.uint32) != sigma.view(np.uint32)).sum())
        rep["n_success"].append(int(r["n_success"][query]))
        if check_o1:
            # O1 consumes the GPU's sampled rows; for JOINT mode base already holds control points
            n_all = len(gp)
            sel = np.arange(n_all) if n_all <= 1000 else np.sort(np.random.default_rng(seed + it).choice(n_all, o1_subsample, replace=False))
            s1 = o1.score(oc, base.astype(np.float64), gp.astype(np.float64).reshape(n_all, -1)[sel])
            st_g = r["status"][query][sel]
            live = st_g != 2
            diff = ((st_g == 0) != (s1["status"] == 0)) & live
            clean = s1["band"] == 0                               # no pair of any point of the path within eps of a threshold
            rep["o1_flag_ambiguous"] += int((diff & ~clean).sum())
            rep["o1_flag_mismatch"] += int((diff & clean).sum())
            rep["o1_checked"] = rep.get("o1_checked", 0) + int(live.sum())
            ok = live & ~diff
            rel = np.abs(r["L"][query][sel][ok] - s1["L"][ok]) / np.maximum(np.abs(s1["L"][ok]), 1e-6)
            if rel.size:
                rep["o1_max_rel"] = max(rep["o1_max_rel"], float(rel.max()))
            okc = ok & clean
            for k in ("Cnf", "Cwf", "cost"):                      # collision sums and the total cost, failed candidates included
                g = r[k][query][sel][okc].astype(np.float64)
                rel = np.abs(g - s1[k][okc]) / np.maximum(np.abs(s1[k][okc]), 1.0)
                if rel.size:
                    rep["o1_max_rel"] = max(rep["o1_max_rel"], float(rel.max()))
    return rep
