#!/usr/bin/env python
"""Cross-check of the MuJoCo-free scene path against MuJoCo itself, where MuJoCo is installed.

The reference takes forward kinematics and contacts from MuJoCo (mj_forward at
/root/reference/include/sspp.h:143 and include/Collision.h:53,86; `ncon > 0` at sspp.h:144,
`contact[i].dist` at Collision.h:93).  This framework restates both (the MJCF-subset loader, the
FP64 oracle O1, the FP32 kernel arithmetic), and neither this container nor the GPU box has
MuJoCo, so parity with MuJoCo proper is pinned only where somebody runs this tool:

    python tools/mujoco_crosscheck.py dump  fixtures/robocrane.xml --mode task --body gripper_collision_with_block/ \
           -n 200 -o tests/golden/mujoco/robocrane_task.json          # needs `import mujoco` (3.2.5 is the reference's pin)
    python tools/mujoco_crosscheck.py check tests/golden/mujoco/robocrane_task.json     # needs only this repo

`dump` samples configurations, runs mj_forward on each and records geom_xpos / geom_xmat of the
primitive geoms, ncon and every contact (geom ids, dist).  `check` replays the same configurations
through the loader + O1 and reports the largest frame deviation, collision-flag disagreements
outside a contact band, and the deviation of each geom pair's signed distance from MuJoCo's
smallest contact dist for that pair.  tests/test_mujoco_goldens.py runs `check` on every golden
file it finds (none are committed unless generated with MuJoCo; the file says which MuJoCo made it).

The dump is plain JSON so that it can travel: {"source", "mujoco_version", "mjcf", "mjcf_sha256",
"mode", "dof", "body", "geom_names", "samples": [{"q", "xpos", "xmat", "ncon", "contacts"}]}.
"""
from __future__ import annotations

import argparse
import hashlib
import json
import sys
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))

SKIPPED_TYPES = {"mesh", "hfield", "ellipsoid", "sdf"}     # not on the scoring path (visual-only in the shipped scenes)


def sample_configs(mode: str, dof: int, n: int, seed: int, lo=None, hi=None) -> np.ndarray:
    rng = np.random.default_rng(seed)
    if mode == "task":
        lo = np.array([0.0, -0.7, 0.0, -1.6] if lo is None else lo, float)
        hi = np.array([0.7, 0.7, 0.6, 1.6] if hi is None else hi, float)
    else:
        lo = np.full(dof, -2.0) if lo is None else np.asarray(lo, float)
        hi = np.full(dof, 2.0) if hi is None else np.asarray(hi, float)
    return rng.uniform(lo, hi, (n, len(lo)))


# --------------------------------------------------------------------------------------------------
# dump: MuJoCo side
# --------------------------------------------------------------------------------------------------
def yaw_quat(yaw: float) -> np.ndarray:
    """(w, x, y, z) of a rotation about z — the pose rule of mj_set_point (include/utility.h:149-175, 198-206)"""
    return np.array([np.cos(yaw / 2), 0.0, 0.0, np.sin(yaw / 2)])


def dump(mjcf: Path, mode: str, dof: int, body: str, qs: np.ndarray) -> dict:
    import mujoco   # deliberately imported here: the rest of the tool works without it

    model = mujoco.MjModel.from_xml_path(str(mjcf))
    data = mujoco.MjData(model)
    type_name = {int(getattr(mujoco.mjtGeom, a)): a[len("mjGEOM_"):].lower() for a in dir(mujoco.mjtGeom) if a.startswith("mjGEOM_")}
    keep = []          # MuJoCo geom ids the loader keeps, in its order (document order = MuJoCo's id order)
    for g in range(model.ngeom):
        t = type_name.get(int(model.geom_type[g]), "?")
        if t in SKIPPED_TYPES and model.geom_contype[g] == 0 and model.geom_conaffinity[g] == 0:
            continue
        keep.append(g)
    local = {g: i for i, g in enumerate(keep)}
    names = [mujoco.mj_id2name(model, mujoco.mjtObj.mjOBJ_GEOM, g) or "" for g in keep]
    qadr = None
    if mode == "task":
        bid = mujoco.mj_name2id(model, mujoco.mjtObj.mjOBJ_BODY, body)
        if bid < 0:
            raise SystemExit(f"body '{body}' not in the model")
        jid = model.body_jntadr[bid]
        if jid < 0 or model.jnt_type[jid] != mujoco.mjtJoint.mjJNT_FREE:
            raise SystemExit(f"body '{body}' has no free joint")
        qadr = int(model.jnt_qposadr[jid])
    samples = []
    for q in qs:
        data.qpos[:] = model.qpos0
        if mode == "task":
            data.qpos[qadr:qadr + 3] = q[:3]
            data.qpos[qadr + 3:qadr + 7] = yaw_quat(q[3])
        else:
            data.qpos[:dof] = q
        mujoco.mj_forward(model, data)
        contacts = []
        for c in data.contact[:data.ncon]:
            if c.geom1 in local and c.geom2 in local:
                contacts.append([local[int(c.geom1)], local[int(c.geom2)], float(c.dist)])
        samples.append(dict(q=[float(x) for x in q], xpos=data.geom_xpos[keep].tolist(),
                            xmat=data.geom_xmat[keep].reshape(-1, 3, 3).tolist(), ncon=int(data.ncon), contacts=contacts))
    return dict(source="mujoco", mujoco_version=mujoco.__version__, mjcf=mjcf.name,
                mjcf_sha256=hashlib.sha256(mjcf.read_bytes()).hexdigest(), mode=mode, dof=dof, body=body,
                geom_names=names, samples=samples)


def dump_selftest(mjcf: Path, mode: str, dof: int, body: str, qs: np.ndarray) -> dict:
    """The same file written from the FP64 oracle instead of MuJoCo: exercises the format and `check`
    on a box without MuJoCo.  Marked source="o1-selftest": it proves nothing about MuJoCo parity."""
    from sspp_b200 import capi
    from oracle import pyoracle
    m = capi.Model(path=str(mjcf))
    o1 = pyoracle.O1(m, 1 if mode == "task" else 0, dof, body)
    pairs = o1.pairs()
    samples = []
    for q in qs:
        xpos, xmat = o1.geom_frames(q)
        contacts = []
        for g1, g2 in pairs:
            d = o1.pair_dist(q, int(g1), int(g2))
            if d < m.arrays()["geom_margin"][[g1, g2]].max():
                contacts.append([int(g1), int(g2), float(d)])
        samples.append(dict(q=[float(x) for x in q], xpos=xpos.tolist(), xmat=xmat.tolist(), ncon=len(contacts), contacts=contacts))
    return dict(source="o1-selftest", mujoco_version=None, mjcf=mjcf.name, mjcf_sha256=hashlib.sha256(mjcf.read_bytes()).hexdigest(),
                mode=mode, dof=dof, body=body, geom_names=[m.geom_name(g) for g in range(m.arrays()["ngeom"])], samples=samples)


# --------------------------------------------------------------------------------------------------
# check: this repo's side
# --------------------------------------------------------------------------------------------------
def check(golden: dict, mjcf_dir: Path, frame_tol=1e-9, dist_tol=1e-6, band=1e-6) -> dict:
    """Replay a dump through the loader + O1.  Returns a report; report["ok"] is the verdict.

    frames:   |xpos - golden| and |xmat - golden| <= frame_tol for every kept geom
    flags:    (ncon > 0) equal, except where the oracle's smallest signed distance minus margin is
              within `band` of zero (contact decided by the last bits)
    dists:    for every geom pair MuJoCo reports, min over its contacts of dist vs the oracle's signed
              distance for the pair, <= dist_tol.  MuJoCo's box-box and cylinder-box routines report
              per-contact-point depths whose minimum is the pair's penetration, which is what O1 states.
    """
    from sspp_b200 import capi
    from oracle import pyoracle
    path = mjcf_dir / golden["mjcf"]
    sha = hashlib.sha256(path.read_bytes()).hexdigest()
    rep = dict(file=str(path), source=golden.get("source"), mujoco_version=golden.get("mujoco_version"), n=len(golden["samples"]),
               sha_matches=(sha == golden["mjcf_sha256"]), max_pos_err=0.0, max_mat_err=0.0, flag_mismatch=0, flag_ambiguous=0,
               max_dist_err=0.0, pairs_missing=0, pairs_extra=0, worst=None)
    m = capi.Model(path=str(path))
    a = m.arrays()
    names = [m.geom_name(g) for g in range(a["ngeom"])]
    rep["geoms_match"] = names == golden["geom_names"]
    if not rep["geoms_match"] or not rep["sha_matches"]:
        rep["ok"] = False
        return rep
    mode = 1 if golden["mode"] == "task" else 0
    o1 = pyoracle.O1(m, mode, golden["dof"], golden["body"])
    active = {tuple(sorted(p)) for p in o1.pairs().tolist()}
    margin = a["geom_margin"]
    for s in golden["samples"]:
        q = np.asarray(s["q"], float)
        xpos, xmat = o1.geom_frames(q)
        rep["max_pos_err"] = max(rep["max_pos_err"], float(np.abs(xpos - np.asarray(s["xpos"])).max()))
        rep["max_mat_err"] = max(rep["max_mat_err"], float(np.abs(xmat - np.asarray(s["xmat"])).max()))
        per_pair = {}
        for g1, g2, d in s["contacts"]:
            k = tuple(sorted((int(g1), int(g2))))
            per_pair[k] = min(per_pair.get(k, np.inf), d)
        closest = np.inf
        mine = set()
        for k in active:
            d = o1.pair_dist(q, k[0], k[1])
            mg = max(margin[k[0]], margin[k[1]])
            closest = min(closest, abs(d - mg))
            if d < mg:
                mine.add(k)
        flag_ref, flag_mine = s["ncon"] > 0, len(mine) > 0
        if flag_ref != flag_mine:
            if closest <= band: rep["flag_ambiguous"] += 1
            else: rep["flag_mismatch"] += 1
        for k, d in per_pair.items():
            if k not in active:
                rep["pairs_extra"] += 1          # MuJoCo made a contact for a pair this loader filters out
                continue
            err = abs(o1.pair_dist(q, k[0], k[1]) - d)
            if err > rep["max_dist_err"]:
                rep["max_dist_err"] = float(err)
                rep["worst"] = dict(q=s["q"], pair=[names[k[0]], names[k[1]]], mujoco=d, oracle=o1.pair_dist(q, k[0], k[1]))
        rep["pairs_missing"] += len([k for k in mine if k not in per_pair and closest > band])
    rep["ok"] = bool(rep["max_pos_err"] <= frame_tol and rep["max_mat_err"] <= frame_tol and rep["flag_mismatch"] == 0 and
                     rep["max_dist_err"] <= dist_tol and rep["pairs_extra"] == 0 and rep["pairs_missing"] == 0)
    return rep


def main(argv=None) -> int:
    ap = argparse.ArgumentParser(description=__doc__.split("\n\n")[0])
    sub = ap.add_subparsers(dest="cmd", required=True)
    d = sub.add_parser("dump", help="record goldens with MuJoCo (or --selftest: with the FP64 oracle)")
    d.add_argument("mjcf", type=Path)
    d.add_argument("--mode", choices=["joint", "task"], default="joint")
    d.add_argument("--dof", type=int, default=None, help="planned coordinates (joint mode; default: all of qpos)")
    d.add_argument("--body", default="", help="free body that is planned (task mode)")
    d.add_argument("-n", type=int, default=200)
    d.add_argument("--seed", type=int, default=20250101)
    d.add_argument("--selftest", action="store_true")
    d.add_argument("-o", "--out", type=Path, required=True)
    c = sub.add_parser("check", help="replay goldens through the loader and the FP64 oracle")
    c.add_argument("golden", type=Path, nargs="+")
    c.add_argument("--mjcf-dir", type=Path, default=ROOT / "fixtures")
    c.add_argument("--dist-tol", type=float, default=1e-6)
    args = ap.parse_args(argv)
    if args.cmd == "dump":
        dof = args.dof
        if args.mode == "task":
            dof = 4
        elif dof is None:
            from sspp_b200 import capi
            dof = int(capi.Model(path=str(args.mjcf)).arrays()["nq"])
        qs = sample_configs(args.mode, dof, args.n, args.seed)
        try:
            out = (dump_selftest if args.selftest else dump)(args.mjcf, args.mode, dof, args.body, qs)
        except ImportError as e:
            print(f"mujoco is not importable here ({e}); nothing written. Use --selftest to exercise the format.", file=sys.stderr)
            return 2
        args.out.parent.mkdir(parents=True, exist_ok=True)
        args.out.write_text(json.dumps(out))
        print(f"wrote {args.out}: {len(out['samples'])} samples, source={out['source']}")
        return 0
    ok = True
    for g in args.golden:
        rep = check(json.loads(g.read_text()), args.mjcf_dir, dist_tol=args.dist_tol)
        print(json.dumps(rep))
        ok &= rep["ok"]
    return 0 if ok else 1


if __name__ == "__main__":
    sys.exit(main())
