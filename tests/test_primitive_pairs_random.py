"""Every primitive pairing the scene compiler accepts, on random scenes: a geom of one type at the end of a three-hinge chain
(yaw, pitch, roll: any orientation), six static geoms of another type at random poses (or a plane under the chain), random
margins.  The FP32 replica (product arithmetic, compile-time pruning, per-point culls) must count the contacts the FP64 oracle -
which tests every filtered pair with its own narrow phase and knows no cull - counts, outside EPS of a decision threshold;
on a GPU the kernel must give the replica's bits.  Semantics restated from MuJoCo's documented behaviour:
contact kept iff signed distance < max(margin_1, margin_2)."""
import numpy as np
import pytest

from oracle import pyoracle
from sspp_b200 import capi

SIZES = dict(sphere=lambda r: f'size="{r.uniform(0.03, 0.12):.4f}"',
             capsule=lambda r: f'size="{r.uniform(0.02, 0.08):.4f} {r.uniform(0.05, 0.45):.4f}"',
             cylinder=lambda r: f'size="{r.uniform(0.03, 0.1):.4f} {r.uniform(0.03, 0.3):.4f}"',
             box=lambda r: f'size="{r.uniform(0.02, 0.3):.4f} {r.uniform(0.02, 0.3):.4f} {r.uniform(0.01, 0.2):.4f}"',
             plane=lambda r: 'size="0 0 0.01"')
PAIRS = [("sphere", "plane"), ("capsule", "plane"), ("cylinder", "plane"), ("box", "plane"), ("sphere", "sphere"), ("sphere", "capsule"),
         ("capsule", "sphere"), ("sphere", "cylinder"), ("cylinder", "sphere"), ("sphere", "box"), ("box", "sphere"), ("capsule", "capsule"),
         ("capsule", "box"), ("box", "capsule"), ("cylinder", "box"), ("box", "cylinder"), ("box", "box")]
EPS = 1e-4


def scene_xml(moving: str, static: str, rng, margin: float) -> str:
    stat = []
    for i in range(1 if static == "plane" else 6):
        pos = "0 0 0" if static == "plane" else f"{rng.uniform(-0.6, 0.6):.3f} {rng.uniform(-0.6, 0.6):.3f} {rng.uniform(0.0, 0.9):.3f}"
        eul = "0 0 0" if static == "plane" else f"{rng.uniform(-180, 180):.1f} {rng.uniform(-180, 180):.1f} {rng.uniform(-180, 180):.1f}"
        stat.append(f'<geom name="s{i}" type="{static}" {SIZES[static](rng)} pos="{pos}" euler="{eul}" margin="{margin}"/>')
    z0 = 0.12 if static == "plane" else 0.5
    return f'''<mujoco><compiler angle="degree"/><worldbody>
{chr(10).join(stat)}
<body name="b1" pos="0 0 {z0}"><joint name="j1" type="hinge" axis="0 0 1"/>
 <body name="b2" pos="0.1 0 0"><joint name="j2" type="hinge" axis="0 1 0"/>
  <body name="b3" pos="0.15 0 0"><joint name="j3" type="hinge" axis="1 0 0"/>
   <geom name="m" type="{moving}" {SIZES[moving](rng)} pos="0.1 0.02 0" margin="{margin}"/>
  </body></body></body></worldbody></mujoco>'''


def build(tmp_path, moving, static, rep):
    margin = (0.0, 0.004, 0.02)[rep]
    rng = np.random.default_rng(1000 * PAIRS.index((moving, static)) + rep)
    path = tmp_path / f"pair_{moving}_{static}_{rep}.xml"
    path.write_text(scene_xml(moving, static, rng, margin))
    cfg = capi.default_config(capi.MODE_JOINT)
    cfg.dof, cfg.degree, cfg.n_points = 3, 3, 8
    cfg.check_points, cfg.n_samples, cfg.n_queries = 16, 16, 1
    pts = rng.uniform(-np.pi, np.pi, (1500, 3)).astype(np.float32)
    return path, cfg, margin, pts


@pytest.mark.parametrize("moving,static", PAIRS)
def test_replica_counts_the_contacts_the_fp64_oracle_counts(tmp_path, moving, static):
    hits = 0
    for rep in range(3):
        path, cfg, margin, pts = build(tmp_path, moving, static, rep)
        model = capi.Model(path=str(path))
        o1 = pyoracle.O1(model, 0, 3, "")
        o2 = pyoracle.O2(str(path), cfg)
        bad = band = 0
        for q in pts.astype(np.float64):
            c1, ncon, md = o1.point_cost(q)
            c2 = o2.point_cost(q)
            hits += c1 != 0
            if c1 != c2:
                if abs(md - margin) < EPS or abs(md) < EPS:
                    band += 1
                else:
                    bad += 1
        assert bad == 0 and band <= 3, (moving, static, margin, bad, band)
    assert hits > 30, "the scenes must produce contacts"


@pytest.mark.gpu
@pytest.mark.parametrize("moving,static", PAIRS)
def test_gpu_gives_the_replicas_bits_on_every_pairing(tmp_path, moving, static):
    for rep in range(3):
        path, cfg, margin, pts = build(tmp_path, moving, static, rep)
        model = capi.Model(path=str(path))
        h = capi.Handle(model, cfg)
        o2 = pyoracle.O2(str(path), cfg)
        g = h.eval_points(pts)
        r = np.array([o2.point_cost(q) for q in pts], np.float32)
        assert np.array_equal(g.view(np.uint32), r.view(np.uint32)), (moving, static, margin)
        h.close()
