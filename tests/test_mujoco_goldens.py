"""MuJoCo cross-check (tools/mujoco_crosscheck.py).

Neither this container nor the GPU box has MuJoCo, so goldens recorded from MuJoCo are
optional inputs: every tests/golden/mujoco/*.json is replayed through the loader and the FP64 oracle
and must pass; with none present the test that needs them is skipped and says so.  The file format
and the checker are exercised either way with a dump written by the oracle itself (source
"o1-selftest", which proves nothing about MuJoCo and is never committed as a golden)."""
import json
import sys
from pathlib import Path

import numpy as np
import pytest

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT / "tools"))
import mujoco_crosscheck as xc   # noqa: E402

GOLDEN_DIR = ROOT / "tests" / "golden" / "mujoco"
GOLDENS = sorted(GOLDEN_DIR.glob("*.json")) if GOLDEN_DIR.exists() else []


@pytest.mark.parametrize("scene,mode,body,dof", [("robocrane.xml", "task", "gripper_collision_with_block/", 4),
                                                  ("stacking.xml", "task", "block1", 4),
                                                  ("robocrane7.xml", "joint", "", 7), ("armlab.xml", "joint", "", 6)])
def test_selftest_dump_round_trips(tmp_path, scene, mode, body, dof):
    qs = xc.sample_configs(mode, dof, 40, seed=3)
    g = xc.dump_selftest(ROOT / "fixtures" / scene, mode, dof, body, qs)
    p = tmp_path / "g.json"
    p.write_text(json.dumps(g))
    rep = xc.check(json.loads(p.read_text()), ROOT / "fixtures")
    assert rep["ok"], rep
    assert sum(s["ncon"] > 0 for s in g["samples"]) > 0          # the sample does reach contacts


def test_checker_notices_wrong_frames_distances_and_stale_files(tmp_path):
    qs = xc.sample_configs("joint", 6, 30, seed=4)
    g = xc.dump_selftest(ROOT / "fixtures" / "armlab.xml", "joint", 6, "", qs)
    bad = json.loads(json.dumps(g)); bad["samples"][3]["xpos"][5][1] += 1e-6
    assert not xc.check(bad, ROOT / "fixtures")["ok"]
    hit = next(i for i, s in enumerate(g["samples"]) if s["contacts"])
    bad = json.loads(json.dumps(g)); bad["samples"][hit]["contacts"][0][2] -= 1e-4
    rep = xc.check(bad, ROOT / "fixtures")
    assert not rep["ok"] and rep["max_dist_err"] > 5e-5 and rep["worst"] is not None
    bad = json.loads(json.dumps(g)); bad["mjcf_sha256"] = "0" * 64
    assert not xc.check(bad, ROOT / "fixtures")["sha_matches"]
    free = next(i for i, s in enumerate(g["samples"]) if not s["contacts"])
    bad = json.loads(json.dumps(g)); bad["samples"][free]["ncon"] = 1
    assert xc.check(bad, ROOT / "fixtures")["flag_mismatch"] == 1


def test_dump_without_mujoco_fails_loudly(tmp_path):
    try:
        import mujoco  # noqa: F401
        pytest.skip("mujoco is importable here: the real dump path applies")
    except ImportError:
        pass
    rc = xc.main(["dump", str(ROOT / "fixtures" / "armlab.xml"), "-n", "2", "-o", str(tmp_path / "x.json")])
    assert rc == 2 and not (tmp_path / "x.json").exists()


@pytest.mark.skipif(not GOLDENS, reason="no MuJoCo goldens under tests/golden/mujoco (MuJoCo is not installed here; "
                                        "see tools/mujoco_crosscheck.py) - parity with MuJoCo proper stays unpinned")
@pytest.mark.parametrize("path", GOLDENS, ids=[p.name for p in GOLDENS])
def test_mujoco_goldens(path):
    g = json.loads(path.read_text())
    assert g["source"] == "mujoco", "only dumps made by MuJoCo belong in tests/golden/mujoco"
    rep = xc.check(g, ROOT / "fixtures")
    assert rep["ok"], rep
