"""tests/golden/flat/*.npz (the scenes the oracles read, oracle/flatmodel.py) are what the product's MJCF loader produces
from fixtures/*.xml today, and the MuJoCo stand-in of oracle/_ref sees the same scene as O1."""
import numpy as np
import pytest

from oracle.flatmodel import FLAT_DIR, FlatModel, _DBL, _INT
from sspp_b200 import capi
from tests.helpers import FIX


@pytest.mark.parametrize("stem", ["robocrane7", "robocrane", "planner", "stacking", "armlab"])
def test_flat_model_files_match_the_loader(stem):
    f = FlatModel.load(stem)
    m = FlatModel.from_capi(capi.Model(path=str(FIX / f"{stem}.xml")))
    for k in _INT + _DBL:
        np.testing.assert_array_equal(f.a[k], m.a[k], err_msg=k)
    assert f.body_names == m.body_names and f.geom_names == m.geom_names
    assert (FLAT_DIR / f"{stem}.npz").stat().st_size < 64 * 1024


def test_mujoco_stand_in_sees_the_scene_o1_sees():
    from oracle import pyoracle, pyref
    if not pyref.available():
        pytest.skip("oracle/_ref/libref.so is not built")
    fm = FlatModel.load("robocrane7")
    o1 = pyoracle.O1(fm, 0, 7)
    rm = pyref.RefModel(fm)
    rng = np.random.default_rng(0)
    for q in rng.uniform(-2, 2, (300, 7)):
        full = fm.a["qpos0"].copy(); full[:7] = q
        assert rm.ncon(full) == o1.point_cost(q)[1]
