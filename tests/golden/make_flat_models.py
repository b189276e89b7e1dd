"""Writes tests/golden/flat/<fixture>.npz: the flat FP64 arrays of every fixture scene, as parsed by the product's MJCF
loader (sspp_b200/csrc/sspp_model.cpp).  The oracles read these files (oracle/flatmodel.py); tests/test_flat_models.py
checks that the loader still reproduces them.  Run from the repo root:  python tests/golden/make_flat_models.py"""
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parents[2]
sys.path.insert(0, str(ROOT))
from oracle.flatmodel import FLAT_DIR, FlatModel      # noqa: E402
from sspp_b200 import capi                            # noqa: E402
from tests.helpers import sdf_case                    # noqa: E402

if __name__ == "__main__":
    sdf_case()                                        # (re)writes fixtures/_cfg3_clutter64.xml from its seeded generator
    FLAT_DIR.mkdir(parents=True, exist_ok=True)
    for xml in sorted((ROOT / "fixtures").glob("*.xml")):
        FlatModel.from_capi(capi.Model(path=str(xml))).save(FLAT_DIR / f"{xml.stem}.npz")
        print("wrote", xml.stem)
