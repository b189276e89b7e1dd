"""Flat FP64 scene for the checker oracles, read from a committed .npz (tests/golden/flat/*.npz).

TEST INFRASTRUCTURE.  The oracles (O1, the _ref stand-in for MuJoCo) consume a scene as plain arrays; this class
serves them from a file so that the CPU legs of bench.py and the `-m "not gpu"` oracle tests do not have to map the
product library just to parse MJCF.  The files are written by tests/golden/make_flat_models.py (which does use the
product's loader, once) and tests/test_flat_models.py checks that the loader still reproduces them."""
from __future__ import annotations

import ctypes as C
from pathlib import Path

import numpy as np

_ip, _dp = C.POINTER(C.c_int32), C.POINTER(C.c_double)
FLAT_DIR = Path(__file__).resolve().parent.parent / "tests" / "golden" / "flat"

_INT = ("body_parent", "body_weld", "body_jtype", "body_qadr", "geom_body", "geom_type", "geom_contype", "geom_conaffinity", "exclude")
_DBL = ("body_pos", "body_quat", "jnt_axis", "jnt_pos", "jnt_range", "qpos0", "geom_size", "geom_pos", "geom_quat", "geom_margin")


class FlatView(C.Structure):       # layout of SsppFlatModelView (include/sspp_cuda.h) == FlatView of o1_oracle.cpp
    _fields_ = [("nbody", C.c_int32), ("ngeom", C.c_int32), ("nq", C.c_int32), ("nexclude", C.c_int32),
                ("body_parent", _ip), ("body_weld", _ip), ("body_jtype", _ip), ("body_qadr", _ip),
                ("body_pos", _dp), ("body_quat", _dp), ("jnt_axis", _dp), ("jnt_pos", _dp), ("jnt_range", _dp),
                ("qpos0", _dp),
                ("geom_body", _ip), ("geom_type", _ip), ("geom_contype", _ip), ("geom_conaffinity", _ip),
                ("geom_size", _dp), ("geom_pos", _dp), ("geom_quat", _dp), ("geom_margin", _dp),
                ("exclude", _ip)]


class FlatModel:
    """same duck type as sspp_b200.capi.Model for what the oracles touch: .view, .body_id(), .body_name(), .geom_name()"""

    def __init__(self, arrays: dict, body_names, geom_names):
        self.a = {}
        for k in _INT:
            self.a[k] = np.ascontiguousarray(arrays[k], np.int32)
        for k in _DBL:
            self.a[k] = np.ascontiguousarray(arrays[k], np.float64)
        self.body_names = [str(s) for s in body_names]
        self.geom_names = [str(s) for s in geom_names]
        v = FlatView()
        v.nbody, v.ngeom, v.nq = len(self.a["body_parent"]), len(self.a["geom_body"]), len(self.a["qpos0"])
        v.nexclude = self.a["exclude"].size // 2
        for k in _INT:
            setattr(v, k, self.a[k].ctypes.data_as(_ip))
        for k in _DBL:
            setattr(v, k, self.a[k].ctypes.data_as(_dp))
        self.view = v

    @classmethod
    def load(cls, name: str) -> "FlatModel":
        """name: fixture stem ('robocrane7') or a path to an .npz"""
        p = Path(name)
        if p.suffix != ".npz":
            p = FLAT_DIR / f"{p.stem}.npz"
        z = np.load(p, allow_pickle=False)
        return cls({k: z[k] for k in _INT + _DBL}, z["body_names"], z["geom_names"])

    @classmethod
    def from_capi(cls, model) -> "FlatModel":
        """copy of a parsed sspp_b200.capi.Model (used by the generator script and by tests that build scenes on the fly)"""
        a = model.arrays()
        return cls(a, [model.body_name(b) for b in range(a["nbody"])], [model.geom_name(g) for g in range(a["ngeom"])])

    def save(self, path) -> None:
        np.savez_compressed(path, body_names=np.array(self.body_names), geom_names=np.array(self.geom_names), **self.a)

    def body_id(self, name: str) -> int:
        return self.body_names.index(name) if name in self.body_names else -1

    def body_name(self, b: int) -> str:
        return self.body_names[b]

    def geom_name(self, g: int) -> str:
        return self.geom_names[g]
