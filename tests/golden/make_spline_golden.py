"""Generate tests/golden/spline_golden.npz by IMPORTING the reference's Python spline twins.

Run here (needs /root/reference):  python tests/golden/make_spline_golden.py
The reference modules sspp/BSplines.py and sspp/CubicPath.py are imported unmodified; casadi and
matplotlib (absent, and only used by the reference's symbolic/plot helpers) are stubbed in
sys.modules.  Outputs of the reference functions on seeded inputs are stored next to the inputs,
together with the reference's only stored artefact scripts/bspline_params.npy (KAT #1).
"""
from __future__ import annotations

import sys
import types
from pathlib import Path

import numpy as np

REF = Path("/root/reference")
OUT = Path(__file__).resolve().parent / "spline_golden.npz"


def import_reference():
    for name in ("casadi", "matplotlib", "matplotlib.pyplot"):
        sys.modules.setdefault(name, types.ModuleType(name))
    sys.path.insert(0, str(REF / "sspp"))
    import BSplines as bs       # noqa
    import CubicPath as cp      # noqa
    return bs, cp


def main():
    bs, cp = import_reference()
    rng = np.random.default_rng(20261104)
    g = {}
    kat = np.load(REF / "scripts" / "bspline_params.npy", allow_pickle=True).item()
    g["kat_knot_vec"] = np.asarray(kat["knot_vec"], float)
    g["kat_ctr_pts"] = np.asarray(kat["ctr_pts"], float)
    g["kat_k"] = np.array(int(kat["k"]))
    shapes = [(7, 9, 2), (10, 7, 3), (5, 4, 2), (4, 3, 3), (3, 4, 2), (6, 2, 1)]
    g["shapes"] = np.array(shapes)
    thetas = np.concatenate(([-0.25, 0.0, 1.0, 1.5], np.linspace(0, 1, 41), rng.uniform(0, 1, 20)))
    g["thetas"] = thetas
    for s, (n, d, k) in enumerate(shapes):
        via = rng.normal(size=(n, d))
        ctrl, t = bs.compute_control_points(via, k)
        g[f"s{s}_via"] = via
        g[f"s{s}_ctrl"] = np.asarray(ctrl)
        g[f"s{s}_knots"] = np.asarray(t, float)
        g[f"s{s}_knot_vector"] = np.asarray(bs.knot_vector(n, k), float)
        g[f"s{s}_curve"] = np.array([bs.bspline(th, t, ctrl, k) for th in thetas])
        g[f"s{s}_dcurve"] = np.array([bs.bspline_derivative(th, t, ctrl, k) for th in thetas])
        g[f"s{s}_B"] = np.array([[bs.B(th, k, i, t) for i in range(n)] for th in thetas])
        g[f"s{s}_dB"] = np.array([[bs.dB(th, k, i, t) for i in range(n)] for th in thetas])
    # the reference's own unit-test cases (sspp/tests/test_BSplines.py:73-94)
    t1 = bs.knot_vector(7, 1)
    g["ramp_curve"] = np.array([bs.bspline(th, t1, np.arange(7.0), 1) for th in np.linspace(0, 1, 25)])
    t3 = bs.knot_vector(6, 3)
    g["const_curve"] = np.array([bs.bspline(th, t3, np.ones(6), 3) for th in np.linspace(0, 0.999, 25)])
    # CubicPath (sspp/CubicPath.py:8-43)
    trip = rng.normal(size=(4, 3, 7))
    us = np.array([-0.5, 0.0, 0.1, 0.25, 0.5, 0.75, 0.9, 1.0, 1.7])
    g["cubic_triples"] = trip
    g["cubic_us"] = us
    pos = np.zeros((4, len(us), 7)); vel = np.zeros_like(pos); acc = np.zeros_like(pos)
    for i, (a, b, c) in enumerate(trip):
        p = cp.CubicPath(); p.plan(a, b, c)
        for j, u in enumerate(us):
            pos[i, j] = p.evaluate(u)
            _, vel[i, j], acc[i, j] = p.evaluate_with_derivatives(u)
    g["cubic_pos"], g["cubic_vel"], g["cubic_acc"] = pos, vel, acc
    np.savez_compressed(OUT, **g)
    print("wrote", OUT, OUT.stat().st_size, "bytes,", len(g), "arrays")


if __name__ == "__main__":
    main()
