"""sspp/BSplines.py and sspp/CubicPath.py of the reference as product functions: same names, argument
order and conventions, over the C ABI (include/sspp_cuda.h) - host FP64 for the scalar algebra, the GPU for batches.

  B, dB, bspline, bspline_derivative   sspp/BSplines.py:11-55   (theta may be a scalar or an array here)
  knot_vector, compute_control_points  sspp/BSplines.py:58-106
  CubicPath                            sspp/CubicPath.py:8-60
  bspline_batch                        additive: many curves x many parameters in one kernel launch

The casadi twins and the SLERP helpers of BSplines.py (:114-208) serve the reference's IK / symbolic scripts, not the
scoring path, and are out of scope."""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import capi

_lib = capi.lib
_dp = C.POINTER(C.c_double)


def _d(a):
    return a.ctypes.data_as(_dp)


def _rows(theta, t, k: int, kind: int) -> np.ndarray:
    t = np.ascontiguousarray(t, np.float64)
    th = np.ascontiguousarray(np.atleast_1d(theta), np.float64)
    n = len(t) - k - 1
    if n < 1:
        raise ValueError("knot vector too short for the order")
    out = np.zeros((len(th), n))
    rc = _lib.ssppcu_bspline_rows(n, k, _d(t), _d(th), len(th), kind, _d(out))
    if rc != capi.OK:
        raise capi.SsppError(rc, "ssppcu_bspline_rows")
    return out


def B(theta, k, i, t):
    """basis function i of order k at theta (BSplines.py:11-29); array theta -> array"""
    r = _rows(theta, t, k, 0)[:, i]
    return float(r[0]) if np.ndim(theta) == 0 else r


def dB(theta, k, i, t):
    r = _rows(theta, t, k, 1)[:, i]
    return float(r[0]) if np.ndim(theta) == 0 else r


def _ordered_sum(rows: np.ndarray, c: np.ndarray) -> np.ndarray:
    """sum_i rows[:, i] * c[i] accumulated in ascending i from zero - the order of the reference's `sum(c[i] * B(...) for i in
    range(n))` (BSplines.py:51,55), so that the result carries the same bits (a BLAS product may associate differently)"""
    acc = np.zeros((rows.shape[0],) + c.shape[1:])
    for i in range(c.shape[0]):
        acc = acc + rows[:, i].reshape((-1,) + (1,) * (c.ndim - 1)) * c[i]
    return acc


def bspline(theta, t, c, k):
    """curve value(s) at theta (BSplines.py:44-51): theta < 0 -> c[0] B_0(0), theta >= 1 -> c[n-1]"""
    c = np.asarray(c, np.float64)
    n = len(t) - k - 1
    v = _ordered_sum(_rows(theta, t, k, 2), c[:n])
    return v[0] if np.ndim(theta) == 0 else v


def bspline_derivative(theta, t, c, k):
    c = np.asarray(c, np.float64)
    n = len(t) - k - 1
    v = _ordered_sum(_rows(theta, t, k, 3), c[:n])
    return v[0] if np.ndim(theta) == 0 else v


def knot_vector(n_control_points: int, k: int) -> np.ndarray:
    t = np.zeros(n_control_points + k + 1)
    rc = _lib.ssppcu_bspline_knot_vector(n_control_points, k, _d(t))
    if rc != capi.OK:
        raise ValueError("knot_vector: need at least k + 1 control points")
    return t


def compute_control_points(via_points, k: int):
    """(control_points [n][d], knot vector) of the order-k B-spline through the via points at theta_i = i/(n-1) (BSplines.py:65-106)"""
    v = np.ascontiguousarray(via_points, np.float64)
    one_d = v.ndim == 1
    v2 = v.reshape(len(v), -1)
    n, d = v2.shape
    ctrl = np.zeros((n, d)); t = np.zeros(n + k + 1)
    rc = _lib.ssppcu_bspline_fit(_d(v2), n, d, k, _d(ctrl), _d(t))
    if rc != capi.OK:
        raise ValueError("compute_control_points: need [n >= max(2, k + 1)][d] via points")
    return (ctrl[:, 0] if one_d else ctrl), t


def bspline_batch(theta, t, c_batch, k: int, derivative: bool = False, device: int = 0) -> np.ndarray:
    """[n_curves][n_theta][d] for control points c_batch [n_curves][n][d]: one GPU launch, FP32"""
    cb = np.ascontiguousarray(c_batch, np.float32)
    if cb.ndim != 3:
        raise ValueError("c_batch must be [n_curves][n][d]")
    n = len(t) - k - 1
    rows = np.ascontiguousarray(_rows(theta, t, k, 3 if derivative else 2), np.float32)
    cb = np.ascontiguousarray(cb[:, :n])
    out = np.zeros((cb.shape[0], rows.shape[0], cb.shape[2]), np.float32)
    err = C.create_string_buffer(512)
    f = C.POINTER(C.c_float)
    rc = _lib.ssppcu_bspline_eval_batch(device, rows.ctypes.data_as(f), rows.shape[0], n, cb.ctypes.data_as(f), cb.shape[0], cb.shape[2],
                                        out.ctypes.data_as(f), err, len(err))
    if rc != capi.OK:
        raise capi.SsppError(rc, err.value.decode())
    return out


def bspline_batch_noalloc(theta, t, c_batch, k: int, derivative: bool = False, device: int = 0) -> np.ndarray:
    """bspline_batch through ssppcu_bspline_eval_batch_noalloc: same output bit for bit, no device allocation (the staging is
    a 4 MiB file-scope device array; larger batches raise SsppError with code ELIMIT).  No counterpart in the reference."""
    cb = np.ascontiguousarray(c_batch, np.float32)
    if cb.ndim != 3:
        raise ValueError("c_batch must be [n_curves][n][d]")
    n = len(t) - k - 1
    rows = np.ascontiguousarray(_rows(theta, t, k, 3 if derivative else 2), np.float32)
    cb = np.ascontiguousarray(cb[:, :n])
    out = np.zeros((cb.shape[0], rows.shape[0], cb.shape[2]), np.float32)
    err = C.create_string_buffer(512)
    f = C.POINTER(C.c_float)
    rc = _lib.ssppcu_bspline_eval_batch_noalloc(device, rows.ctypes.data_as(f), rows.shape[0], n, cb.ctypes.data_as(f), cb.shape[0],
                                                cb.shape[2], out.ctypes.data_as(f), err, len(err))
    if rc != capi.OK:
        raise capi.SsppError(rc, err.value.decode())
    return out


class CubicPath:
    """closed-form cubic through start, via (at u = 1/2) and end with zero start velocity (CubicPath.py:8-60)"""

    def __init__(self):
        self.coefficients = None
        self._pts = None

    def plan(self, start, via, end):
        s, v, e = (np.ascontiguousarray(x, np.float64).ravel() for x in (start, via, end))
        if not (len(s) == len(v) == len(e)):
            raise ValueError("start, via and end must have the same dimension")
        self._pts = (s, v, e)
        self.a = 2 * (e + 3 * s - 4 * v)
        self.b = 4 * (v - s - self.a / 8)
        self.c = 0
        self.d = s
        return True

    def _eval(self, u, want):
        if self._pts is None:
            raise RuntimeError("CubicPath.plan() has not been called")
        s, v, e = self._pts
        uu = np.ascontiguousarray(np.atleast_1d(u), np.float64)
        outs = [np.zeros((len(uu), len(s))) if w else None for w in want]
        rc = _lib.ssppcu_cubic_path(_d(s), _d(v), _d(e), len(s), _d(uu), len(uu), *[None if o is None else _d(o) for o in outs])
        if rc != capi.OK:
            raise capi.SsppError(rc, "ssppcu_cubic_path")
        return [None if o is None else (o[0] if np.ndim(u) == 0 else o) for o in outs]

    def evaluate(self, u):
        return self._eval(u, (True, False, False))[0]

    def evaluate_with_derivatives(self, u):
        return tuple(self._eval(u, (True, True, True)))
