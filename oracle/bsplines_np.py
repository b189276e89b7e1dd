"""TEST INFRASTRUCTURE, NOT PRODUCT CODE.  numpy restatement of the reference's Python spline twins.

Rows P1/P2 of SURVEY 8(a) (paths under /root/reference):
  basis(), basis_deriv()      sspp/BSplines.py:11-42   Cox-de Boor with HALF-OPEN spans (sums to 0 at theta = 1)
  curve(), curve_deriv()      sspp/BSplines.py:44-55   theta < 0 -> c[0]*B(0), theta >= 1 -> c[n-1]
  uniform_knots()             sspp/BSplines.py:58-62   clamped uniform knot vector
  fit_control_points()        sspp/BSplines.py:65-106  collocation at i/(n-1), unit first/last rows, least squares
  cubic_plan(), cubic_eval()  sspp/CubicPath.py:8-60   closed-form cubic through start / via(1/2) / end

Pinned: tests/test_oracle_spline.py checks every function against tests/golden/spline_golden.npz,
which tests/golden/make_spline_golden.py produced by importing the reference modules themselves,
and against the reference's stored artefact scripts/bspline_params.npy (KAT #1).

Written as a bottom-up table (degree 0 row, then raise the degree) instead of the reference's
top-down recursion; the arithmetic per entry is the same two-term Cox-de Boor blend.
"""
from __future__ import annotations

import numpy as np


def _degree_table(theta: float, k: int, t: np.ndarray) -> list[np.ndarray]:
    """rows[j][i] = B_{i,j}(theta) for j = 0..k, i = 0..len(t)-j-2"""
    t = np.asarray(t, float)
    rows = [np.array([1.0 if (t[i] <= theta < t[i + 1]) else 0.0 for i in range(len(t) - 1)])]
    for j in range(1, k + 1):
        prev = rows[-1]
        cur = np.zeros(len(t) - j - 1)
        for i in range(len(cur)):
            left = 0.0 if t[i + j] == t[i] else (theta - t[i]) / (t[i + j] - t[i]) * prev[i]
            right = 0.0 if t[i + j + 1] == t[i + 1] else (t[i + j + 1] - theta) / (t[i + j + 1] - t[i + 1]) * prev[i + 1]
            cur[i] = left + right
        rows.append(cur)
    return rows


def basis(theta: float, k: int, t) -> np.ndarray:
    """all n = len(t)-k-1 basis values of degree k at theta"""
    return _degree_table(theta, k, t)[k]


def basis_deriv(theta: float, k: int, t) -> np.ndarray:
    t = np.asarray(t, float)
    n = len(t) - k - 1
    if k == 0:
        return np.zeros(n)
    low = _degree_table(theta, k - 1, t)[k - 1]
    out = np.zeros(n)
    for i in range(n):
        a = 0.0 if t[i + k] == t[i] else k / (t[i + k] - t[i]) * low[i]
        b = 0.0 if t[i + k + 1] == t[i + 1] else -k / (t[i + k + 1] - t[i + 1]) * low[i + 1]
        out[i] = a + b
    return out


def curve(theta: float, t, c, k: int) -> np.ndarray:
    c = np.asarray(c, float)
    n = len(t) - k - 1
    if theta < 0:
        return c[0] * basis(0.0, k, t)[0]
    if theta >= 1:
        return c[n - 1]
    b = basis(theta, k, t)
    out = np.zeros_like(c[0])
    for i in range(n):          # same left-to-right accumulation as the reference's sum()
        out = out + c[i] * b[i]
    return out


def curve_deriv(theta: float, t, c, k: int) -> np.ndarray:
    c = np.asarray(c, float)
    n = len(t) - k - 1
    b = basis_deriv(theta, k, t)
    out = np.zeros_like(c[0])
    for i in range(n):
        out = out + c[i] * b[i]
    return out


def uniform_knots(n_ctrl: int, k: int) -> np.ndarray:
    inner = np.linspace(0, 1, n_ctrl + k + 1 - 2 * k)
    return np.concatenate((np.zeros(k), inner, np.ones(k)))


def fit_matrix(n: int, k: int) -> np.ndarray:
    t = uniform_knots(n, k)
    A = np.array([basis(i / (n - 1), k, t) for i in range(n)])
    A[0, 0] = 1.0
    A[n - 1, n - 1] = 1.0
    return A


def fit_control_points(via, k: int):
    via = np.asarray(via, float)
    n = len(via)
    ctrl = np.linalg.lstsq(fit_matrix(n, k), via, rcond=None)[0]
    return ctrl, uniform_knots(n, k)


def cubic_plan(start, via, end):
    start, via, end = (np.asarray(x, float) for x in (start, via, end))
    a = 2 * (end + 3 * start - 4 * via)
    b = 4 * (via - start - a / 8)
    return a, b, 0.0, start


def cubic_eval(coef, u: float):
    a, b, c, d = coef
    u = float(np.clip(u, 0, 1))
    return a * u**3 + b * u**2 + c * u + d, 3 * a * u**2 + 2 * b * u + c, 6 * a * u + 2 * b
