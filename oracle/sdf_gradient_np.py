"""CPU restatement (numpy, FP64) of the signed-distance refinement step.  TEST INFRASTRUCTURE ONLY: imported by
tests/ (and nothing in the product path); parity for this step is UNPINNED against MuJoCo (the reference computes the
witness direction with mj_geomDistance, which is not available here) and anchored on the reference's own formula.

Reference: /root/reference/scripts/sdf_gradient/main.py
  get_closest_points       :37-63  min over (collision geom x environment geom) of the signed distance and
                                   grad = min_dist * unit(fromto[3:6] - fromto[0:3])          (:60-61)
  get_collision_gradient   :66-81  col_grad_i = (1/len(u_list)) * sum_u grad(u) * B_i(u)
On the baked grid the pair minimum is the minimum over the moving body's sample spheres of field - radius, and the
witness direction (from the moving geom towards the obstacle) is minus the gradient of the trilinear interpolant.
Planner B pose rule (include/utility.h:149-175, 198-206): q = (x, y, z, yaw), R = Rz(yaw).
"""
from __future__ import annotations

import numpy as np


def locate(dims, lo, cell, x):
    """cell index and fractions of world point x (points outside read the boundary cell) - sspp_math.h sdf_locate"""
    n = np.asarray(dims)
    g = np.clip((np.asarray(x, float) - lo) / cell, 0.0, n - 1.0)
    i = np.minimum(g.astype(int), n - 2)
    return i, g - i


def field_and_gradient(nodes, lo, cell, x):
    """trilinear value and its exact gradient at x; nodes [nz][ny][nx]"""
    nz, ny, nx = nodes.shape
    (i, j, k), (fx, fy, fz) = locate((nx, ny, nz), lo, cell, x)
    v = nodes[k:k + 2, j:j + 2, i:i + 2].astype(float)          # [z][y][x]
    wx, wy, wz = np.array([1 - fx, fx]), np.array([1 - fy, fy]), np.array([1 - fz, fz])
    val = np.einsum("zyx,z,y,x->", v, wz, wy, wx)
    d = np.array([-1.0, 1.0])
    g = np.array([np.einsum("zyx,z,y,x->", v, wz, wy, d) / cell[0],
                  np.einsum("zyx,z,y,x->", v, wz, d, wx) / cell[1],
                  np.einsum("zyx,z,y,x->", v, d, wy, wx) / cell[2]])
    return val, g


def closest_sample(nodes, lo, cell, spheres, q):
    """(clearance, field gradient) of the sample sphere closest to the scene at q = (x, y, z, yaw); spheres [n][5]"""
    c, s = np.cos(q[3]), np.sin(q[3])
    R = np.array([[c, -s, 0.0], [s, c, 0.0], [0.0, 0.0, 1.0]])
    best, best_g = np.inf, np.zeros(3)
    for sp in np.asarray(spheres, float):
        assert sp[4] == 0, "restated for a single planned body"
        val, g = field_and_gradient(nodes, lo, cell, np.asarray(q[:3], float) + R @ sp[:3])
        if val - sp[3] < best:
            best, best_g = val - sp[3], g
    return best, best_g


def sdf_gradient(nodes, lo, cell, spheres, pts, W):
    """pts [K][4] point matrix (or control points), W [rows][K] collocation (or basis) rows ->
    (grad [K][3], clearance [rows], dirs [rows][3])"""
    pts, W = np.asarray(pts, float), np.asarray(W, float)
    rows = W.shape[0]
    dist, dirs = np.zeros(rows), np.zeros((rows, 3))
    for m in range(rows):
        q = W[m] @ pts
        cl, g = closest_sample(nodes, lo, cell, spheres, q)
        n = np.linalg.norm(g)
        dist[m] = cl
        dirs[m] = -cl * g / n if n > 0 else 0.0
    return (dirs.T @ W).T / rows, dist, dirs
