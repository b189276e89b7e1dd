"""ctypes wrappers of the two checker libraries (oracle/_build/libo1.so, libo2.so).

TEST INFRASTRUCTURE — only tests/, __graft_entry__.smoke() and bench.py (cpu_baseline and
--impl reference legs) import this module.  Nothing under sspp_b200/ does.
"""
from __future__ import annotations

import ctypes as C
import subprocess
from pathlib import Path

import numpy as np

_HERE = Path(__file__).resolve().parent
_BUILD = _HERE / "_build"


def build(force: bool = False) -> None:
    if force or not (_BUILD / "libo1.so").exists() or not (_BUILD / "libo2.so").exists():
        subprocess.run(["make", "-C", str(_HERE)] + (["-B"] if force else []), check=True, capture_output=True)


def _lib(name: str) -> C.CDLL:
    p = _BUILD / name
    if not p.exists():
        build()
    return C.CDLL(str(p))


_dp = C.POINTER(C.c_double)


class O1Cfg(C.Structure):
    _fields_ = [("mode", C.c_int), ("D", C.c_int), ("K", C.c_int), ("p", C.c_int), ("M", C.c_int),
                ("n_free", C.c_int), ("free_begin", C.c_int),
                ("w_collision", C.c_double), ("eval_zmin", C.c_double), ("floor_margin", C.c_double),
                ("floor_scale", C.c_double), ("limit_scale", C.c_double), ("sample_zmin", C.c_double),
                ("dist_zmin", C.c_double), ("elite_frac", C.c_double), ("inc", C.c_double), ("dec", C.c_double),
                ("sigma_min", C.c_double), ("sigma_max", C.c_double), ("sigma_floor", C.c_double),
                ("var_beta", C.c_double), ("mean_lr", C.c_double), ("lo", C.c_double * 9), ("hi", C.c_double * 9)]


def o1_cfg_from(pc) -> O1Cfg:
    """O1 configuration from a sspp_b200.capi.PlannerConfig (field-by-field, no product code involved)."""
    c = O1Cfg()
    task = pc.mode == 1
    c.mode, c.D, c.K, c.p, c.M = pc.mode, pc.dof, pc.n_points, pc.degree, pc.check_points
    c.n_free = pc.n_points - 2 if task else pc.n_points - 2 * pc.degree
    c.free_begin = 1 if task else pc.degree
    for f in ("w_collision", "eval_zmin", "floor_margin", "floor_scale", "limit_scale", "sample_zmin", "dist_zmin",
              "inc", "dec", "sigma_floor", "var_beta", "mean_lr"):
        setattr(c, f, getattr(pc, f))
    c.elite_frac, c.sigma_min, c.sigma_max = pc.elite_fraction, pc.stddev_min, pc.stddev_max
    for d in range(9):
        c.lo[d], c.hi[d] = pc.lo[d], pc.hi[d]
    return c


def _d(a):
    return a.ctypes.data_as(_dp)


class O1:
    """FP64 restatement of the reference semantics over a flat model view."""

    def __init__(self, model, mode: int, dof: int, body: str = ""):
        self.lib = _lib("libo1.so")
        L = self.lib
        L.o1_create.restype = C.c_void_p
        L.o1_create.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int]
        L.o1_point_cost.restype = C.c_double
        L.o1_point_cost.argtypes = [C.c_void_p, _dp, C.POINTER(C.c_int32), _dp]
        L.o1_pair_dist.restype = C.c_double
        L.o1_pair_dist.argtypes = [C.c_void_p, _dp, C.c_int, C.c_int]
        L.o1_geom_frames.argtypes = [C.c_void_p, _dp, _dp, _dp]
        L.o1_num_pairs.argtypes = [C.c_void_p]
        L.o1_get_pairs.argtypes = [C.c_void_p, C.c_void_p]
        L.o1_destroy.argtypes = [C.c_void_p]
        L.o1_score.argtypes = [C.c_void_p, C.POINTER(O1Cfg), _dp, _dp, C.c_int, _dp, _dp, _dp, _dp, C.c_void_p, _dp, _dp, C.c_void_p]
        L.o1_elites.argtypes = [_dp, C.c_void_p, C.c_int, C.c_double, C.c_void_p, _dp, C.POINTER(C.c_int32)]
        L.o1_update.argtypes = [C.POINTER(O1Cfg), _dp, _dp, _dp, _dp, C.c_int]
        L.o1_sample.argtypes = [C.POINTER(O1Cfg), C.c_uint64, C.c_uint32, C.c_uint32, C.c_int, C.c_int, _dp, _dp, _dp]
        L.o1_normal.restype = C.c_double
        L.o1_normal.argtypes = [C.c_uint64, C.c_uint32, C.c_uint32, C.c_uint32, C.c_uint32, C.c_uint32]
        L.o1_philox.argtypes = [C.c_void_p, C.c_uint32, C.c_uint32]
        L.o1_knots.argtypes = [C.c_int, C.c_int, _dp]
        L.o1_interpolate.argtypes = [_dp, C.c_int, C.c_int, C.c_int, _dp]
        L.o1_spline_eval.argtypes = [_dp, C.c_int, C.c_int, C.c_int, C.c_double, _dp]
        L.o1_spline_eval_knots.argtypes = [_dp, C.c_int, C.c_int, C.c_int, _dp, C.c_double, _dp]
        L.o1_basis_row.argtypes = [C.c_int, C.c_int, _dp, C.c_double, _dp]
        L.o1_set_sdf.argtypes = [C.c_void_p, C.c_void_p, _dp, _dp, C.c_void_p, C.c_int, C.c_double]
        L.o1_get_sdf.argtypes = [C.c_void_p, _dp]
        self.model = model
        self.mode, self.dof = mode, dof
        plan_body = model.body_id(body) if mode == 1 else -1
        if mode == 1 and plan_body < 0:
            raise ValueError(f"body '{body}' not found")
        self.h = L.o1_create(C.addressof(model.view), mode, dof, plan_body)
        self.ngeom = model.view.ngeom

    def __del__(self):
        if getattr(self, "h", None):
            self.lib.o1_destroy(self.h)
            self.h = None

    def set_sdf(self, dims, lo, hi, spheres, margin: float = 0.0):
        """switch to the SDF collision mode: FP64 grid of the static partners, sample spheres as an input"""
        d = np.ascontiguousarray(dims, np.int32); lo = np.ascontiguousarray(lo, np.float64); hi = np.ascontiguousarray(hi, np.float64)
        sp = np.ascontiguousarray(spheres, np.float32)
        self._sdf_dims = tuple(int(x) for x in d)
        self.lib.o1_set_sdf(self.h, d.ctypes.data_as(C.c_void_p), _d(lo), _d(hi), sp.ctypes.data_as(C.c_void_p), len(sp), float(margin))

    def sdf_nodes(self) -> np.ndarray:
        nx, ny, nz = self._sdf_dims
        out = np.zeros((nz, ny, nx))
        self.lib.o1_get_sdf(self.h, _d(out))
        return out

    def pairs(self) -> np.ndarray:
        n = self.lib.o1_num_pairs(self.h)
        out = np.zeros((n, 2), np.int32)
        self.lib.o1_get_pairs(self.h, out.ctypes.data_as(C.c_void_p))
        return out

    def point_cost(self, q):
        q = np.ascontiguousarray(q, np.float64)
        nc, md = C.c_int32(), C.c_double()
        c = self.lib.o1_point_cost(self.h, _d(q), C.byref(nc), C.byref(md))
        return c, nc.value, md.value

    def pair_dist(self, q, g1: int, g2: int) -> float:
        q = np.ascontiguousarray(q, np.float64)
        return self.lib.o1_pair_dist(self.h, _d(q), g1, g2)

    def geom_frames(self, q):
        q = np.ascontiguousarray(q, np.float64)
        xpos = np.zeros((self.ngeom, 3)); xmat = np.zeros((self.ngeom, 3, 3))
        self.lib.o1_geom_frames(self.h, _d(q), _d(xpos), _d(xmat))
        return xpos, xmat

    def score(self, cfg: O1Cfg, base, params, want_points: bool = False):
        base = np.ascontiguousarray(base, np.float64)
        params = np.ascontiguousarray(params, np.float64).reshape(-1, cfg.n_free * cfg.D)
        n = len(params)
        L = np.zeros(n); Cnf = np.zeros(n); Cwf = np.zeros(n); cost = np.zeros(n); md = np.zeros(n)
        st = np.zeros(n, np.int32)
        pts = np.zeros((n, cfg.M + 1, cfg.D)) if want_points else None
        band = np.zeros(n, np.int32)      # decisions within EPS_BAND of a threshold, per candidate
        self.lib.o1_score(self.h, C.byref(cfg), _d(base), _d(params), n, _d(L), _d(Cnf), _d(Cwf), _d(cost),
                          st.ctypes.data_as(C.c_void_p), _d(md), None if pts is None else _d(pts), band.ctypes.data_as(C.c_void_p))
        return dict(L=L, Cnf=Cnf, Cwf=Cwf, cost=cost, status=st, min_dist=md, pts=pts, band=band)

    def elites(self, cost, status, frac: float):
        cost = np.ascontiguousarray(cost, np.float64); status = np.ascontiguousarray(status, np.int32)
        n = len(cost)
        idx = np.zeros(n, np.int32); w = np.zeros(n); ns = C.c_int32()
        k = self.lib.o1_elites(_d(cost), status.ctypes.data_as(C.c_void_p), n, frac, idx.ctypes.data_as(C.c_void_p), _d(w), C.byref(ns))
        return idx[:k], w[:k], ns.value

    def update(self, cfg: O1Cfg, mean, sigma, elite_params, w):
        mean = np.ascontiguousarray(mean, np.float64).copy().ravel()
        sigma = np.ascontiguousarray(sigma, np.float64).copy().ravel()
        ep = np.ascontiguousarray(elite_params, np.float64).reshape(len(w), -1) if len(w) else np.zeros((0, len(mean)))
        w = np.ascontiguousarray(w, np.float64)
        self.lib.o1_update(C.byref(cfg), _d(mean), _d(sigma), _d(ep), _d(w), len(w))
        return mean, sigma

    def sample(self, cfg: O1Cfg, seed: int, it: int, query: int, cand0: int, n: int, mean, sigma):
        mean = np.ascontiguousarray(mean, np.float64).ravel(); sigma = np.ascontiguousarray(sigma, np.float64).ravel()
        out = np.zeros((n, cfg.n_free * cfg.D))
        self.lib.o1_sample(C.byref(cfg), C.c_uint64(seed), it, query, cand0, n, _d(mean), _d(sigma), _d(out))
        return out

    # spline pieces
    def knots(self, K: int, p: int) -> np.ndarray:
        t = np.zeros(K + p + 1)
        self.lib.o1_knots(K, p, _d(t))
        return t

    def interpolate(self, V, p: int) -> np.ndarray:
        V = np.ascontiguousarray(V, np.float64)
        K, D = V.shape
        out = np.zeros((K, D))
        self.lib.o1_interpolate(_d(V), K, D, p, _d(out))
        return out

    def spline_eval(self, ctrl, p: int, u: float) -> np.ndarray:
        ctrl = np.ascontiguousarray(ctrl, np.float64)
        K, D = ctrl.shape
        out = np.zeros(D)
        self.lib.o1_spline_eval(_d(ctrl), K, D, p, float(u), _d(out))
        return out

    def spline_eval_knots(self, ctrl, p: int, knots, u: float) -> np.ndarray:
        ctrl = np.ascontiguousarray(ctrl, np.float64); knots = np.ascontiguousarray(knots, np.float64)
        K, D = ctrl.shape
        assert len(knots) == K + p + 1
        out = np.zeros(D)
        self.lib.o1_spline_eval_knots(_d(ctrl), K, D, p, _d(knots), float(u), _d(out))
        return out

    def basis_row(self, K: int, p: int, knots, u: float) -> np.ndarray:
        knots = np.ascontiguousarray(knots, np.float64)
        out = np.zeros(K)
        self.lib.o1_basis_row(K, p, _d(knots), float(u), _d(out))
        return out

    def philox(self, ctr, k0: int, k1: int) -> np.ndarray:
        c = np.ascontiguousarray(ctr, np.uint32).copy()
        self.lib.o1_philox(c.ctypes.data_as(C.c_void_p), k0, k1)
        return c

    def normal(self, seed, it, query, cand, stream, j) -> float:
        return self.lib.o1_normal(C.c_uint64(seed), it, query, cand, stream, j)


class O2:
    """FP32 bit replica of one fused iteration (product arithmetic header compiled for the host)."""

    def __init__(self, mjcf_path: str, planner_cfg):
        self.lib = _lib("libo2.so")
        L = self.lib
        L.o2_create.restype = C.c_void_p
        L.o2_create.argtypes = [C.c_char_p, C.c_void_p, C.c_char_p, C.c_size_t]
        L.o2_destroy.argtypes = [C.c_void_p]
        L.o2_n_cand.argtypes = [C.c_void_p]
        L.o2_point_cost.restype = C.c_float
        L.o2_point_cost.argtypes = [C.c_void_p, C.c_void_p]
        L.o2_normal.restype = C.c_float
        L.o2_normal.argtypes = [C.c_uint64, C.c_uint32, C.c_uint32, C.c_uint32, C.c_uint32, C.c_uint32]
        L.o2_log.restype = C.c_float
        L.o2_log.argtypes = [C.c_float]
        L.o2_sincos.argtypes = [C.c_float, C.POINTER(C.c_float), C.POINTER(C.c_float)]
        L.o2_philox.argtypes = [C.c_void_p, C.c_uint32, C.c_uint32]
        L.o2_iterate.argtypes = [C.c_void_p] + [C.c_void_p] * 7 + [C.c_uint64, C.c_uint32, C.c_uint32, C.c_int, C.c_int, C.c_int] + [C.c_void_p] * 11
        L.o2_sdf_nodes.restype = C.POINTER(C.c_float)
        L.o2_sdf_nodes.argtypes = [C.c_void_p, C.POINTER(C.c_int32 * 3)]
        L.o2_sdf_sample.restype = C.c_float
        L.o2_sdf_sample.argtypes = [C.c_void_p, C.c_void_p]
        L.o2_merge_update.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_uint64, C.c_uint32] + [C.c_void_p] * 7
        err = C.create_string_buffer(1024)
        self.cfg = planner_cfg
        self.h = L.o2_create(str(mjcf_path).encode(), C.addressof(planner_cfg), err, len(err))
        if not self.h:
            raise RuntimeError(err.value.decode())
        task = planner_cfg.mode == 1
        self.D, self.K, self.M = planner_cfg.dof, planner_cfg.n_points, planner_cfg.check_points
        self.n_free = self.K - 2 if task else self.K - 2 * planner_cfg.degree
        self.nfd = self.n_free * self.D
        self.n_cand = L.o2_n_cand(self.h)

    def __del__(self):
        if getattr(self, "h", None):
            self.lib.o2_destroy(self.h)
            self.h = None

    def point_cost(self, q) -> float:
        q = np.ascontiguousarray(q, np.float32)
        return self.lib.o2_point_cost(self.h, q.ctypes.data_as(C.c_void_p))

    def sdf_nodes(self) -> np.ndarray:
        dims = (C.c_int32 * 3)()
        p = self.lib.o2_sdf_nodes(self.h, C.byref(dims))
        return np.ctypeslib.as_array(p, shape=(dims[2], dims[1], dims[0])).copy()

    def spheres(self) -> np.ndarray:
        self.lib.o2_spheres.argtypes = [C.c_void_p, C.c_void_p, C.c_int]
        n = self.lib.o2_spheres(self.h, None, 0)
        out = np.zeros((max(n, 1), 5), np.float32)
        self.lib.o2_spheres(self.h, out.ctypes.data_as(C.c_void_p), n)
        return out[:n]

    def sdf_sample(self, x) -> float:
        x = np.ascontiguousarray(x, np.float32)
        return self.lib.o2_sdf_sample(self.h, x.ctypes.data_as(C.c_void_p))

    def normal(self, seed, it, query, cand, stream, j) -> float:
        return self.lib.o2_normal(C.c_uint64(seed), it, query, cand, stream, j)

    def log(self, x: float) -> float:
        return self.lib.o2_log(C.c_float(x))

    def sincos(self, x: float):
        s, c = C.c_float(), C.c_float()
        self.lib.o2_sincos(C.c_float(x), C.byref(s), C.byref(c))
        return s.value, c.value

    def philox(self, ctr, k0: int, k1: int) -> np.ndarray:
        c = np.ascontiguousarray(ctr, np.uint32).copy()
        self.lib.o2_philox(c.ctypes.data_as(C.c_void_p), k0, k1)
        return c

    def iterate(self, psi_col, psi_arc, base, mean, sigma, best, best_valid: int, seed: int, it: int, query: int = 0,
                cand_offset: int = 0, n_local: int | None = None, do_elites: bool = True, want_points: bool = False):
        """One iteration of one query.  Returns a dict; mean/sigma/best in it are the updated state."""
        n = self.n_cand if n_local is None else n_local
        f = lambda a: np.ascontiguousarray(a, np.float32).copy()
        psi_col, base, mean, sigma, best = f(psi_col), f(base), f(mean).ravel(), f(sigma).ravel(), f(best).ravel()
        psi_arc = f(psi_arc) if psi_arc is not None else np.zeros(1, np.float32)
        bv = np.array([best_valid], np.int32)
        out = dict(params=np.zeros((n, self.n_free, self.D), np.float32), L=np.zeros(n, np.float32),
                   Cnf=np.zeros(n, np.float32), Cwf=np.zeros(n, np.float32), cost=np.zeros(n, np.float32),
                   status=np.zeros(n, np.uint8), keys=np.zeros(n, np.uint64), weights=np.zeros(max(n, 1), np.float32))
        ns, bi = np.zeros(1, np.int32), np.zeros(1, np.int32)
        pts = np.zeros((n, self.M + 1, self.D), np.float32) if want_points else None
        p = lambda a: None if a is None else a.ctypes.data_as(C.c_void_p)
        k = self.lib.o2_iterate(self.h, p(psi_col), p(psi_arc), p(base), p(mean), p(sigma), p(best), p(bv),
                                C.c_uint64(seed), it, query, cand_offset, n, int(do_elites),
                                p(out["params"]), p(out["L"]), p(out["Cnf"]), p(out["Cwf"]), p(out["cost"]),
                                p(out["status"]), p(out["keys"]), p(out["weights"]), p(ns), p(bi), p(pts))
        out.update(n_elite=k, n_success=int(ns[0]), best_idx=int(bi[0]), mean=mean.reshape(self.n_free, self.D),
                   sigma=sigma.reshape(self.n_free, self.D), best=best.reshape(self.n_free, self.D),
                   best_valid=int(bv[0]), pts=pts)
        out["weights"] = out["weights"][:k]
        return out

    def merge_update(self, keys, seed: int, it: int, mean, sigma, best, best_valid: int):
        """population-sharded K6/K7 over the gathered keys of all shards (restates the merge kernels)"""
        f = lambda a: np.ascontiguousarray(a, np.float32).copy().ravel()
        keys = np.ascontiguousarray(keys, np.uint64).copy().ravel()
        mean, sigma, best = f(mean), f(sigma), f(best)
        bv = np.array([best_valid], np.int32); ns = np.zeros(1, np.int32); bi = np.zeros(1, np.int32)
        w = np.zeros(max(len(keys), 1), np.float32)
        p = lambda a: a.ctypes.data_as(C.c_void_p)
        k = self.lib.o2_merge_update(self.h, p(keys), len(keys), C.c_uint64(seed), it, p(mean), p(sigma), p(best), p(bv), p(w), p(ns), p(bi))
        return dict(keys=keys, n_elite=k, weights=w[:k], n_success=int(ns[0]), best_idx=int(bi[0]),
                    mean=mean.reshape(self.n_free, self.D), sigma=sigma.reshape(self.n_free, self.D),
                    best=best.reshape(self.n_free, self.D), best_valid=int(bv[0]))
