"""ctypes wrapper of oracle/_ref/libref.so — the reference's own hot-path headers compiled unmodified against the
third-party stand-ins of oracle/shim/ (see oracle/ref_driver.cpp).

TEST INFRASTRUCTURE — only tests/ and bench.py's CPU legs import this module.  The library is built in the development
container (where /root/reference exists) by `make -C oracle`; on the GPU box the prebuilt file is used as shipped.
"""
from __future__ import annotations

import ctypes as C
import subprocess
from pathlib import Path

import numpy as np

_HERE = Path(__file__).resolve().parent
LIB_PATH = _HERE / "_ref" / "libref.so"
_dp = C.POINTER(C.c_double)
MODE_JOINT, MODE_TASK = 0, 1


def available() -> bool:
    if not LIB_PATH.exists() and Path("/root/reference/include/sspp.h").exists():
        subprocess.run(["make", "-C", str(_HERE)], check=False, capture_output=True)
    return LIB_PATH.exists()


def _d(a):
    return a.ctypes.data_as(_dp)


class DistCfg(C.Structure):
    _fields_ = [("sigma_floor", C.c_double), ("var_beta", C.c_double), ("mean_lr", C.c_double), ("stddev_min", C.c_double),
                ("stddev_max", C.c_double), ("z_min", C.c_double), ("lo", C.c_double * 4), ("hi", C.c_double * 4)]


class TspArgs(C.Structure):
    """arguments of tsp::TaskSpacePlanner's constructor (include/sspp/tsp.h:12-33), in its order and with its defaults"""
    _fields_ = [("stddev_initial", C.c_double), ("stddev_min", C.c_double), ("stddev_max", C.c_double), ("inc", C.c_double),
                ("dec", C.c_double), ("elite_fraction", C.c_double), ("sample_count", C.c_int), ("check_points", C.c_int),
                ("init_points", C.c_int), ("collision_weight", C.c_double), ("z_min", C.c_double), ("lo", C.c_double * 4),
                ("hi", C.c_double * 4), ("sigma_floor", C.c_double), ("var_ema_beta", C.c_double), ("mean_lr", C.c_double),
                ("floor_margin", C.c_double), ("floor_penalty_scale", C.c_double)]


_lib = None


def lib() -> C.CDLL:
    global _lib
    if _lib is None:
        if not available():
            raise FileNotFoundError(f"{LIB_PATH} is missing and cannot be built here (no /root/reference)")
        L = C.CDLL(str(LIB_PATH))
        L.ref_model_create.restype = C.c_void_p
        L.ref_model_create.argtypes = [C.c_void_p, C.POINTER(C.c_char_p), C.POINTER(C.c_char_p)]
        L.ref_model_destroy.argtypes = [C.c_void_p]
        L.ref_model_ncon.argtypes = [C.c_void_p, _dp]
        L.ref_tape_arm.argtypes = [C.c_int, C.c_uint64, C.c_uint32, C.c_uint32, C.c_uint32, C.c_int, C.c_int]
        L.ref_tape_counts.argtypes = [C.POINTER(C.c_long)] * 3
        L.ref_elites.argtypes = [_dp, C.c_int, C.c_double, C.c_void_p, _dp]
        L.ref_dist_reset.argtypes = [C.POINTER(DistCfg), C.c_int, _dp, C.c_double, _dp, _dp]
        L.ref_dist_update.argtypes = [C.POINTER(DistCfg), C.c_int, C.c_int, _dp, C.c_int, C.c_void_p, _dp, C.c_double, C.c_double, _dp, _dp]
        L.ref_wrap_angle_diff.restype = C.c_double
        L.ref_wrap_angle_diff.argtypes = [C.c_double] * 4
        L.ref_sample_sets.argtypes = [_dp, _dp, C.c_int, _dp, _dp, C.c_double, C.c_int, _dp]
        L.ref_floor_penalty.restype = C.c_double
        L.ref_floor_penalty.argtypes = [C.c_double, C.c_double, C.c_double, _dp]
        L.ref_from_vias.argtypes = [_dp, _dp, C.c_int, _dp, C.c_int, _dp, _dp, _dp, _dp]
        L.ref_point_costs.argtypes = [C.c_void_p, C.c_char_p, _dp, C.c_int, C.c_int, _dp]
        L.ref_check_points.argtypes = [C.c_void_p, C.c_char_p, _dp, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, _dp]
        L.ref_tsp_create.restype = C.c_void_p
        L.ref_tsp_create.argtypes = [C.c_void_p, C.c_char_p, C.POINTER(TspArgs)]
        L.ref_tsp_destroy.argtypes = [C.c_void_p]
        L.ref_tsp_plan.argtypes = [C.c_void_p, _dp, _dp, C.c_int, C.c_int, _dp, _dp, _dp, _dp, C.c_void_p, _dp, _dp, C.c_int, _dp, _dp]
        L.ref_tsp_run.restype = C.c_long
        L.ref_tsp_run.argtypes = [C.c_void_p, _dp, _dp, C.c_int, C.c_int]
        L.ref_sspp_create.restype = C.c_void_p
        L.ref_sspp_create.argtypes = [C.c_void_p, C.c_int]
        L.ref_sspp_destroy.argtypes = [C.c_void_p]
        L.ref_sspp_init_path.argtypes = [C.c_void_p, _dp, _dp, C.c_int, _dp, _dp]
        L.ref_sspp_candidates.argtypes = [C.c_void_p, _dp, _dp, C.c_double, _dp, C.c_int, C.c_int, C.c_int, _dp, C.c_void_p, _dp]
        L.ref_sspp_plan.argtypes = [C.c_void_p, _dp, _dp, C.c_double, _dp, C.c_int, C.c_int, C.c_int, C.c_int, _dp, _dp, C.POINTER(C.c_int)]
        L.ref_set_threads.argtypes = [C.c_int]
        _lib = L
    return _lib


def set_threads(n: int) -> None:
    lib().ref_set_threads(n)


def max_threads() -> int:
    return int(lib().ref_max_threads())


def arm_tape(mode: int, seed: int, it: int, query: int = 0, cand0: int = 0, streams_per_row: int = 4, rows: int = 1) -> None:
    """the Philox counters the next std::normal_distribution constructions map to (oracle/shim/random)"""
    lib().ref_tape_arm(mode, C.c_uint64(seed), it, query, cand0, streams_per_row, rows)


def tape_counts():
    a, b, c = C.c_long(), C.c_long(), C.c_long()
    lib().ref_tape_counts(C.byref(a), C.byref(b), C.byref(c))
    return dict(streams=a.value, normals=b.value, uniforms=c.value)


class RefModel:
    """mjModel of the MuJoCo stand-in, built from a flat model (oracle.flatmodel.FlatModel or sspp_b200.capi.Model)"""

    def __init__(self, flat):
        self.flat = flat
        nb, ng = flat.view.nbody, flat.view.ngeom
        self._bn = (C.c_char_p * nb)(*[flat.body_name(b).encode() for b in range(nb)])
        self._gn = (C.c_char_p * ng)(*[flat.geom_name(g).encode() for g in range(ng)])
        self.h = lib().ref_model_create(C.addressof(flat.view), self._bn, self._gn)
        self.nq = flat.view.nq

    def __del__(self):
        if getattr(self, "h", None) and _lib is not None:
            _lib.ref_model_destroy(self.h)
            self.h = None

    def ncon(self, qpos_full) -> int:
        q = np.ascontiguousarray(qpos_full, np.float64)
        return int(lib().ref_model_ncon(self.h, _d(q)))

    def point_costs(self, body: str, pts, use_center: bool = True) -> np.ndarray:
        """Collision<Point>::collision_point_cost (Collision.h:84-103) at task-space points [n][4]"""
        pts = np.ascontiguousarray(pts, np.float64).reshape(-1, 4)
        out = np.zeros(len(pts))
        lib().ref_point_costs(self.h, body.encode(), _d(pts), len(pts), int(use_center), _d(out))
        return out


    def check_points(self, body: str, pts):
        """Collision<Point>::check_collision_point (Collision.h:51-82) at task-space points [n][4] -> (hit, geom1, geom2, dist):
        the function's return value and the ids / distance it prints for the first penetrating contact"""
        pts = np.ascontiguousarray(pts, np.float64).reshape(-1, 4)
        n = len(pts)
        hit = np.zeros(n, np.int32); g1 = np.zeros(n, np.int32); g2 = np.zeros(n, np.int32); dist = np.zeros(n)
        v = C.c_void_p
        lib().ref_check_points(self.h, body.encode(), _d(pts), n, hit.ctypes.data_as(v), g1.ctypes.data_as(v), g2.ctypes.data_as(v), _d(dist))
        return hit.astype(bool), g1, g2, dist


# ---- CES pieces (tsp_elites.h, tsp_distribution.h, tsp_sampler.h, tsp_evaluator.h, tsp_path_model.h) ----
def elites(cost, frac: float):
    """EliteSelector::select + weights over the successes' total costs -> (indices into cost, weights)"""
    cost = np.ascontiguousarray(cost, np.float64)
    idx = np.zeros(max(len(cost), 1), np.int64); w = np.zeros(max(len(cost), 1))
    k = lib().ref_elites(_d(cost), len(cost), frac, idx.ctypes.data_as(C.c_void_p), _d(w))
    return idx[:k], w[:k]


def dist_cfg(sigma_floor, var_beta, mean_lr, stddev_min, stddev_max, z_min, lo, hi) -> DistCfg:
    c = DistCfg(sigma_floor, var_beta, mean_lr, stddev_min, stddev_max, z_min)
    for d in range(4):
        c.lo[d], c.hi[d] = lo[d], hi[d]
    return c


def dist_reset(cfg: DistCfg, mean0, s0: float):
    mean0 = np.ascontiguousarray(mean0, np.float64).reshape(-1, 4)
    m, s = np.zeros_like(mean0), np.zeros_like(mean0)
    lib().ref_dist_reset(C.byref(cfg), len(mean0), _d(mean0), s0, _d(m), _d(s))
    return m, s


def dist_update(cfg: DistCfg, mean, sigma, vias, elite_idx, w, inc: float, dec: float):
    """Distribution::update + adapt as Planner::plan orders them (tsp_planner.h:122-141); no elites: adapt(false)"""
    mean = np.ascontiguousarray(mean, np.float64).reshape(-1, 4).copy()
    sigma = np.ascontiguousarray(sigma, np.float64).reshape(-1, 4).copy()
    Kv = len(mean)
    vias = np.ascontiguousarray(vias, np.float64).reshape(-1, Kv, 4) if len(elite_idx) else np.zeros((1, Kv, 4))
    ei = np.ascontiguousarray(elite_idx, np.int64); w = np.ascontiguousarray(w, np.float64)
    lib().ref_dist_update(C.byref(cfg), Kv, len(vias), _d(vias), len(ei), ei.ctypes.data_as(C.c_void_p), _d(w), inc, dec, _d(mean), _d(sigma))
    return mean, sigma


def wrap_angle_diff(a, b, lo, hi) -> float:
    return lib().ref_wrap_angle_diff(a, b, lo, hi)


def sample_sets(lo, hi, mean, sigma, z_min: float, n: int) -> np.ndarray:
    """Sampler::sample_set n times in sequence (arm the tape first) -> [n][Kv][4]"""
    lo = np.ascontiguousarray(lo, np.float64); hi = np.ascontiguousarray(hi, np.float64)
    mean = np.ascontiguousarray(mean, np.float64).reshape(-1, 4); sigma = np.ascontiguousarray(sigma, np.float64).reshape(-1, 4)
    out = np.zeros((n, len(mean), 4))
    lib().ref_sample_sets(_d(lo), _d(hi), len(mean), _d(mean), _d(sigma), z_min, n, _d(out))
    return out


def floor_penalty(z_min, margin, scale, p) -> float:
    p = np.ascontiguousarray(p, np.float64)
    return lib().ref_floor_penalty(z_min, margin, scale, _d(p))


def from_vias(start, end, total_points: int, vias, u):
    """PathModel::initLinear + fromVias -> (ctrl [total][4], knots, points at u)"""
    start = np.ascontiguousarray(start, np.float64); end = np.ascontiguousarray(end, np.float64)
    vias = np.ascontiguousarray(vias, np.float64).reshape(-1, 4); u = np.ascontiguousarray(u, np.float64)
    assert len(vias) == total_points - 2
    ctrl = np.zeros((total_points, 4)); knots = np.zeros(total_points + 2 + 1); pts = np.zeros((len(u), 4))
    lib().ref_from_vias(_d(start), _d(end), total_points, _d(vias), len(u), _d(u), _d(ctrl), _d(knots), _d(pts))
    return ctrl, knots, pts


class RefTsp:
    """tsp::TaskSpacePlanner (include/sspp/tsp.h over tsp_planner.h) on the MuJoCo stand-in"""

    def __init__(self, model: RefModel, body: str, args: TspArgs):
        self.model, self.args = model, args
        self.h = lib().ref_tsp_create(model.h, body.encode(), C.byref(args))
        self.Kv = args.init_points - 2

    def __del__(self):
        if getattr(self, "h", None) and _lib is not None:
            _lib.ref_tsp_destroy(self.h)
            self.h = None

    def plan(self, start, end, iterate: bool, u=None):
        """one plan() call with ONE OpenMP thread (deterministic order).  Returns a dict with per-seed arrays in seed order"""
        set_threads(1)
        start = np.ascontiguousarray(start, np.float64); end = np.ascontiguousarray(end, np.float64)
        cap = self.args.sample_count + 2
        vias = np.zeros((cap, max(self.Kv, 1), 4)); L = np.zeros(cap); Cnf = np.zeros(cap); Cwf = np.zeros(cap)
        st = np.zeros(cap, np.int32); m0 = np.zeros(4); s0 = np.zeros(4)
        u = np.ascontiguousarray(np.linspace(0, 1, 11) if u is None else u, np.float64)
        path = np.zeros((len(u), 4))
        n = lib().ref_tsp_plan(self.h, _d(start), _d(end), int(iterate), cap, _d(vias), _d(L), _d(Cnf), _d(Cwf),
                               st.ctypes.data_as(C.c_void_p), _d(m0), _d(s0), len(u), _d(u), _d(path))
        if n < 0:
            raise RuntimeError(f"ref_tsp_plan failed ({n})")
        return dict(n=n, vias=vias[:n, :self.Kv], L=L[:n], Cnf=Cnf[:n], Cwf=Cwf[:n], status=st[:n], mean0=m0, sigma0=s0, path=path)

    def run(self, start, end, iters: int, cold: bool = True) -> int:
        start = np.ascontiguousarray(start, np.float64); end = np.ascontiguousarray(end, np.float64)
        return int(lib().ref_tsp_run(self.h, _d(start), _d(end), iters, int(cold)))


class RefSspp:
    """sspp::SamplingPathPlanner<dof> (include/sspp.h) on the MuJoCo stand-in; dof in {3, 6, 7, 9} (sspp_bindings.cpp:58-61)"""

    def __init__(self, model: RefModel, dof: int):
        self.model, self.dof = model, dof
        self.h = lib().ref_sspp_create(model.h, dof)
        if not self.h:
            raise ValueError(f"the reference binds no SamplingPathPlanner{dof}")

    def __del__(self):
        if getattr(self, "h", None) and _lib is not None:
            _lib.ref_sspp_destroy(self.h)
            self.h = None

    def init_path(self, start, end, K: int):
        start = np.ascontiguousarray(start, np.float64); end = np.ascontiguousarray(end, np.float64)
        ctrl = np.zeros((K, self.dof)); knots = np.zeros(K + 3 + 1)
        lib().ref_sspp_init_path(self.h, _d(start), _d(end), K, _d(ctrl), _d(knots))
        return ctrl, knots

    def candidates(self, start, end, sigma: float, limits, n: int, M: int, K: int):
        """sampleWithNoise + checkCollision + computeArcLength per candidate, in candidate order (one thread; arm the tape first)"""
        set_threads(1)
        start = np.ascontiguousarray(start, np.float64); end = np.ascontiguousarray(end, np.float64); limits = np.ascontiguousarray(limits, np.float64)
        ctrl = np.zeros((n, K, self.dof)); col = np.zeros(n, np.int32); arc = np.zeros(n)
        lib().ref_sspp_candidates(self.h, _d(start), _d(end), sigma, _d(limits), n, M, K, _d(ctrl), col.ctypes.data_as(C.c_void_p), _d(arc))
        return dict(ctrl=ctrl, collides=col.astype(bool), arc=arc)

    def plan(self, start, end, sigma: float, limits, n: int, M: int, K: int, threads: int = 1):
        """plan() itself (sspp.h:194-225): successes' control points in arrival order, the best path's control points"""
        set_threads(threads)
        start = np.ascontiguousarray(start, np.float64); end = np.ascontiguousarray(end, np.float64); limits = np.ascontiguousarray(limits, np.float64)
        succ = np.zeros((n, K, self.dof)); best = np.zeros((K, self.dof)); found = C.c_int()
        ns = lib().ref_sspp_plan(self.h, _d(start), _d(end), sigma, _d(limits), n, M, K, n, _d(succ), _d(best), C.byref(found))
        return dict(n_success=ns, succ_ctrl=succ[:ns], best_ctrl=best, found=bool(found.value))
