"""ctypes binding of the C ABI in include/sspp_cuda.h (libsspp_cuda.so).

This is the thinnest possible host-side layer: it only marshals arguments.  The shared library
is built in-tree by ``__graft_entry__.build()`` (nvcc, sm_100a).  If it is missing the import
fails loudly — there is no Python or CPU fallback for the scoring path.
"""
from __future__ import annotations

import ctypes as C
import os
from pathlib import Path

import numpy as np

_HERE = Path(__file__).resolve().parent
LIB_PATH = Path(os.environ.get("SSPP_CUDA_LIB", _HERE / "csrc" / "libsspp_cuda.so"))   # override: kernel experiments only

MAX_DOF = 9
MODE_JOINT, MODE_TASK = 0, 1
COLLIDE_PRIMITIVES, COLLIDE_SDF = 0, 1
ST_CONVERGED, ST_FAILED, ST_ABSENT = 0, 1, 2
OK, EINVAL, EMODEL, ECUDA, ELIMIT = 0, -1, -2, -3, -4


class SsppError(RuntimeError):
    def __init__(self, code: int, msg: str):
        super().__init__(msg)
        self.code = code


def _load() -> C.CDLL:
    if not LIB_PATH.exists():
        raise ImportError(
            f"{LIB_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
            "(nvcc, sm_100a). sspp_b200 has no CPU fallback.")
    return C.CDLL(str(LIB_PATH), mode=os.RTLD_GLOBAL if hasattr(os, "RTLD_GLOBAL") else 0)


_dp = C.POINTER(C.c_double)
_ip = C.POINTER(C.c_int32)
_fp = C.POINTER(C.c_float)


class FlatModelView(C.Structure):
    _fields_ = [("nbody", C.c_int32), ("ngeom", C.c_int32), ("nq", C.c_int32), ("nexclude", C.c_int32),
                ("body_parent", _ip), ("body_weld", _ip), ("body_jtype", _ip), ("body_qadr", _ip),
                ("body_pos", _dp), ("body_quat", _dp), ("jnt_axis", _dp), ("jnt_pos", _dp), ("jnt_range", _dp),
                ("qpos0", _dp),
                ("geom_body", _ip), ("geom_type", _ip), ("geom_contype", _ip), ("geom_conaffinity", _ip),
                ("geom_size", _dp), ("geom_pos", _dp), ("geom_quat", _dp), ("geom_margin", _dp),
                ("exclude", _ip)]


class PlannerConfig(C.Structure):
    _fields_ = [("mode", C.c_int32), ("dof", C.c_int32), ("degree", C.c_int32), ("n_points", C.c_int32),
                ("check_points", C.c_int32), ("n_samples", C.c_int32), ("include_mean", C.c_int32),
                ("include_best", C.c_int32), ("do_update", C.c_int32), ("n_queries", C.c_int32),
                ("device", C.c_int32), ("shard_offset", C.c_int32), ("shard_global", C.c_int32),
                ("body_name", C.c_char_p),
                ("w_collision", C.c_double), ("elite_fraction", C.c_double), ("inc", C.c_double), ("dec", C.c_double),
                ("sigma_floor", C.c_double), ("var_beta", C.c_double), ("mean_lr", C.c_double),
                ("stddev_min", C.c_double), ("stddev_max", C.c_double),
                ("eval_zmin", C.c_double), ("floor_margin", C.c_double), ("floor_scale", C.c_double),
                ("limit_scale", C.c_double), ("sample_zmin", C.c_double), ("dist_zmin", C.c_double),
                ("lo", C.c_double * MAX_DOF), ("hi", C.c_double * MAX_DOF),
                ("collision_mode", C.c_int32), ("use_center_dist", C.c_int32), ("sdf_dims", C.c_int32 * 3), ("tile_hint", C.c_int32),
                ("sdf_lo", C.c_double * 3), ("sdf_hi", C.c_double * 3), ("sdf_sphere_ratio", C.c_double)]


class Dims(C.Structure):
    _fields_ = [(n, C.c_int32) for n in ("n_queries", "n_cand", "n_free", "dof", "n_points", "nfd", "n_pad",
                                          "rows_collision", "rows_arc", "first_free_row", "threads", "cand_per_cta",
                                          "ctas_per_query")] + [("smem_bytes", C.c_int64)] + \
               [(n, C.c_int32) for n in ("n_mbody", "n_mgeom", "n_sgeom", "n_pair", "n_mmpair", "chain", "variant", "grid")]


class DeviceView(C.Structure):
    _fields_ = [(n, C.c_void_p) for n in ("L", "Cnf", "Cwf", "cost", "status", "params", "mean", "sigma", "best",
                                          "best_valid", "keys", "elite_w", "n_success", "n_elite", "best_idx",
                                          "best_cost", "any_success", "stream")]


# every symbol include/sspp_cuda.h declares (tests check the library exports all of them)
SYMBOLS = [
    "ssppcu_model_load", "ssppcu_model_parse", "ssppcu_model_free", "ssppcu_model_view", "ssppcu_model_body_id",
    "ssppcu_model_body_name", "ssppcu_model_geom_name", "ssppcu_model_body_point", "ssppcu_model_pair_counts",
    "ssppcu_default_config", "ssppcu_create", "ssppcu_destroy", "ssppcu_last_error", "ssppcu_dims",
    "ssppcu_set_stream", "ssppcu_set_queries", "ssppcu_reset_distribution", "ssppcu_set_distribution",
    "ssppcu_get_distribution", "ssppcu_set_joint_noise", "ssppcu_iterate", "ssppcu_sync", "ssppcu_get_results",
    "ssppcu_get_params", "ssppcu_get_best", "ssppcu_get_elites", "ssppcu_device_view", "ssppcu_eval_points",
    "ssppcu_get_collocation", "ssppcu_get_base", "ssppcu_shard_keys", "ssppcu_merge_update", "ssppcu_version",
    "ssppcu_spline_setup", "ssppcu_collocation_matrix", "ssppcu_save_state", "ssppcu_restore_state",
    "ssppcu_fp32_peak", "ssppcu_get_merged", "ssppcu_get_sdf", "ssppcu_get_spheres", "ssppcu_phase_probe", "ssppcu_iterate_restored",
    "ssppcu_exchange_export", "ssppcu_exchange_attach_ipc", "ssppcu_exchange_attach_ptrs", "ssppcu_iterate_exchange",
    "ssppcu_exchange_status", "ssppcu_sdf_gradient", "ssppcu_set_result_mirror", "ssppcu_work_counters",
    "ssppcu_plan_step", "ssppcu_set_success_rows", "ssppcu_get_success_rows", "ssppcu_candidate_ctrl",
    "ssppcu_bspline_knot_vector", "ssppcu_bspline_rows", "ssppcu_bspline_fit", "ssppcu_cubic_path", "ssppcu_bspline_eval_batch",
    "ssppcu_bspline_eval_batch_noalloc", "ssppcu_check_points",
]
WORK_COUNTERS = ("points", "fk_hinge_axis", "fk_hinge_gen", "fk_slide", "fk_free4", "fk_r0", "fk_anchor", "geom_pos", "geom_rot_box",
                 "geom_rot_axis", "broad", "body_broad", "np_plane", "np_sphere_box", "np_box_cull", "np_caps_box_early", "np_caps_box", "np_caps_newton",
                 "np_box_box_face", "np_box_box", "np_cyl_box_face", "np_cyl_box", "np_other", "contacts", "mm_tests", "mm_narrow",
                 "sdf_samples")

lib = _load()
lib.ssppcu_last_error.restype = C.c_char_p
lib.ssppcu_model_body_name.restype = C.c_char_p
lib.ssppcu_model_geom_name.restype = C.c_char_p
lib.ssppcu_model_load.argtypes = [C.c_char_p, C.POINTER(C.c_void_p), C.c_char_p, C.c_size_t]
lib.ssppcu_model_parse.argtypes = [C.c_char_p, C.POINTER(C.c_void_p), C.c_char_p, C.c_size_t]
lib.ssppcu_model_free.argtypes = [C.c_void_p]
lib.ssppcu_model_view.argtypes = [C.c_void_p, C.POINTER(FlatModelView)]
lib.ssppcu_model_body_id.argtypes = [C.c_void_p, C.c_char_p]
lib.ssppcu_model_body_name.argtypes = [C.c_void_p, C.c_int]
lib.ssppcu_model_geom_name.argtypes = [C.c_void_p, C.c_int]
lib.ssppcu_model_body_point.argtypes = [C.c_void_p, C.c_char_p, _dp]
lib.ssppcu_model_pair_counts.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_char_p, _ip, _ip, _ip, C.c_char_p, C.c_size_t]
lib.ssppcu_default_config.argtypes = [C.c_int, C.POINTER(PlannerConfig)]
lib.ssppcu_default_config.restype = None
lib.ssppcu_create.argtypes = [C.c_void_p, C.POINTER(PlannerConfig), C.POINTER(C.c_void_p), C.c_char_p, C.c_size_t]
lib.ssppcu_destroy.argtypes = [C.c_void_p]
lib.ssppcu_destroy.restype = None
lib.ssppcu_model_free.restype = None
lib.ssppcu_last_error.argtypes = [C.c_void_p]
lib.ssppcu_dims.argtypes = [C.c_void_p, C.POINTER(Dims)]
lib.ssppcu_set_stream.argtypes = [C.c_void_p, C.c_void_p]
lib.ssppcu_set_result_mirror.argtypes = [C.c_void_p, C.c_int]
lib.ssppcu_set_queries.argtypes = [C.c_void_p, _dp, _dp]
lib.ssppcu_reset_distribution.argtypes = [C.c_void_p, C.c_double]
lib.ssppcu_set_distribution.argtypes = [C.c_void_p, _dp, _dp]
lib.ssppcu_get_distribution.argtypes = [C.c_void_p, _dp, _dp]
lib.ssppcu_set_joint_noise.argtypes = [C.c_void_p, C.c_double, _dp]
lib.ssppcu_iterate.argtypes = [C.c_void_p, C.c_uint64, C.c_uint32, C.c_int]
lib.ssppcu_iterate_restored.argtypes = [C.c_void_p, C.c_uint64, C.c_uint32, C.c_int]
lib.ssppcu_sync.argtypes = [C.c_void_p]
lib.ssppcu_get_results.argtypes = [C.c_void_p] + [C.c_void_p] * 10
lib.ssppcu_get_params.argtypes = [C.c_void_p, C.c_void_p]
lib.ssppcu_get_best.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p]
lib.ssppcu_plan_step.argtypes = [C.c_void_p, _dp, _dp, C.c_double, _dp, C.c_uint64, C.c_uint32, C.c_int]
lib.ssppcu_get_elites.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int]
lib.ssppcu_device_view.argtypes = [C.c_void_p, C.POINTER(DeviceView)]
lib.ssppcu_eval_points.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p]
lib.ssppcu_get_collocation.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p]
lib.ssppcu_get_base.argtypes = [C.c_void_p, C.c_void_p]
lib.ssppcu_shard_keys.argtypes = [C.c_void_p, C.POINTER(C.c_void_p), _ip]
lib.ssppcu_merge_update.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_uint64, C.c_uint32]
lib.ssppcu_get_merged.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int]
lib.ssppcu_exchange_export.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_void_p]
lib.ssppcu_exchange_attach_ipc.argtypes = [C.c_void_p, C.c_void_p]
lib.ssppcu_exchange_attach_ptrs.argtypes = [C.c_void_p, C.POINTER(C.c_void_p)]
lib.ssppcu_iterate_exchange.argtypes = [C.c_void_p, C.c_uint64, C.c_uint32]
lib.ssppcu_exchange_status.argtypes = [C.c_void_p]
lib.ssppcu_get_sdf.argtypes = [C.c_void_p, C.c_void_p, _ip, _dp, _dp]
lib.ssppcu_get_spheres.argtypes = [C.c_void_p, C.c_void_p, C.c_int]
lib.ssppcu_sdf_gradient.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p]
lib.ssppcu_save_state.argtypes = [C.c_void_p]
lib.ssppcu_restore_state.argtypes = [C.c_void_p]
lib.ssppcu_fp32_peak.argtypes = [C.c_int, _dp, _dp]
lib.ssppcu_spline_setup.argtypes = [C.c_int, C.c_int, _dp, _dp]
lib.ssppcu_collocation_matrix.argtypes = [C.c_int, C.c_int, _dp, C.c_int, C.c_int, _dp]


def _ptr(a: np.ndarray | None):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


def _dptr(a: np.ndarray):
    return a.ctypes.data_as(_dp)


def default_config(mode: int) -> PlannerConfig:
    cfg = PlannerConfig()
    lib.ssppcu_default_config(mode, C.byref(cfg))
    return cfg


def fp32_peak(device: int = 0) -> tuple[float, float]:
    """(measured FP32 FMA TFLOP/s, SM clock MHz from the device properties)"""
    t, c = C.c_double(), C.c_double()
    rc = lib.ssppcu_fp32_peak(device, C.byref(t), C.byref(c))
    if rc != OK:
        raise SsppError(rc, "FP32 probe failed (no CUDA device?)")
    return t.value, c.value


def spline_setup(n_points: int, degree: int):
    """(knots, Ainv) of the interpolating spline through n_points uniform parameters (host FP64)."""
    knots = np.zeros(n_points + degree + 1); ainv = np.zeros((n_points, n_points))
    rc = lib.ssppcu_spline_setup(n_points, degree, _dptr(knots), _dptr(ainv))
    if rc != OK:
        raise SsppError(rc, "invalid spline shape")
    return knots, ainv


def collocation_matrix(n_points: int, degree: int, grid, via_form: bool) -> np.ndarray:
    g = np.ascontiguousarray(grid, np.float64)
    psi = np.zeros((len(g), n_points))
    rc = lib.ssppcu_collocation_matrix(n_points, degree, _dptr(g), len(g), int(via_form), _dptr(psi))
    if rc != OK:
        raise SsppError(rc, "invalid spline shape")
    return psi


def planner_grids(cfg: "PlannerConfig"):
    """host restatement of the grids ssppcu_create builds: (psi_col, psi_arc) as float32"""
    M, task = cfg.check_points, cfg.mode == MODE_TASK
    if task:
        gcol = np.arange(M + 1) * (1.0 / M)
        return collocation_matrix(cfg.n_points, cfg.degree, gcol, True).astype(np.float32), None
    gcol = np.arange(M + 1) / M
    garc = np.arange(M) / (M - 1)
    return (collocation_matrix(cfg.n_points, cfg.degree, gcol, False).astype(np.float32),
            collocation_matrix(cfg.n_points, cfg.degree, garc, False).astype(np.float32))


def base_points(cfg: "PlannerConfig", start, end) -> np.ndarray:
    """point matrix [K][D] of one query as ssppcu_set_queries builds it (float32)"""
    K = cfg.n_points
    t = (np.arange(K) / (K - 1))[:, None]
    V = (1 - t) * np.asarray(start, np.float64)[None, :] + t * np.asarray(end, np.float64)[None, :]
    if cfg.mode == MODE_TASK:
        return V.astype(np.float32)
    _, ainv = spline_setup(K, cfg.degree)
    out = np.zeros_like(V)
    for i in range(K):           # same accumulation order as the C side
        for d in range(V.shape[1]):
            acc = 0.0
            for k in range(K):
                acc += ainv[i, k] * V[k, d]
            out[i, d] = acc
    return out.astype(np.float32)


class Model:
    """Flat FP64 model parsed from an MJCF file or string (no MuJoCo involved)."""

    def __init__(self, path: str | None = None, text: str | None = None):
        self._h = C.c_void_p()
        err = C.create_string_buffer(1024)
        if path is not None:
            rc = lib.ssppcu_model_load(os.fsencode(path), C.byref(self._h), err, len(err))
        else:
            rc = lib.ssppcu_model_parse((text or "").encode(), C.byref(self._h), err, len(err))
        if rc != OK:
            raise SsppError(rc, err.value.decode())
        self.path = path
        self.view = FlatModelView()
        lib.ssppcu_model_view(self._h, C.byref(self.view))

    def __del__(self):
        h = getattr(self, "_h", None)
        if h and lib is not None:      # `lib` is already None during interpreter shutdown
            lib.ssppcu_model_free(h)
            self._h = None

    def body_id(self, name: str) -> int:
        return lib.ssppcu_model_body_id(self._h, name.encode())

    def body_name(self, b: int) -> str:
        return (lib.ssppcu_model_body_name(self._h, b) or b"").decode()

    def geom_name(self, g: int) -> str:
        return (lib.ssppcu_model_geom_name(self._h, g) or b"").decode()

    def body_point(self, name: str) -> np.ndarray:
        out = np.zeros(4)
        rc = lib.ssppcu_model_body_point(self._h, name.encode(), _dptr(out))
        if rc != OK:
            raise SsppError(rc, f"body '{name}' is not a free body of the model")
        return out

    def pair_counts(self, mode: int, dof: int, body: str = "") -> tuple[int, int, int]:
        a, b, c = C.c_int32(), C.c_int32(), C.c_int32()
        err = C.create_string_buffer(1024)
        rc = lib.ssppcu_model_pair_counts(self._h, mode, dof, body.encode(), C.byref(a), C.byref(b), C.byref(c), err, len(err))
        if rc != OK:
            raise SsppError(rc, err.value.decode())
        return a.value, b.value, c.value

    def arrays(self) -> dict:
        """numpy copies of the flat arrays"""
        v = self.view
        nb, ng = v.nbody, v.ngeom

        def ai(p, n): return np.ctypeslib.as_array(p, shape=(n,)).copy() if n else np.zeros(0, np.int32)
        def ad(p, n, k): return np.ctypeslib.as_array(p, shape=(n * k,)).reshape(n, k).copy() if n else np.zeros((0, k))
        return dict(nbody=nb, ngeom=ng, nq=v.nq,
                    body_parent=ai(v.body_parent, nb), body_weld=ai(v.body_weld, nb), body_jtype=ai(v.body_jtype, nb),
                    body_qadr=ai(v.body_qadr, nb), body_pos=ad(v.body_pos, nb, 3), body_quat=ad(v.body_quat, nb, 4),
                    jnt_axis=ad(v.jnt_axis, nb, 3), jnt_pos=ad(v.jnt_pos, nb, 3), jnt_range=ad(v.jnt_range, nb, 2),
                    qpos0=ad(v.qpos0, v.nq, 1).ravel(),
                    geom_body=ai(v.geom_body, ng), geom_type=ai(v.geom_type, ng), geom_contype=ai(v.geom_contype, ng),
                    geom_conaffinity=ai(v.geom_conaffinity, ng), geom_size=ad(v.geom_size, ng, 3),
                    geom_pos=ad(v.geom_pos, ng, 3), geom_quat=ad(v.geom_quat, ng, 4),
                    geom_margin=ad(v.geom_margin, ng, 1).ravel(),
                    exclude=ai(v.exclude, 2 * v.nexclude).reshape(-1, 2))


class _Results(dict):
    """the dict Handle.results returns; keeps the ctypes pointers of its arrays so that a refill converts nothing"""
    ptrs: list


class _Best(tuple):
    """the (best, best_valid) pair Handle.best returns; keeps the ctypes pointers of its arrays like _Results"""
    ptrs: tuple


class Handle:
    """One ssppcu_t: a scene specialised for one planner shape, resident on one GPU."""

    def __init__(self, model: Model, cfg: PlannerConfig):
        self._h = C.c_void_p()
        self.model = model
        self._body = cfg.body_name            # keep the bytes alive
        err = C.create_string_buffer(1024)
        rc = lib.ssppcu_create(model._h, C.byref(cfg), C.byref(self._h), err, len(err))
        if rc != OK:
            raise SsppError(rc, err.value.decode())
        self.dims = Dims()
        lib.ssppcu_dims(self._h, C.byref(self.dims))

    def close(self):
        if getattr(self, "_h", None) and lib is not None:
            lib.ssppcu_destroy(self._h)
            self._h = None

    __del__ = close

    def _ck(self, rc: int):
        if rc < 0:
            raise SsppError(rc, lib.ssppcu_last_error(self._h).decode())
        return rc

    def set_stream(self, stream_ptr: int | None):
        self._ck(lib.ssppcu_set_stream(self._h, stream_ptr))

    def set_result_mirror(self, on: bool):
        """on (default): the cooperative kernel writes its results into the pinned host mirror itself; off: device only"""
        self._ck(lib.ssppcu_set_result_mirror(self._h, int(bool(on))))

    def set_queries(self, start, end):
        d = self.dims
        s = np.ascontiguousarray(np.asarray(start, np.float64).reshape(d.n_queries, d.dof))
        e = np.ascontiguousarray(np.asarray(end, np.float64).reshape(d.n_queries, d.dof))
        self._ck(lib.ssppcu_set_queries(self._h, _dptr(s), _dptr(e)))

    def reset_distribution(self, sigma0: float):
        self._ck(lib.ssppcu_reset_distribution(self._h, float(sigma0)))

    def set_distribution(self, mean=None, sigma=None):
        d = self.dims
        m = None if mean is None else np.ascontiguousarray(np.asarray(mean, np.float64).reshape(d.n_queries, d.nfd))
        s = None if sigma is None else np.ascontiguousarray(np.asarray(sigma, np.float64).reshape(d.n_queries, d.nfd))
        self._ck(lib.ssppcu_set_distribution(self._h, None if m is None else _dptr(m), None if s is None else _dptr(s)))

    def get_distribution(self):
        d = self.dims
        m = np.zeros((d.n_queries, d.n_free, d.dof)); s = np.zeros_like(m)
        self._ck(lib.ssppcu_get_distribution(self._h, _dptr(m), _dptr(s)))
        return m, s

    def set_joint_noise(self, sigma: float, limits):
        lim = np.ascontiguousarray(np.asarray(limits, np.float64).reshape(self.dims.dof))
        self._ck(lib.ssppcu_set_joint_noise(self._h, float(sigma), _dptr(lim)))

    def iterate(self, seed: int, iter0: int = 0, n_iters: int = 1, restored: bool = False):
        """n_iters fused iterations; restored=True: restore_state() folded into the same launch"""
        fn = lib.ssppcu_iterate_restored if restored else lib.ssppcu_iterate
        self._ck(fn(self._h, C.c_uint64(seed), C.c_uint32(iter0), int(n_iters)))

    def plan_step(self, start, end, sigma: float, limits=None, seed: int = 0, iter0: int = 0, n_iters: int = 1):
        """set_queries + set_joint_noise / reset_distribution + iterate in one call and one launch (ssppcu_plan_step)"""
        s = np.ascontiguousarray(start, np.float64); e = np.ascontiguousarray(end, np.float64)
        lim = None if limits is None else np.ascontiguousarray(limits, np.float64)
        self._ck(lib.ssppcu_plan_step(self._h, _dptr(s), _dptr(e), float(sigma), None if lim is None else _dptr(lim), seed, iter0, n_iters))

    def bind_plan_step(self, start, end, limits=None):
        """plan_step over caller-owned host buffers whose addresses are taken ONCE, the way a C caller keeps its pointers (the
        ndarray -> ctypes pointer conversion costs microseconds per argument in Python).  start / end / limits must be C-contiguous
        float64 arrays; the returned callable (sigma, seed, iter0, n_iters) reads their CURRENT contents on every call."""
        for a in (start, end) + (() if limits is None else (limits,)):
            if not (isinstance(a, np.ndarray) and a.dtype == np.float64 and a.flags.c_contiguous):
                raise ValueError("bind_plan_step needs C-contiguous float64 ndarrays")
        ps, pe, pl = _dptr(start), _dptr(end), (None if limits is None else _dptr(limits))
        keep = (start, end, limits)
        fn, hh, ck = lib.ssppcu_plan_step, self._h, self._ck

        def step(sigma: float, seed: int = 0, iter0: int = 0, n_iters: int = 1, _keep=keep):
            ck(fn(hh, ps, pe, sigma, pl, seed, iter0, n_iters))
        return step

    def set_success_rows(self, on: bool):
        lib.ssppcu_set_success_rows.argtypes = [C.c_void_p, C.c_int]
        self._ck(lib.ssppcu_set_success_rows(self._h, int(on)))

    def success_rows(self, query: int = 0):
        """(rows [n][nfd], candidate indices [n]) of the converged candidates of the last launch, candidate order; views into
        the handle's pinned block, valid until the next launch"""
        rows, idx, n = C.POINTER(C.c_float)(), C.POINTER(C.c_int32)(), C.c_int32()
        lib.ssppcu_get_success_rows.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p]
        self._ck(lib.ssppcu_get_success_rows(self._h, query, C.byref(rows), C.byref(idx), C.byref(n)))
        d = self.dims
        if n.value == 0:
            return np.zeros((0, d.nfd), np.float32), np.zeros(0, np.int32)
        return np.ctypeslib.as_array(rows, shape=(n.value, d.nfd)), np.ctypeslib.as_array(idx, shape=(n.value,))

    def candidate_ctrl(self, idx: int, query: int = 0) -> np.ndarray:
        d = self.dims
        out = np.zeros((d.n_points, d.dof), np.float32)
        lib.ssppcu_candidate_ctrl.argtypes = [C.c_void_p, C.c_int, C.c_uint32, C.c_void_p]
        self._ck(lib.ssppcu_candidate_ctrl(self._h, query, idx, out.ctypes.data_as(C.c_void_p)))
        return out

    def sync(self):
        self._ck(lib.ssppcu_sync(self._h))

    def phase_probe(self, enable: bool = True, read: bool = False):
        """diagnostics of the cooperative kernel: per-CTA %globaltimer stamps [grid][12] of the last launch's first
        iteration (start, scored, past barrier 1, ranked, past barrier 2, updated, rank keys staged, rank counted; fused key
        exchange: shard ordered, published, peers' flags seen, merge ranked - and `updated` = finished); None unless `read`"""
        lib.ssppcu_phase_probe.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_int]
        grid = lib.ssppcu_phase_probe(self._h, int(enable), None, 0)
        if grid < 0:
            self._ck(grid)
        if not read:
            return None
        out = np.zeros((grid, 12), np.uint64)
        self._ck(min(0, lib.ssppcu_phase_probe(self._h, int(enable), out.ctypes.data_as(C.c_void_p), grid)))
        return out

    def work_counters(self) -> dict:
        """executed work of the last iteration's candidates (untimed diagnostic kernel): counter name -> total"""
        out = np.zeros(len(WORK_COUNTERS), np.uint64)
        lib.ssppcu_work_counters.argtypes = [C.c_void_p, C.c_void_p, C.c_int]
        n = lib.ssppcu_work_counters(self._h, out.ctypes.data_as(C.c_void_p), len(out))
        if n < 0:
            self._ck(n)
        return {k: int(v) for k, v in zip(WORK_COUNTERS, out)}

    def sdf_gradient(self, pts, weights=None):
        """SDF mode refinement step: (grad [n_points][3], clearance [rows], dirs [rows][3]); see ssppcu_sdf_gradient"""
        pts = np.ascontiguousarray(pts, np.float32)
        K = self.dims.n_points
        if weights is None:
            rows, W = self.collocation()[0].shape[0], None
        else:
            W = np.ascontiguousarray(weights, np.float32); rows = W.shape[0]
            assert W.shape[1] == K
        assert pts.shape == (K, self.dims.dof)
        grad = np.zeros((K, 3), np.float32); dist = np.zeros(rows, np.float32); dirs = np.zeros((rows, 3), np.float32)
        self._ck(lib.ssppcu_sdf_gradient(self._h, _ptr(pts), _ptr(W), rows, _ptr(grad), _ptr(dist), _ptr(dirs)))
        return grad, dist, dirs

    def save_state(self):
        self._ck(lib.ssppcu_save_state(self._h))

    def restore_state(self):
        self._ck(lib.ssppcu_restore_state(self._h))

    _RES_KEYS = ("L", "Cnf", "Cwf", "cost", "status", "n_success", "n_elite", "best_idx", "best_cost", "any_success")

    def results(self, out: dict | None = None) -> dict:
        """every candidate's L / Cnf / Cwf / cost / status and the per-query scalars.  `out`: a dict returned by an earlier
        call on this handle, filled again in place (caller-owned buffers, as a C caller of ssppcu_get_results would keep)"""
        if out is None:
            d = self.dims
            Q, N = d.n_queries, d.n_cand
            out = _Results(L=np.empty((Q, N), np.float32), Cnf=np.empty((Q, N), np.float32), Cwf=np.empty((Q, N), np.float32),
                       cost=np.empty((Q, N), np.float32), status=np.empty((Q, N), np.uint8),
                       n_success=np.empty(Q, np.int32), n_elite=np.empty(Q, np.int32), best_idx=np.empty(Q, np.int32),
                       best_cost=np.empty(Q, np.float32), any_success=np.empty(Q, np.int32))
            out.ptrs = [_ptr(out[k]) for k in self._RES_KEYS]
        self._ck(lib.ssppcu_get_results(self._h, *out.ptrs))
        return out

    def params(self) -> np.ndarray:
        d = self.dims
        p = np.zeros((d.n_queries, d.n_cand, d.n_free, d.dof), np.float32)
        self._ck(lib.ssppcu_get_params(self._h, _ptr(p)))
        return p

    def best(self, out: tuple | None = None):
        """forwarded best candidate of every query and its validity flag; `out`: the pair of an earlier call, refilled in place"""
        if out is None:
            d = self.dims
            out = (np.zeros((d.n_queries, d.n_free, d.dof), np.float32), np.zeros(d.n_queries, np.int32))
        if not isinstance(out, _Best):
            out = _Best(out)
            out.ptrs = (_ptr(out[0]), _ptr(out[1]))
        self._ck(lib.ssppcu_get_best(self._h, *out.ptrs))
        return out

    def check_points(self, pts):
        """Collision.h:51-82 for a batch of configurations: (geom1, geom2, dist) per point, geom ids of the penetrating contact
        with the smallest geom pair, -1 / -1 / 0 when free"""
        p = np.ascontiguousarray(pts, np.float32).reshape(-1, self.dims.dof)
        g1 = np.zeros(len(p), np.int32); g2 = np.zeros(len(p), np.int32); dist = np.zeros(len(p), np.float32)
        lib.ssppcu_check_points.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p]
        self._ck(lib.ssppcu_check_points(self._h, _ptr(p), len(p), _ptr(g1), _ptr(g2), _ptr(dist)))
        return g1, g2, dist

    def elites(self, query: int = 0, cap: int | None = None):
        cap = self.dims.n_cand if cap is None else cap
        idx = np.zeros(cap, np.uint32); cost = np.zeros(cap, np.float32); w = np.zeros(cap, np.float32)
        n = self._ck(lib.ssppcu_get_elites(self._h, query, _ptr(idx), _ptr(cost), _ptr(w), cap))
        return idx[:n], cost[:n], w[:n]

    def device_view(self) -> DeviceView:
        v = DeviceView()
        self._ck(lib.ssppcu_device_view(self._h, C.byref(v)))
        return v

    def eval_points(self, pts) -> np.ndarray:
        p = np.ascontiguousarray(np.asarray(pts, np.float32).reshape(-1, self.dims.dof))
        out = np.zeros(len(p), np.float32)
        self._ck(lib.ssppcu_eval_points(self._h, _ptr(p), len(p), _ptr(out)))
        return out

    def sdf(self):
        """(nodes [nz][ny][nx] float32, lo[3], cell[3]) of an SDF-mode handle"""
        dims = (C.c_int32 * 3)(); lo = np.zeros(3); cell = np.zeros(3)
        self._ck(lib.ssppcu_get_sdf(self._h, None, dims, _dptr(lo), _dptr(cell)))
        nodes = np.zeros((dims[2], dims[1], dims[0]), np.float32)
        self._ck(lib.ssppcu_get_sdf(self._h, _ptr(nodes), dims, _dptr(lo), _dptr(cell)))
        return nodes, lo, cell

    def spheres(self) -> np.ndarray:
        n = lib.ssppcu_get_spheres(self._h, None, 0)
        out = np.zeros((max(n, 1), 5), np.float32)
        lib.ssppcu_get_spheres(self._h, _ptr(out), n)
        return out[:n]

    def collocation(self):
        d = self.dims
        pc = np.zeros((d.rows_collision, d.n_points), np.float32)
        pa = np.zeros((d.rows_arc, d.n_points), np.float32)
        self._ck(lib.ssppcu_get_collocation(self._h, _ptr(pc), _ptr(pa)))
        return pc, pa

    def base(self) -> np.ndarray:
        d = self.dims
        b = np.zeros((d.n_queries, d.n_points, d.dof), np.float32)
        self._ck(lib.ssppcu_get_base(self._h, _ptr(b)))
        return b

    def shard_keys(self) -> tuple[int, int]:
        p = C.c_void_p(); n = C.c_int32()
        self._ck(lib.ssppcu_shard_keys(self._h, C.byref(p), C.byref(n)))
        return p.value, n.value

    def merge_update(self, dev_keys_ptr: int, n_keys: int, n_shards: int, seed: int, it: int):
        self._ck(lib.ssppcu_merge_update(self._h, C.c_void_p(dev_keys_ptr), int(n_keys), int(n_shards), C.c_uint64(seed), C.c_uint32(it)))

    # population shard with the key exchange fused into the kernel (ssppcu_exchange_*, include/sspp_cuda.h)
    def exchange_export(self, world: int, rank: int) -> tuple[int, bytes]:
        """allocate this rank's exchange block; returns (device address, 64-byte CUDA IPC handle)"""
        blk = C.c_void_p(); ipc = C.create_string_buffer(64)
        self._ck(lib.ssppcu_exchange_export(self._h, int(world), int(rank), C.byref(blk), ipc))
        return blk.value, ipc.raw

    def exchange_attach_ipc(self, handles: list[bytes]):
        buf = b"".join(handles)
        self._ck(lib.ssppcu_exchange_attach_ipc(self._h, C.c_char_p(buf)))

    def exchange_attach_ptrs(self, blocks: list[int]):
        arr = (C.c_void_p * len(blocks))(*blocks)
        self._ck(lib.ssppcu_exchange_attach_ptrs(self._h, arr))

    def iterate_exchange(self, seed: int, it: int):
        self._ck(lib.ssppcu_iterate_exchange(self._h, C.c_uint64(seed), C.c_uint32(it)))

    def exchange_status(self):
        self._ck(lib.ssppcu_exchange_status(self._h))

    def merged(self, cap: int):
        idx = np.zeros(cap, np.uint32); cost = np.zeros(cap, np.float32); w = np.zeros(cap, np.float32)
        n = self._ck(lib.ssppcu_get_merged(self._h, _ptr(idx), _ptr(cost), _ptr(w), cap))
        return idx[:n], cost[:n], w[:n]
