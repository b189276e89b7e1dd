#!/usr/bin/env python
"""bench.py — scored collocation points/s of the fused candidate-scoring iteration on B200.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--workload cfg1|cfg2|cfg3|cfg4|cfg5]
  (N > 1: python -m torch.distributed.run --nproc-per-node N ... bench.py --gpus N ...)

Workloads (BASELINE.json configs):
  cfg2 (default)  robocrane 7-DoF fixture, ONE planning query, 4096 candidates x 64 collocation points x 10 control points.
                  One STEP = one planner iteration of that query from its initial sampling state, ONE fused launch
                  (re-arm the state -> Philox perturb -> spline -> kinematics -> collision -> cost reduce -> elites -> update).
                  The state is re-armed every step so that every step scores the same statistical workload.
  cfg1            the reference's own CPU-runnable case: KAT #1 (scripts/bspline_params.npy reproduced by the product's
                  BSplines / CubicPath functions) + one Planner-A query with scripts/main.py:168-173 parameters (100 x 100 x 7).
  cfg3            stacking scene + 64 clutter boxes, SDF-grid collision, 65536 candidates x 64 points: one GPU holds the whole
                  population; N > 1 shards ONE population over the ranks (8192 * N candidates, key exchange fused into the kernel).
  cfg4            1024 independent ICRA-style queries (robocrane.xml gripper), a step = a whole 60-iteration plan of each.
  cfg5            pairwise path-cost matrix over 256 waypoints: 32 640 Planner-B plans, a step = the whole matrix.
N > 1 (cfg1/2/4): independent planning queries shard across GPUs, no data-path collective ("scaling": "weak"):
value = points of all ranks / max-over-ranks time.  At N > 1 the default line also carries a `popshard` record: one SDF
population sharded over the N ranks through both transports, checked bit for bit against the unsharded handle.

Prints ONE JSON line (rank 0).  `value`: device-resident throughput (CUDA events on the launch stream); `e2e`: through the C ABI
with host buffers in and every candidate's result out; `e2e_plan` (cfg1/cfg2): through the drop-in Python call
`_sspp.SamplingPathPlanner7.plan()` with the returned list of splines materialised.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time
import zlib
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent
sys.path.insert(0, str(ROOT))

SEED = 0x53535050
METRIC = "scored collocation points/sec"
UNIT = "points/s"
WORKLOADS = ("cfg1", "cfg2", "cfg3", "cfg4", "cfg5")


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=20)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="cfg2", choices=WORKLOADS)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-popshard", action="store_true", help="N > 1: skip the population-sharded record of the default line")
    return ap.parse_args()


def workload_case(name: str, rank: int = 0):
    """the synthetic inputs of SURVEY 8(d); rank selects this rank's independent query/queries"""
    from tests.helpers import joint_case, task_case, sdf_case, Q_START, Q_END
    rng = np.random.default_rng(1000 + rank)
    if name in ("cfg1", "cfg2"):
        if name == "cfg2":
            case = joint_case(n_samples=4096, check_points=64, n_points=10, do_update=True)
            desc = dict(workload="cfg2: robocrane7 fixture, 1 query x 4096 candidates x 64 collocation points x 10 control points (cubic), "
                                 "sigma 0.08*pi, CES update on, state re-armed every step",
                        candidates=4096, collocation_points=64, control_points=10, dof=7, queries_per_gpu=1)
        else:
            case = joint_case(n_samples=100, check_points=100, n_points=7, do_update=False)
            desc = dict(workload="cfg1: robocrane7 fixture among planner.xml's obstacles, ONE Planner-A query with scripts/main.py:168-173 "
                                 "parameters (100 candidates x 100 collocation points x 7 control points, sigma 0.08, limits pi), start/end "
                                 "from scripts/bspline_params.npy; plus KAT #1 pass/fail",
                        candidates=100, collocation_points=100, control_points=7, dof=7, queries_per_gpu=1)
        if rank > 0:   # another start/goal pair near the KAT configurations
            case.start = (Q_START + rng.normal(0, 0.05, 7))[None, :]
            case.end = (Q_END + rng.normal(0, 0.05, 7))[None, :]
        return case, desc
    if name == "cfg3":
        case = sdf_case(n_samples=65536 - 2, check_points=64, n_points=5, dims=(128, 128, 128))
        desc = dict(workload="cfg3: stacking.xml + 64 seeded clutter boxes, block1 moving, SDF grid 128^3, 65536 candidates x 64 collocation "
                             "points x 5 points (3 vias, quadratic), CES update on, state re-armed every step",
                    candidates=65536, collocation_points=64, control_points=5, dof=4, queries_per_gpu=1)
        return case, desc
    Q = 1024
    s = np.column_stack([rng.uniform(0.2, 0.6, Q), rng.uniform(-0.4, 0.4, Q), rng.uniform(0.14, 0.4, Q), rng.uniform(-1.5, 1.5, Q)])
    e = np.column_stack([rng.uniform(0.2, 0.6, Q), rng.uniform(-0.4, 0.4, Q), rng.uniform(0.14, 0.4, Q), rng.uniform(-1.5, 1.5, Q)])
    case = task_case(n_queries=Q, start=s, end=e)
    desc = dict(workload="cfg4: robocrane.xml gripper, 1024 ICRA-style queries x 17 candidates x 40 collocation points, 1 via (quadratic), "
                         "a step = the 60-iteration cold-start plan of every query",
                candidates=17, collocation_points=40, control_points=3, dof=4, queries_per_gpu=Q)
    return case, desc


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled during the timed region"""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int):
        self.rows, self.proc = [], None
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-i", str(index),
                                          "-lms", "100"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([x.strip() for x in line.split(",")])

    def stop(self) -> dict:
        if not self.proc:
            return dict(sm_mhz=None, sm_max_mhz=None, reasons=["nvidia-smi unavailable"])
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except subprocess.TimeoutExpired:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        for r in self.rows:
            try:
                sm.append(float(r[0])); mx.append(float(r[1]))
            except (ValueError, IndexError):
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), r[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        return dict(sm_mhz=float(np.median(sm)) if sm else None, sm_max_mhz=max(mx) if mx else None,
                    reasons=sorted(reasons), samples=len(sm))


# ------------------------------------------------------------------------------------------
# CPU arm.  Planner A / B shapes: the REFERENCE'S OWN plan() (oracle/_ref/libref.so = sspp.h / sspp/tsp*.h compiled
# unmodified; Eigen and MuJoCo replaced by the stand-ins of oracle/shim, MuJoCo's kinematics + contacts = oracle O1's FP64
# restatement) when that library was shipped, else - and for the SDF mode, which the reference does not have - the O1 port.
# Neither maps the product library: scenes come from tests/golden/flat/*.npz.
# ------------------------------------------------------------------------------------------
def host_threads() -> int:
    """all the host threads the process may use (torchrun exports OMP_NUM_THREADS=1 to its workers; the OpenMP runtimes the
    oracles are linked to are told explicitly)"""
    n = len(os.sched_getaffinity(0))
    try:
        import ctypes
        gomp = ctypes.CDLL("libgomp.so.1")
        gomp.omp_set_num_threads(n)
        n = int(gomp.omp_get_max_threads())
    except OSError:
        n = int(os.environ.get("OMP_NUM_THREADS", n))
    return n


def set_threads(n: int):
    import ctypes
    try:
        ctypes.CDLL("libgomp.so.1").omp_set_num_threads(n)
    except OSError:
        pass
    from oracle import pyref
    if pyref.available():
        pyref.set_threads(n)


class CpuArm:
    def __init__(self, case, workload: str):
        from oracle import pyoracle, pyref            # the one place bench.py executes oracle/: as the CPU baseline
        from oracle.flatmodel import FlatModel
        self.case, self.workload = case, workload
        c = case.cfg
        self.task = c.mode == 1
        self.sdf = c.collision_mode == 1
        fm = FlatModel.load(case.mjcf.stem)
        self.o1 = pyoracle.O1(fm, c.mode, c.dof, case.body)
        self.oc = pyoracle.o1_cfg_from(c)
        if self.sdf:                                  # sample spheres are an input of the SDF mode: from the replica's host build
            sp = pyoracle.O2(str(case.mjcf), c).spheres()
            self.o1.set_sdf(list(c.sdf_dims), list(c.sdf_lo), list(c.sdf_hi), sp, 0.0)
        self.nfree = c.n_points - 2 if self.task else c.n_points - 2 * c.degree
        self.fb = 1 if self.task else c.degree
        K = c.n_points
        self.base = []
        for q in range(len(case.start)):
            V = np.array([(1 - i / (K - 1)) * case.start[q] + (i / (K - 1)) * case.end[q] for i in range(K)], np.float64)
            self.base.append(V if self.task else self.o1.interpolate(V, c.degree))
        self.mean = [b[self.fb:self.fb + self.nfree].copy() for b in self.base]
        if self.task:
            self.sigma = [np.full_like(m, case.sigma0) for m in self.mean]
        else:
            self.sigma = [np.tile(case.joint_sigma * case.limits, (self.nfree, 1)) for _ in self.mean]
        self.cores = host_threads()
        self.M = c.check_points
        self.ref = None
        if pyref.available() and not self.sdf:
            rm = pyref.RefModel(fm)
            if self.task:
                from tests.ref_helpers import tsp_args_from
                self.ref = pyref.RefTsp(rm, case.body, tsp_args_from(case))
            elif c.dof in (3, 6, 7, 9):
                self.ref = pyref.RefSspp(rm, c.dof)
        self.kind = "reference" if self.ref is not None else "port"

    def describe(self) -> str:
        if self.ref is not None:
            what = "tsp::TaskSpacePlanner::plan (include/sspp/tsp.h)" if self.task else "sspp::SamplingPathPlanner::plan (include/sspp.h)"
            return (f"the reference's own {what}, headers compiled unmodified into oracle/_ref/libref.so; its third-party calls go to stand-ins: "
                    "Eigen splines restated, mj_forward = oracle O1's FP64 kinematics + narrow phase (positions and contacts only, no "
                    "dynamics - cheaper than MuJoCo's), Philox normals")
        return "oracle O1 (FP64 restatement of the reference CPU path" + (", SDF-grid collision mode, which the reference does not have)" if self.sdf else ")")

    # --- one bounded sample of the workload: n_cand candidates of query `query`; returns scored points (N x M) ---
    def step(self, it: int, n_cand: int, query: int = 0) -> int:
        c = self.case
        if self.ref is not None and not self.task:    # Planner A: plan() = sample + collision-check every candidate (early exit) + best path
            from oracle import pyref
            pyref.arm_tape(pyref.MODE_JOINT, SEED, it, query, 0)
            self.ref.plan(c.start[query], c.end[query], c.joint_sigma, c.limits, n_cand, self.M, c.cfg.n_points, threads=self.threads)
            return n_cand * self.M
        o1, oc = self.o1, self.oc
        p = o1.sample(oc, SEED, it, query, 0, n_cand, self.mean[query], self.sigma[query])
        s = o1.score(oc, self.base[query], p)
        idx, w, ns = o1.elites(s["cost"], s["status"], c.cfg.elite_fraction)
        o1.update(oc, self.mean[query], self.sigma[query], p[idx], w)     # result discarded: state re-armed every step
        return n_cand * self.M

    def plan(self, query: int, iters: int, n_cand: int) -> int:
        """one whole cold-start plan of a query (main_icra_benchmark.cpp:186-204: queries one after another, OpenMP inside plan())"""
        c = self.case
        if self.ref is not None and self.task:
            from oracle import pyref
            pyref.arm_tape(pyref.MODE_TASK, SEED, 0, query, 2, 3 if c.cfg.lo[3] == c.cfg.hi[3] else 4, max(self.nfree, 1))
            return self.ref.run(c.start[query], c.end[query], iters, cold=True)
        o1, oc = self.o1, self.oc
        mean, sigma = self.mean[query].copy(), self.sigma[query].copy()
        for it in range(iters):
            p = o1.sample(oc, SEED, it, query, 0, n_cand, mean, sigma)
            s = o1.score(oc, self.base[query], p)
            idx, w, ns = o1.elites(s["cost"], s["status"], c.cfg.elite_fraction)
            if ns > 0:
                m, sg = o1.update(oc, mean, sigma, p[idx], w)
                mean, sigma = m.reshape(mean.shape), sg.reshape(sigma.shape)
        return iters * n_cand * self.M

    threads = 1

    def use_threads(self, n: int):
        self.threads = n
        set_threads(n)

    def unit(self, n: int, i: int) -> int:
        """the bounded sample the workload is timed on: cfg4/cfg5 whole plans of n queries, otherwise n candidates of the query"""
        if self.workload in ("cfg4", "cfg5"):
            Q = len(self.case.start)
            return sum(self.plan((i * n + q) % Q, 60, self.case.cfg.n_samples) for q in range(n))
        return self.step(i, n)

    def calibrate(self, budget_s: float) -> int:
        """units per sample so that one sample takes about budget_s"""
        n = 1 if self.workload in ("cfg4", "cfg5") else 8
        cap = len(self.case.start) if self.workload in ("cfg4", "cfg5") else self.case.cfg.n_samples
        while True:
            t0 = time.perf_counter(); self.unit(n, 0); dt = time.perf_counter() - t0
            if dt >= 0.25 * budget_s or n >= cap:
                break
            n = min(cap, max(n + 1, int(n * min(8.0, 0.6 * budget_s / max(dt, 1e-6)))))
        return n

    def sample_text(self, n: int) -> str:
        c = self.case.cfg
        if self.workload in ("cfg4", "cfg5"):
            return (f"each sample = {n} of {len(self.case.start)} queries x 60 iterations x {c.n_samples + 2} candidates x {self.M} points "
                    "(cold-start plans, queries in sequence)")
        return f"each sample = {n} of {c.n_samples} candidates x {self.M} points of the same query"


def cpu_baseline(case, workload: str, budget_s: float = 18.0) -> dict:
    """median of >= 5 repetitions on all host threads + a 1-thread figure"""
    arm = CpuArm(case, workload)
    arm.use_threads(arm.cores)
    n = arm.calibrate(budget_s * 0.5 / 5)
    vals = []
    for r in range(5):
        t0 = time.perf_counter(); pts = arm.unit(n, 1 + r); vals.append(pts / (time.perf_counter() - t0))
    arm.use_threads(1)
    n1 = max(1, n // max(1, min(arm.cores, 8)))
    v1 = []
    for r in range(3):
        t0 = time.perf_counter(); pts = arm.unit(n1, 1 + r); v1.append(pts / (time.perf_counter() - t0))
    arm.use_threads(arm.cores)
    return dict(value=float(np.median(vals)), unit=UNIT, cores=arm.cores, kind=arm.kind, reps=len(vals),
                spread=float((max(vals) - min(vals)) / np.median(vals)), value_1thread=float(np.median(v1)),
                sample=f"{arm.describe()}; {arm.sample_text(n)}; median of {len(vals)} repetitions, OpenMP {arm.cores} threads")


def run_reference(args, rank: int, world: int):
    """--impl reference: the CPU path on the host cores, a bounded sample per step."""
    if rank != 0:
        return
    wl = "cfg4" if args.workload == "cfg5" else args.workload
    case, desc = workload_case(wl, 0)
    if args.workload == "cfg5":
        desc = dict(desc, workload="cfg5 (CPU arm): the same Planner-B plans as cfg4, one after another - the cost matrix is 32 640 of them")
    arm = CpuArm(case, args.workload)
    arm.use_threads(arm.cores)
    total = args.steps + args.warmup
    n = arm.calibrate(min(2.0, 75.0 / max(total, 1)))
    for i in range(args.warmup):
        arm.unit(n, i)
    t0 = time.perf_counter()
    pts = 0
    for i in range(args.steps):
        pts += arm.unit(n, args.warmup + i)
    dt = time.perf_counter() - t0
    val = pts / dt
    sample = f"{arm.describe()}; {arm.sample_text(n)}; OpenMP {arm.cores} threads"
    line = dict(impl="reference", metric=METRIC, value=val, unit=UNIT, n_gpus=args.gpus, steps=args.steps, warmup=args.warmup,
                ms_per_step=dt / max(args.steps, 1) * 1e3, higher_is_better=True, scaling="weak", vs_baseline=None, dtype="f64",
                data="synthetic", config=desc,
                cpu_baseline=dict(value=val, unit=UNIT, cores=arm.cores, kind=arm.kind, sample=sample),
                e2e=dict(value=val, unit=UNIT, h2d_bytes_per_step=0, d2h_bytes_per_step=0))
    emit(line)


# ------------------------------------------------------------------------------------------
# GPU arm
# ------------------------------------------------------------------------------------------
class Dist:
    def __init__(self, rank, world, local_rank):
        import torch
        import torch.distributed as dist
        self.torch, self.dist, self.rank, self.world, self.local = torch, dist, rank, world, local_rank
        torch.cuda.set_device(local_rank)
        if world > 1:
            dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))

    def barrier(self):
        if self.world > 1:
            self.dist.barrier()
        self.torch.cuda.synchronize()

    def max_over_ranks(self, x: float) -> float:
        if self.world == 1:
            return x
        t = self.torch.tensor([x], dtype=self.torch.float64, device=self.torch.device("cuda", self.local))
        self.dist.all_reduce(t, op=self.dist.ReduceOp.MAX)
        return float(t.item())

    def all_true(self, ok: bool) -> bool:
        return self.max_over_ranks(0.0 if ok else 1.0) == 0.0


def timed_windows(D: Dist, fn, est_s: float, n_windows: int = 5, min_s: float = 0.05, min_steps: int = 1):
    """>= n_windows host-clock windows of >= min_s each (barrier + synchronize on both sides, max over ranks); returns
    (steps per window, [seconds per step of every window])"""
    n = max(min_steps, int(np.ceil(min_s / max(est_s, 1e-7))))
    out, k = [], 0
    for w in range(n_windows):
        D.barrier()
        t0 = time.perf_counter()
        for i in range(n):
            fn(k); k += 1
        D.barrier()
        out.append(D.max_over_ranks(time.perf_counter() - t0) / n)
    return n, out


def kat1() -> dict:
    """cfg1 (i): scripts/main_bspline.py:176-209 through the product's BSplines / CubicPath functions reproduces the stored
    scripts/bspline_params.npy (kept in tests/golden/spline_golden.npz) from its own end points"""
    from sspp_b200 import splines as bs
    g = np.load(ROOT / "tests" / "golden" / "spline_golden.npz")
    t, c, k = g["kat_knot_vec"], g["kat_ctr_pts"], int(g["kat_k"])
    A = np.array([[bs.B(i / 6, k, j, t) for j in range(7)] for i in range(7)])
    A[0, 0] = A[6, 6] = 1.0
    via = A @ c
    path = bs.CubicPath(); path.plan(via[0], via[3], via[6])
    ctrl, t2 = bs.compute_control_points(path.evaluate(np.arange(7) / 6), k)
    err = float(np.abs(ctrl - c).max())
    return dict(passed=bool(err < 1e-12 and np.array_equal(t2, t)), max_abs_err=err)


def roofline_record(h, case, args, D, pts_per_step, k_ms, k_ms_isolated, inner, workload) -> dict:
    from sspp_b200 import capi, roofline
    c, d = case.cfg, h.dims
    fp32_peak, _ = capi.fp32_peak(D.local)
    peaks = json.loads((ROOT / "MEASURED_PEAKS.json").read_text()) if (ROOT / "MEASURED_PEAKS.json").exists() else {}
    # F_executed: one untimed iteration from the re-armed state, counted by the diagnostic kernel
    h.iterate(SEED, args.warmup, 1, restored=True)
    wc = h.work_counters()
    ex = roofline.executed_flop(wc, n_cand_total=d.n_queries * d.n_cand, check_points=c.check_points, dof=c.dof, n_points=c.n_points,
                                degree=c.degree, mode=c.mode, limit_penalty=c.limit_scale > 0)
    pps = pts_per_step / (k_ms * 1e-3)
    ach = pps * ex["flop_per_point_executed"] * 1e-12
    rec = dict(bound="fp32", achieved=ach, peak=fp32_peak, unit="TFLOP/s", frac=ach / fp32_peak, kernel="sspp_iterate_kernel",
               kernel_ms=k_ms, kernel_ms_isolated=k_ms_isolated, flop_per_point_executed=ex["flop_per_point_executed"],
               executed=dict({k: round(v, 2) for k, v in ex.items()}, counters_iteration="first iteration from the re-armed state"),
               peak_source="FMA probe measured live by ssppcu_fp32_peak (MEASURED_PEAKS.json has no FP32 entry)")
    if not (c.collision_mode == 1):
        wm = roofline.work_model(case.model, c.mode, c.dof, c.n_points, case.body)
        rec.update(flop_per_point_nominal=wm["flop_per_point"], frac_nominal=pps * wm["flop_per_point"] * 1e-12 / fp32_peak)
    rec["note"] = ("frac = EXECUTED floating-point work (work counters x operation counts of sspp_math.h, sspp_b200/roofline.py) over the FMA "
                   "peak; frac_nominal charges every filtered pair its full narrow phase (what MuJoCo would test) although the compile step "
                   "prunes pairs and the broad phase culls most - it is not a physical utilisation.  The kernel is instruction-issue bound "
                   "with a minority of FP32 instructions: see issue_active / pipe figures (ncu, profiles/)")
    kf = ROOT / "profiles" / "r02_iterate_keys.json"
    if kf.exists():
        keys = json.loads(kf.read_text()).get(workload)
        if keys:
            rec["ncu"] = keys                          # smsp__issue_active, pipe utilisation, executed instructions, DRAM bytes per launch
            rec["traffic"] = keys.get("dram_bytes_per_launch")
    rec.setdefault("traffic", None)
    hbm_bytes = d.n_queries * d.n_cand * inner * roofline.BYTES_PER_CANDIDATE
    hbm_ach = hbm_bytes / (k_ms * 1e-3) * 1e-9
    rec["hbm"] = dict(achieved=hbm_ach, peak=peaks.get("hbm_gbs"), unit="GB/s", frac=(hbm_ach / peaks["hbm_gbs"]) if peaks.get("hbm_gbs") else None,
                      peak_source="MEASURED_PEAKS.json hbm_gbs" if peaks else "absent")
    return rec


def run_handle_workload(args, D: Dist, workload: str) -> dict | None:
    """cfg1 / cfg2 / cfg3 on one GPU per rank / cfg4: a step = one launch of the handle from its re-armed state"""
    torch = D.torch
    from sspp_b200 import capi
    case, desc = workload_case(workload, D.rank)
    case.cfg.device = D.local
    h = capi.Handle(case.model, case.cfg)
    stream = torch.cuda.Stream()
    h.set_stream(stream.cuda_stream)
    case.prepare(h)
    h.save_state()
    d = h.dims
    inner = 60 if workload == "cfg4" else 1
    pts_per_step = d.n_queries * d.n_cand * case.cfg.check_points * inner
    launches_per_step = 1 if d.ctas_per_query == 1 else inner
    joint = case.cfg.mode == 0

    def step(i: int):
        h.iterate(SEED, i * inner, inner, restored=True)      # re-arm from the snapshot + iterate: ONE launch

    clocks = ClockSampler(D.local)
    h.set_result_mirror(False)                       # value / roofline: results stay in HBM, nothing is written to the host
    for i in range(max(args.warmup, 3)):
        step(i)
    # ---- value: K steps, device resident, CUDA events on the launch stream ----
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    D.barrier()
    ev0.record(stream)
    for i in range(args.steps):
        step(args.warmup + i)
    ev1.record(stream)
    D.barrier()
    ms_local = ev0.elapsed_time(ev1)
    ms_total = D.max_over_ranks(ms_local)
    value = pts_per_step * D.world * args.steps / (ms_total * 1e-3)
    # ---- the fused kernel alone, one event pair per launch ----
    nk = min(args.steps, 100)
    evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(nk)]
    for i, (a, b) in enumerate(evs):
        a.record(stream); step(args.warmup + i); b.record(stream)
    torch.cuda.synchronize()
    k_ms_isolated = float(np.mean([a.elapsed_time(b) for a, b in evs]))      # one event pair per launch (adds the event gaps)
    k_ms = ms_local / args.steps      # the kernel's time per step over the timed region (a step is this kernel only)
    # ---- e2e: through the C ABI, host buffers in, every candidate's result out ----
    h.set_result_mirror(True)                        # the library's default: the kernel fills the pinned result mirror itself
    res, best = h.results(), h.best()                # caller-owned host buffers, refilled by every step
    sig = case.joint_sigma if joint else case.sigma0

    # caller-owned host input buffers; their addresses are taken once, as a C caller keeps its pointers (every call reads their
    # current contents: nothing about the inputs is cached on the device or in the library between steps)
    in_start = np.ascontiguousarray(case.start, np.float64); in_end = np.ascontiguousarray(case.end, np.float64)
    in_lim = np.ascontiguousarray(case.limits, np.float64) if joint else None
    bound_step = h.bind_plan_step(in_start, in_end, in_lim)

    def e2e_step(i: int):
        nonlocal res, best
        bound_step(sig, SEED, i * inner, inner)      # host start/end (+ noise scale) in
        res = h.results(res)                         # L, Cnf, Cwf, cost, status per candidate + per-query scalars out
        best = h.best(best)
    for i in range(max(args.warmup, 3)):             # untimed: the mirroring instantiation of the kernel is loaded by its first launch
        e2e_step(i)
    t0 = time.perf_counter(); e2e_step(0); est = time.perf_counter() - t0
    n_w, wins = timed_windows(D, lambda i: e2e_step(args.warmup + i), est)
    e2e_s = float(np.median(wins))
    nfd = d.nfd
    words = d.n_queries * (d.n_points * d.dof + 3 * nfd + 2)               # point matrices + initial state: in the kernel's parameter block
    h2d = words * 4 if words <= 256 else d.n_queries * (d.n_points * d.dof + 2 * nfd) * 4   # ... or through the pinned staging ring
    d2h = d.n_queries * (d.n_cand * (4 * 4 + 1) + 5 * 4 + nfd * 4 + 4)
    e2e = dict(value=pts_per_step * D.world / e2e_s, unit=UNIT, h2d_bytes_per_step=h2d, d2h_bytes_per_step=d2h,
               steps_per_window=n_w, windows=len(wins), spread=float((max(wins) - min(wins)) / e2e_s), us_per_step=e2e_s * 1e6,
               api="sspp_b200.capi (ctypes over the C ABI): ssppcu_plan_step (start / end / noise scale from pageable numpy arrays; they ride in "
                   "the kernel's parameter block when they fit, else through the pinned staging ring) + ssppcu_get_results + ssppcu_get_best; the "
                   "result block reaches its mapped pinned mirror from inside the kernel (cooperative shapes) or by one D2H copy and is copied "
                   "out to the caller's arrays; median of the windows, max over ranks per window")
    # ---- e2e_plan: the drop-in Python call of the reference's API (JOINT shapes) ----
    e2e_plan = None
    if joint and case.cfg.dof in (3, 6, 7, 9):
        from sspp_b200 import _sspp
        planner = getattr(_sspp, f"SamplingPathPlanner{case.cfg.dof}")(str(case.mjcf), device=D.local)
        planner.set_verbose(False)
        planner.set_seed(SEED)
        st, en, lim = case.start[0], case.end[0], case.limits
        N, M, K = case.cfg.n_samples, case.cfg.check_points, case.cfg.n_points
        n_succ = 0

        def plan_step(i: int):
            nonlocal n_succ
            ok, paths = planner.plan(st, en, case.joint_sigma, lim, N, M, K)     # (bool, list[SplineD]) as src/sspp_bindings.cpp:43-50
            n_succ = len(paths)
        for i in range(3):
            plan_step(i)
        t0 = time.perf_counter(); plan_step(0); est = time.perf_counter() - t0
        n_p, pw = timed_windows(D, plan_step, est)
        ps = float(np.median(pw))
        e2e_plan = dict(value=N * M * D.world / ps, unit=UNIT, us_per_call=ps * 1e6, steps_per_window=n_p, windows=len(pw),
                        spread=float((max(pw) - min(pw)) / ps), splines_returned=n_succ,
                        h2d_bytes_per_step=h2d, d2h_bytes_per_step=d2h + 4 + n_succ * (nfd * 4 + 4),
                        api=f"sspp_b200._sspp.SamplingPathPlanner{case.cfg.dof}(xml).plan(start, end, sigma, limits, {N}, {M}, {K}) -> "
                            "(bool, list[Spline]) with the list materialised (classic Planner A: no CES update); only the collision-free "
                            "candidates' rows cross PCIe (compaction kernel into pinned memory), GIL released around launch + wait")
    clk = clocks.stop()
    if D.rank != 0:
        h.close()
        return None
    roof = roofline_record(h, case, args, D, pts_per_step, k_ms, k_ms_isolated, inner, workload)
    line = dict(metric=METRIC, value=value, unit=UNIT, n_gpus=D.world, steps=args.steps, warmup=args.warmup,
                ms_per_step=ms_total / args.steps, higher_is_better=True, scaling="weak", vs_baseline=None, dtype="f32",
                data="synthetic", config=dict(desc, l2="no flush: a step reads <64 KiB of constants and writes ~100 KiB of results; compute bound"
                                             if workload != "cfg3" else "no flush: the SDF grid (65.5 MB of packed cells) is meant to stay L2 resident",
                                             parallelism=f"independent queries x{D.world}", threads=d.threads,
                                             cand_per_cta=d.cand_per_cta, ctas_per_query=d.ctas_per_query, smem_bytes=d.smem_bytes),
                clocks=clk, e2e=e2e, gpu_launches=launches_per_step * args.steps, roofline=roof)
    if e2e_plan:
        line["e2e_plan"] = e2e_plan
    if workload == "cfg1":
        line["kat1"] = kat1()
    line["last_step"] = dict(n_success=int(res["n_success"][0]), n_elite=int(res["n_elite"][0]), best_cost=float(res["best_cost"][0]))
    h.close()
    return line


def popshard_record(args, D: Dist, per_rank: int = 8192, iters: int = 50) -> dict | None:
    """cfg3 over the ranks: ONE SDF population of per_rank * N candidates through the fused key exchange (product) and through
    NCCL all-gather + merge kernels (baseline), both checked bit for bit against the unsharded handle, with phase stamps"""
    torch, dist = D.torch, D.dist
    from sspp_b200 import capi
    from sspp_b200.sharding import DeviceShard, PopulationExchange, PopulationSharded, block_range
    from tests.helpers import sdf_case
    N = per_rank * D.world

    def mk(n, off=0, glob=0):
        c = sdf_case(n_samples=n, check_points=64, n_points=5, dims=(128, 128, 128))
        c.cfg.device = D.local
        c.cfg.shard_offset, c.cfg.shard_global = off, glob
        h = capi.Handle(c.model, c.cfg); c.prepare(h)
        return h
    lo, hi = block_range(N, D.rank, D.world)
    full = mk(N - 2)

    def same_as_full(hs) -> bool:
        fi, fc, fw = full.elites(0); mi, mc, mw = hs.merged(N)
        m0, s0 = full.get_distribution(); m1, s1 = hs.get_distribution()
        r0, r1 = full.results(), hs.results()
        return bool(np.array_equal(mi[:len(fi)], fi) and np.array_equal(mw[:len(fw)].view(np.uint32), fw.view(np.uint32)) and np.array_equal(m0, m1)
                    and np.array_equal(s0, s1) and r0["n_success"][0] == r1["n_success"][0] and r0["best_idx"][0] == r1["best_idx"][0])

    def time_loop(fn, stream):
        ev = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
        D.barrier()
        with torch.cuda.stream(stream):
            ev[0].record()
            for it in range(3, 3 + iters):
                fn(it)
            ev[1].record()
        D.barrier()
        return D.max_over_ranks(ev[0].elapsed_time(ev[1]) / iters)
    # baseline transport
    ps = PopulationSharded(DeviceShard(mk(hi - lo, lo, N), torch.cuda.Stream()))
    ok_nccl = True
    for it in range(3):
        ps.iterate(SEED, it); full.iterate(SEED, it, 1); torch.cuda.synchronize()
        ok_nccl &= same_as_full(ps.b.h)
    ms_nccl = time_loop(lambda it: ps.iterate(SEED, it), ps.b.stream)
    ps.b.h.close()
    # product: exchange fused into the kernel
    full.close(); full = mk(N - 2)
    px = PopulationExchange(mk(hi - lo, lo, N), torch.cuda.Stream())
    full.set_stream(px.stream.cuda_stream)
    ok_fused = True
    for it in range(3):
        px.iterate(SEED, it); full.iterate(SEED, it, 1); px.check(); torch.cuda.synchronize()
        ok_fused &= same_as_full(px.h)
    ms_fused = time_loop(lambda it: px.iterate(SEED, it), px.stream)
    px.check()
    px.h.phase_probe(True)
    tl = []
    for it in range(60, 68):
        D.barrier(); px.iterate(SEED, it); px.check()
        st = px.h.phase_probe(True, read=True).astype(np.int64)
        tl.append((st - st[:, 0].min()) * 1e-3)
    px.h.phase_probe(False)
    A = np.stack(tl[2:])
    names = {1: "scored", 3: "shard_ranked", 9: "published", 10: "flags_seen", 11: "merged", 5: "finished"}
    phases = {n: [round(float(A[:, :, i].min(axis=1).mean()), 1), round(float(np.median(A[:, :, i], axis=1).mean()), 1),
                  round(float(A[:, :, i].max(axis=1).mean()), 1)] for i, n in names.items()}
    ms_one = time_loop(lambda it: full.iterate(SEED, it, 1), px.stream)       # every rank times its own unsharded copy; max over ranks
    ok_nccl, ok_fused = D.all_true(ok_nccl), D.all_true(ok_fused)
    px.h.close(); full.close()
    if D.rank != 0:
        return None
    return dict(workload=f"cfg3 population sharded: ONE query, {N} candidates x 64 points ({per_rank} per rank), cluttered stacking scene, SDF grid 128^3",
                candidates=N, ranks=D.world, ms_per_iter=ms_fused, points_per_s=N * 64 / (ms_fused * 1e-3),
                ms_per_iter_nccl_allgather=ms_nccl, ms_per_iter_1gpu=ms_one, speedup_vs_1gpu=ms_one / ms_fused,
                strong_scaling_efficiency=ms_one / ms_fused / D.world,
                bit_identical_to_unsharded=bool(ok_fused), bit_identical_to_unsharded_nccl=bool(ok_nccl),
                phases_us_min_med_max_rank0=phases, iterations_timed=iters,
                bound="the slowest CTA of the scoring phase on the slowest rank (tile cost imbalance), then merge + regenerate; the key "
                      "exchange itself is the gap published -> flags_seen")


def run_cfg3_sharded(args, D: Dist) -> dict | None:
    rec = popshard_record(args, D, iters=max(10, args.steps))
    clk = ClockSampler(D.local).stop()
    if rec is None:
        return None
    return dict(metric=METRIC, value=rec["points_per_s"], unit=UNIT, n_gpus=D.world, steps=rec["iterations_timed"], warmup=3,
                ms_per_step=rec["ms_per_iter"], higher_is_better=True, scaling="weak", vs_baseline=None, dtype="f32", data="synthetic",
                config=dict(workload=rec["workload"], parallelism=f"population sharded x{D.world} (8192 candidates per rank), key exchange fused into "
                            "the kernel over NVLink peer memory", l2="no flush: SDF grid meant to stay L2 resident"),
                clocks=clk, gpu_launches=rec["iterations_timed"], popshard=rec,
                e2e=dict(value=rec["points_per_s"], unit=UNIT, h2d_bytes_per_step=0, d2h_bytes_per_step=0,
                         note="device-resident loop (the sharded state never leaves the GPUs between iterations)"))


def run_cfg5(args, D: Dist) -> dict | None:
    torch, dist = D.torch, D.dist
    from sspp_b200 import cost_matrix as cm
    from tests.helpers import task_case
    W, BATCH, ITERS = 256, 1020, 60
    case = task_case(n_queries=BATCH)
    case.cfg.device = D.local
    planner = cm.HandlePlanner(case.model, case.cfg, case.sigma0, ITERS)
    wp = cm.sample_waypoints(W)
    dev = torch.device("cuda", D.local)
    clocks = ClockSampler(D.local)

    def build():
        return cm.cost_matrix(planner, wp, BATCH, D.rank, D.world, device=dev if D.world > 1 else None)
    m = build()
    reps = max(3, min(args.steps, 10))
    times = []
    for _ in range(reps):
        D.barrier()
        t0 = time.perf_counter()
        m = build()
        D.barrier()
        times.append(D.max_over_ranks(time.perf_counter() - t0))
    clk = clocks.stop()
    n_pairs = W * (W - 1) // 2
    s = float(np.median(times))
    d = planner.h.dims
    pts = n_pairs * ITERS * d.n_cand * case.cfg.check_points
    launches = planner.launches // (reps + 1)
    planner.close()
    if D.rank != 0:
        return None
    off = m[np.triu_indices(W, 1)]
    return dict(metric=METRIC, value=pts / s, unit=UNIT, n_gpus=D.world, steps=reps, warmup=1, ms_per_step=s * 1e3, higher_is_better=True,
                scaling="strong", vs_baseline=None, dtype="f32", data="synthetic",
                config=dict(workload=f"cfg5: {W} waypoints, {n_pairs} Planner-B plans x {ITERS} iterations x {d.n_cand} candidates x "
                                     f"{case.cfg.check_points} collocation points, robocrane.xml gripper, batches of {BATCH} pairs; a step = the whole matrix, "
                                     "host waypoints in -> matrix on every rank's host",
                            parallelism=f"pair batches x{D.world}, one all-gather of the costs at the end"),
                clocks=clk, gpu_launches=launches * reps, plans_per_s=n_pairs / s, seconds_per_matrix=s, spread=float((max(times) - min(times)) / s),
                finite_entries=int(np.isfinite(off).sum()), symmetric=bool(np.array_equal(m, m.T)), matrix_crc=zlib.crc32(m.tobytes()),
                e2e=dict(value=pts / s, unit=UNIT, h2d_bytes_per_step=int(n_pairs * 2 * 4 * 8), d2h_bytes_per_step=int(n_pairs * 4),
                         note="the timed region IS end to end: host waypoints -> host matrix"))


def run_ours(args, rank: int, world: int, local_rank: int):
    import torch
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the scoring path has no CPU fallback (use --impl reference for the CPU arm)")
    D = Dist(rank, world, local_rank)
    if args.workload == "cfg5":
        line = run_cfg5(args, D)
    elif args.workload == "cfg3" and world > 1:
        line = run_cfg3_sharded(args, D)
    else:
        line = run_handle_workload(args, D, args.workload)
        if world > 1 and args.workload == "cfg2" and not args.no_popshard:
            rec = popshard_record(args, D)
            if line is not None:
                line["popshard"] = rec
    if line is not None:
        if world == 1 and not args.no_cpu_baseline and args.workload != "cfg5":
            case, _ = workload_case(args.workload, 0)
            line["cpu_baseline"] = cpu_baseline(case, args.workload)
        emit(line)
    if world > 1:
        D.dist.destroy_process_group()


_JSON_FD = 1


def emit(line: dict):
    """the ONE JSON line, on the process's original stdout"""
    os.write(_JSON_FD, (json.dumps(line) + "\n").encode())


def main():
    global _JSON_FD
    args = parse_args()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if world == 1 and args.gpus > 1:
        # launched without torchrun: re-exec under it, one rank per GPU
        cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={args.gpus}",
               "--master-addr", "127.0.0.1", "--master-port", os.environ.get("MASTER_PORT", "29531"), __file__] + sys.argv[1:]
        raise SystemExit(subprocess.call(cmd))
    # stdout carries the JSON line and nothing else: libraries that print there (NCCL writes its version line to stdout
    # when the box exports NCCL_DEBUG=VERSION; the reference's plan() prints a summary line) are sent to stderr
    sys.stdout.flush()
    _JSON_FD = os.dup(1)
    os.dup2(2, 1)
    if args.impl == "reference":
        run_reference(args, rank, world)
        return
    run_ours(args, rank, world, local_rank)


if __name__ == "__main__":
    main()
