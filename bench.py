#!/usr/bin/env python
"""bench.py — scored collocation points/s of the fused candidate-scoring iteration on B200.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--workload cfg2|cfg4]
  (N > 1: python -m torch.distributed.run --nproc-per-node N ... bench.py --gpus N ...)

Workload at N = 1 (BASELINE.json configs[1]): robocrane 7-DoF fixture, one planning query,
4096 candidates x 64 collocation points x 10 control points.  One STEP = one planner iteration of
that query from its initial sampling state, ONE fused launch (ssppcu_iterate_restored): re-arm the
state from its snapshot -> Philox perturb -> spline -> kinematics -> collision -> cost reduce -> elites -> update.
The state is re-armed every step so that every step scores the same statistical workload (about half
of the candidates collide); Philox counters differ per step.
N > 1: independent planning queries shard across GPUs (no data-path collective): every rank runs
the workload on its own query ("scaling": "weak"); value = points of all ranks / max-over-ranks time.

Prints ONE JSON line (rank 0).  `value` is device-resident throughput; `e2e` goes through the
host API with host buffers (H2D of the query, D2H of every candidate's result) inside the timed region.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent
sys.path.insert(0, str(ROOT))

SEED = 0x53535050
METRIC = "scored collocation points/sec"
UNIT = "points/s"


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=20)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="cfg2", choices=["cfg2", "cfg4"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    return ap.parse_args()


def workload_case(name: str, rank: int = 0):
    """the synthetic inputs of SURVEY 8(d); rank selects this rank's independent query/queries"""
    from tests.helpers import joint_case, task_case, Q_START, Q_END
    rng = np.random.default_rng(1000 + rank)
    if name == "cfg2":
        case = joint_case(n_samples=4096, check_points=64, n_points=10, do_update=True)
        if rank > 0:   # another start/goal pair near the KAT configurations
            case.start = (Q_START + rng.normal(0, 0.05, 7))[None, :]
            case.end = (Q_END + rng.normal(0, 0.05, 7))[None, :]
        desc = dict(workload="cfg2: robocrane7 fixture, 1 query x 4096 candidates x 64 collocation points x 10 control points (cubic), "
                             "sigma 0.08*pi, CES update on, state re-armed every step",
                    candidates=4096, collocation_points=64, control_points=10, dof=7, queries_per_gpu=1)
        return case, desc
    Q = 1024
    s = np.column_stack([rng.uniform(0.2, 0.6, Q), rng.uniform(-0.4, 0.4, Q), rng.uniform(0.14, 0.4, Q), rng.uniform(-1.5, 1.5, Q)])
    e = np.column_stack([rng.uniform(0.2, 0.6, Q), rng.uniform(-0.4, 0.4, Q), rng.uniform(0.14, 0.4, Q), rng.uniform(-1.5, 1.5, Q)])
    case = task_case(n_queries=Q, start=s, end=e)
    desc = dict(workload="cfg4: robocrane.xml gripper, 1024 ICRA-style queries x 17 candidates x 40 collocation points, 1 via (quadratic)",
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
# CPU arm: the FP64 oracle O1 (restatement of the reference's CPU path) on the host cores.
# The reference itself needs MuJoCo + Eigen, absent here and on the GPU box -> kind "port".
# ------------------------------------------------------------------------------------------
class CpuArm:
    def __init__(self, case):
        from oracle import pyoracle            # the one place bench.py executes oracle/: as the CPU baseline
        from sspp_b200 import capi
        self.case = case
        self.o1 = pyoracle.O1(case.model, case.cfg.mode, case.cfg.dof, case.body)
        self.oc = pyoracle.o1_cfg_from(case.cfg)
        c = case.cfg
        task = c.mode == 1
        self.nfree = c.n_points - 2 if task else c.n_points - 2 * c.degree
        fb = 1 if task else c.degree
        self.base = [capi.base_points(c, case.start[q], case.end[q]).astype(np.float64) for q in range(len(case.start))]
        self.mean = [b[fb:fb + self.nfree].copy() for b in self.base]
        if task:
            self.sigma = [np.full_like(m, case.sigma0) for m in self.mean]
        else:
            self.sigma = [np.tile(case.joint_sigma * case.limits, (self.nfree, 1)) for _ in self.mean]
        # all the host threads the process may use: torchrun exports OMP_NUM_THREADS=1 to its workers, which would
        # leave the CPU arm single-threaded; the OpenMP runtime the oracle is linked to is told explicitly
        self.cores = len(os.sched_getaffinity(0))
        try:
            import ctypes
            gomp = ctypes.CDLL("libgomp.so.1")
            gomp.omp_set_num_threads(self.cores)
            self.cores = int(gomp.omp_get_max_threads())
        except OSError:
            self.cores = int(os.environ.get("OMP_NUM_THREADS", self.cores))
        self.M = c.check_points

    def step(self, it: int, n_cand: int, query: int = 0) -> int:
        """sample + score + elites + update of n_cand candidates; returns scored points"""
        o1, oc = self.o1, self.oc
        p = o1.sample(oc, SEED, it, query, 0, n_cand, self.mean[query], self.sigma[query])
        s = o1.score(oc, self.base[query], p)
        idx, w, ns = o1.elites(s["cost"], s["status"], self.case.cfg.elite_fraction)
        o1.update(oc, self.mean[query], self.sigma[query], p[idx], w)     # result discarded: state re-armed every step
        return n_cand * self.M

    def plan(self, query: int, iters: int, n_cand: int) -> int:
        """one whole plan of a query: `iters` iterations on its own distribution (cold start), the candidates of an
        iteration in parallel inside the oracle - how the reference's benchmark drivers spend their time
        (main_icra_benchmark.cpp:186-204: queries one after another, OpenMP inside plan())"""
        o1, oc = self.o1, self.oc
        mean, sigma = self.mean[query].copy(), self.sigma[query].copy()
        for it in range(iters):
            p = o1.sample(oc, SEED, it, query, 0, n_cand, mean, sigma)
            s = o1.score(oc, self.base[query], p)
            idx, w, ns = o1.elites(s["cost"], s["status"], self.case.cfg.elite_fraction)
            if ns > 0:
                m, sg = o1.update(oc, mean, sigma, p[idx], w)
                mean, sigma = m.reshape(mean.shape), sg.reshape(sigma.shape)
        return iters * n_cand * self.M

    def calibrate(self) -> float:
        n = 8
        t0 = time.perf_counter(); self.step(0, n); dt = time.perf_counter() - t0
        while dt < 0.3 and n < 4096:
            n *= 4
            t0 = time.perf_counter(); self.step(0, n); dt = time.perf_counter() - t0
        return n * self.M / dt


def cpu_baseline(case, budget_s: float = 15.0) -> dict:
    arm = CpuArm(case)
    rate = arm.calibrate()
    n = int(max(8, min(case.cfg.n_samples, rate * budget_s / arm.M)))
    t0 = time.perf_counter(); pts = arm.step(1, n); dt = time.perf_counter() - t0
    return dict(value=pts / dt, unit=UNIT, cores=arm.cores, kind="port",
                sample=f"oracle O1 (FP64 restatement of the reference CPU path; the reference needs MuJoCo+Eigen, absent): "
                       f"{n} of {case.cfg.n_samples} candidates x {arm.M} points of the same query, {dt:.1f} s, OpenMP {arm.cores} threads")


def run_reference(args, rank: int, world: int):
    """--impl reference: the CPU path on the host cores, bounded sample per step."""
    if rank != 0:
        return
    case, desc = workload_case(args.workload, 0)
    arm = CpuArm(case)
    rate = arm.calibrate()
    total = args.steps + args.warmup
    if args.workload == "cfg4":
        # a step of our arm is a whole 60-iteration plan of every query: the CPU step is the same plan for a bounded
        # number of the queries, one query after another
        iters, nc, Q = 60, case.cfg.n_samples, len(case.start)
        t0 = time.perf_counter(); arm.plan(0, iters, nc); per_plan = time.perf_counter() - t0
        nq = int(max(1, min(Q, 75.0 / max(total, 1) / max(per_plan, 1e-6))))
        for i in range(args.warmup):
            for q in range(nq):
                arm.plan(q, iters, nc)
        t0 = time.perf_counter()
        pts = 0
        for i in range(args.steps):
            for q in range(nq):
                pts += arm.plan((i * nq + q) % Q, iters, nc)
        dt = time.perf_counter() - t0
        val = pts / dt
        sample = (f"each step = {nq} of {Q} queries x {iters} iterations x {nc} candidates x {arm.M} points (cold-start plan: sample + spline + "
                  f"placement + collision + cost + elites + update), queries in sequence, oracle O1 port, OpenMP {arm.cores} threads; "
                  f"{dt / max(args.steps * nq, 1) * 1e3:.2f} ms per plan")
        n = nc
    else:
        n = int(max(1, min(case.cfg.n_samples, rate * 75.0 / max(total, 1) / arm.M)))
        for i in range(args.warmup):
            arm.step(i, n)
        t0 = time.perf_counter()
        pts = 0
        for i in range(args.steps):
            pts += arm.step(args.warmup + i, n)
        dt = time.perf_counter() - t0
        val = pts / dt
        sample = (f"each step = {n} of {case.cfg.n_samples} candidates x {arm.M} points (sample + spline + FK + collision + cost + elites + update), "
                  f"oracle O1 port, OpenMP {arm.cores} threads")
    line = dict(impl="reference", metric=METRIC, value=val, unit=UNIT, n_gpus=args.gpus, steps=args.steps, warmup=args.warmup,
                ms_per_step=dt / max(args.steps, 1) * 1e3, higher_is_better=True, scaling="weak", vs_baseline=None, dtype="f64",
                data="synthetic", config=desc,
                cpu_baseline=dict(value=val, unit=UNIT, cores=arm.cores, kind="port", sample=sample),
                e2e=dict(value=val, unit=UNIT, h2d_bytes_per_step=0, d2h_bytes_per_step=0))
    emit(line)


# ------------------------------------------------------------------------------------------
# GPU arm
# ------------------------------------------------------------------------------------------
def run_ours(args, rank: int, world: int, local_rank: int):
    import torch
    import torch.distributed as dist
    from sspp_b200 import capi, roofline

    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the scoring path has no CPU fallback (use --impl reference for the CPU arm)")
    torch.cuda.set_device(local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(x: float) -> float:
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    case, desc = workload_case(args.workload, rank)
    case.cfg.device = local_rank
    h = capi.Handle(case.model, case.cfg)
    stream = torch.cuda.Stream()
    h.set_stream(stream.cuda_stream)
    case.prepare(h)
    h.save_state()
    d = h.dims
    cfg4 = args.workload == "cfg4"
    inner = 60 if cfg4 else 1                       # cfg4: a step is a whole 60-iteration plan of every query
    pts_per_step = d.n_queries * d.n_cand * case.cfg.check_points * inner
    launches_per_step = 1 if d.ctas_per_query == 1 else inner

    def step(i: int):
        h.iterate(SEED, i * inner, inner, restored=True)      # re-arm from the snapshot + iterate: ONE launch

    clocks = ClockSampler(local_rank)
    h.set_result_mirror(False)                       # value / roofline: results stay in HBM, nothing is written to the host
    for i in range(max(args.warmup, 3)):
        step(i)
    # ---- value: K steps, device resident, CUDA events on the launch stream ----
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    ev0.record(stream)
    for i in range(args.steps):
        step(args.warmup + i)
    ev1.record(stream)
    barrier()
    ms_local = ev0.elapsed_time(ev1)
    ms_total = max_over_ranks(ms_local)
    value = pts_per_step * world * args.steps / (ms_total * 1e-3)
    # ---- roofline pass: the fused kernel alone, one event pair per launch ----
    nk = min(args.steps, 100)
    evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(nk)]
    for i, (a, b) in enumerate(evs):
        a.record(stream); h.iterate(SEED, (args.warmup + i) * inner, inner, restored=True); b.record(stream)
    torch.cuda.synchronize()
    k_ms_isolated = float(np.mean([a.elapsed_time(b) for a, b in evs]))      # one event pair per launch (adds the event gaps)
    k_ms = ms_local / args.steps      # the kernel's time per step over the timed region (events on the launch stream; a step is this kernel only:
                                      # `launches_per_step` launches of it, and work per step is counted the same way)
    # ---- e2e: through the host API, host buffers in, every candidate's result out ----
    n_e2e = min(args.steps, 200)
    h.set_result_mirror(True)                        # the library's default: the kernel fills the pinned result mirror itself
    res, best = h.results(), h.best()                # caller-owned host buffers, refilled by every step
    for i in range(max(args.warmup, 3)):             # untimed: the mirroring instantiation of the kernel is loaded by its first launch
        case.prepare(h)
        h.iterate(SEED, i * inner, inner)
        res = h.results(res)
        best = h.best(best)
    barrier()
    t0 = time.perf_counter()
    for i in range(n_e2e):
        case.prepare(h)                              # H2D: start/end -> point matrix, mean, sigma
        h.iterate(SEED, (args.warmup + i) * inner, inner)
        res = h.results(res)                         # D2H: L, Cnf, Cwf, cost, status per candidate + per-query scalars
        best = h.best(best)
    barrier()
    e2e_s = max_over_ranks(time.perf_counter() - t0)
    clk = clocks.stop()
    nfd = d.nfd
    h2d = d.n_queries * (d.n_points * d.dof + 2 * nfd) * 4
    d2h = d.n_queries * (d.n_cand * (4 * 4 + 1) + 5 * 4 + nfd * 4 + 4)
    e2e = dict(value=pts_per_step * world * n_e2e / e2e_s, unit=UNIT, h2d_bytes_per_step=h2d, d2h_bytes_per_step=d2h,
               steps=n_e2e, api="sspp_b200.capi (ctypes over the C ABI): set_queries + set_joint_noise/reset_distribution + iterate + get_results + get_best; "
                   "host arrays are pageable numpy buffers, the library stages inputs through its pinned ring; the result block reaches its "
                   "mapped pinned mirror from inside the kernel (cooperative shapes) or by one D2H copy, and is copied out to the caller's arrays")
    if rank != 0:
        h.close()
        return
    # ---- roofline (FP32 issue bound) ----
    wm = roofline.work_model(case.model, case.cfg.mode, case.cfg.dof, case.cfg.n_points, case.body)
    fp32_peak, _ = capi.fp32_peak(local_rank)
    peaks = {}
    pk = ROOT / "MEASURED_PEAKS.json"
    if pk.exists():
        peaks = json.loads(pk.read_text())
    ach = pts_per_step * wm["flop_per_point"] / (k_ms * 1e-3) * 1e-12
    hbm_bytes = d.n_queries * d.n_cand * inner * roofline.BYTES_PER_CANDIDATE
    hbm_ach = hbm_bytes / (k_ms * 1e-3) * 1e-9
    traffic = None                                  # DRAM bytes per launch of the kernel, from the committed ncu --set full capture
    tf = ROOT / "profiles" / "roofline_traffic.json"
    if tf.exists():
        traffic = json.loads(tf.read_text()).get(args.workload, {}).get("dram_bytes_per_launch")
    roof = dict(bound="fp32", achieved=ach, peak=fp32_peak, unit="TFLOP/s", frac=ach / fp32_peak, traffic=traffic,
                kernel="sspp_iterate_kernel", kernel_ms=k_ms, kernel_ms_isolated=k_ms_isolated, flop_per_point=wm["flop_per_point"],
                note="nominal FLOP model of sspp_b200/roofline.py (every filtered pair charged its full narrow phase; the kernel culls "
                     "most by bounding spheres) over the FMA peak measured live by ssppcu_fp32_peak; MEASURED_PEAKS.json has no FP32 entry",
                hbm=dict(achieved=hbm_ach, peak=peaks.get("hbm_gbs"), unit="GB/s",
                         frac=(hbm_ach / peaks["hbm_gbs"]) if peaks.get("hbm_gbs") else None,
                         peak_source="MEASURED_PEAKS.json hbm_gbs (of measured)" if peaks else "absent"))
    line = dict(metric=METRIC, value=value, unit=UNIT, n_gpus=world, steps=args.steps, warmup=args.warmup,
                ms_per_step=ms_total / args.steps, higher_is_better=True, scaling="weak", vs_baseline=None, dtype="f32",
                data="synthetic", config=dict(desc, l2="no flush: a step reads <64 KiB of constants and writes ~100 KiB of results; compute bound",
                                             parallelism=f"independent queries x{world}", threads=d.threads,
                                             cand_per_cta=d.cand_per_cta, ctas_per_query=d.ctas_per_query, smem_bytes=d.smem_bytes),
                clocks=clk, e2e=e2e, gpu_launches=launches_per_step * args.steps, roofline=roof)
    if world == 1 and not args.no_cpu_baseline:
        line["cpu_baseline"] = cpu_baseline(case)
    # elites of the last step as a sanity field
    line["last_step"] = dict(n_success=int(res["n_success"][0]), n_elite=int(res["n_elite"][0]), best_cost=float(res["best_cost"][0]))
    emit(line)
    h.close()


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
    # when the box exports NCCL_DEBUG=VERSION, and honours NCCL_DEBUG_FILE only above that level) are sent to stderr
    sys.stdout.flush()
    _JSON_FD = os.dup(1)
    os.dup2(2, 1)
    if args.impl == "reference":
        run_reference(args, rank, world)
        return
    run_ours(args, rank, world, local_rank)
    if world > 1:
        import torch.distributed as dist
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
