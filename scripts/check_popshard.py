"""torchrun check on N GPUs: a population sharded over the ranks reaches the state of the unsharded population,
bit for bit, with both transports (all-gather + merge kernels; key exchange fused into the iteration kernel), and what
an iteration costs either way.  POP_CASE=sdf runs the cfg 3 scene (cluttered stacking scene, SDF grid 128^3) instead
of the robocrane task scene; POP_PER_RANK (default 8192) candidates per rank.
POP_SOAK=<iterations> (default 0) appends a soak of the fused exchange: that many iterations back to back, the state re-armed
every 50 iterations, bit identity with the unsharded handle verified every 1000 iterations (flag protocol / epoch wrap / halves)."""
import os, sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import numpy as np, torch, torch.distributed as dist
from sspp_b200 import capi
from sspp_b200.sharding import DeviceShard, PopulationExchange, PopulationSharded, block_range
from tests.helpers import sdf_case, task_case, SEED

rank, world, lr = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(lr)
dist.init_process_group("nccl", device_id=torch.device("cuda", lr))
N = int(os.environ.get("POP_PER_RANK", "8192")) * world
SDF = os.environ.get("POP_CASE", "task") == "sdf"
def mk(n, off=0, glob=0):
    c = sdf_case(n_samples=n, check_points=64, n_points=5, dims=(128, 128, 128)) if SDF else task_case(n_samples=n, check_points=64, n_points=5)
    c.cfg.device = lr
    c.cfg.shard_offset, c.cfg.shard_global = off, glob
    h = capi.Handle(c.model, c.cfg); c.prepare(h); return h
lo, hi = block_range(N, rank, world)
ps = PopulationSharded(DeviceShard(mk(hi - lo, lo, N), torch.cuda.Stream()))
full = mk(N - 2)
ok = True
for it in range(3):
    ps.iterate(SEED, it); full.iterate(SEED, it, 1)
    torch.cuda.synchronize()
    fi, fc, fw = full.elites(0); mi, mc, mw = ps.b.h.merged(N)
    m0, s0 = full.get_distribution(); m1, s1 = ps.b.h.get_distribution()
    ok &= bool(np.array_equal(mi, fi) and np.array_equal(m0, m1) and np.array_equal(s0, s1))
ev = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
torch.cuda.synchronize(); dist.barrier()
with torch.cuda.stream(ps.b.stream):
    ev[0].record()
    for it in range(3, 53): ps.iterate(SEED, it)
    ev[1].record()
torch.cuda.synchronize()
def report(what, good, ms):
    rows = [None] * world
    dist.all_gather_object(rows, (bool(good), float(ms)))
    if rank == 0:
        print(f"{what}: {'OK' if all(r[0] for r in rows) else 'FAIL ' + str([i for i, r in enumerate(rows) if not r[0]])} on {world} ranks, "
              f"{max(r[1] for r in rows):.3f} ms per iteration (max over ranks) for {N} candidates x 64 points", flush=True)
report("all-gather + merge", ok, ev[0].elapsed_time(ev[1]) / 50)

# the same population with the exchange fused into the kernel (peer memory, no collective in the loop)
px = PopulationExchange(mk(hi - lo, lo, N), torch.cuda.Stream())
full2 = mk(N - 2); full2.set_stream(px.stream.cuda_stream)
ok2 = True
for it in range(3):
    px.iterate(SEED, it); full2.iterate(SEED, it, 1)
    px.check(); torch.cuda.synchronize()
    fi, fc, fw = full2.elites(0); mi, mc, mw = px.h.merged(N)
    m0, s0 = full2.get_distribution(); m1, s1 = px.h.get_distribution()
    r0, r1 = full2.results(), px.h.results()
    ok2 &= bool(np.array_equal(mi[:len(fi)], fi) and np.array_equal(mw[:len(fw)], fw) and np.array_equal(m0, m1) and np.array_equal(s0, s1)
                and r0["n_success"][0] == r1["n_success"][0] and r0["best_idx"][0] == r1["best_idx"][0])
torch.cuda.synchronize(); dist.barrier()
with torch.cuda.stream(px.stream):
    ev[0].record()
    for it in range(3, 53): px.iterate(SEED, it)
    ev[1].record()
torch.cuda.synchronize(); px.check()
report("fused exchange", ok2, ev[0].elapsed_time(ev[1]) / 50)
px.h.phase_probe(True)
tl = []
for it in range(53, 63):
    dist.barrier(); px.iterate(SEED, it); px.check()
    st = px.h.phase_probe(True, read=True).astype(np.int64)
    tl.append((st - st[:, 0].min()) * 1e-3)
px.h.phase_probe(False)
A = np.stack(tl[2:])
names = {1: "scored", 3: "ranked", 8: "shard_ordered", 9: "published", 10: "flags_seen", 11: "merge_ranked", 5: "finished"}
if rank == 0:
    print("fused-exchange phases on rank 0 (us since launch start, min/median/max over CTAs): " +
          ", ".join(f"{n} {A[:, :, i].min(axis=1).mean():.1f}/{np.median(A[:, :, i], axis=1).mean():.1f}/{A[:, :, i].max(axis=1).mean():.1f}" for i, n in names.items()), flush=True)
SOAK = int(os.environ.get("POP_SOAK", "0"))
if SOAK:
    import time
    c0 = sdf_case(n_samples=8, check_points=64, n_points=5, dims=(128, 128, 128)) if SDF else task_case(n_samples=8, check_points=64, n_points=5)
    ok3, checked, t_s = True, 0, time.perf_counter()
    for it in range(SOAK):
        if it % 50 == 0:
            c0.prepare(px.h); c0.prepare(full2)
        px.iterate(SEED, 1000 + it); full2.iterate(SEED, 1000 + it, 1)
        if it % 1000 == 999 or it == SOAK - 1:
            px.check(); torch.cuda.synchronize()
            fi, fc, fw = full2.elites(0); mi, mc, mw = px.h.merged(N)
            m0, s0 = full2.get_distribution(); m1, s1 = px.h.get_distribution()
            ok3 &= bool(np.array_equal(mi[:len(fi)], fi) and np.array_equal(mw[:len(fw)], fw) and np.array_equal(m0, m1) and np.array_equal(s0, s1))
            checked += 1
    px.check(); torch.cuda.synchronize()
    rows = [None] * world
    dist.all_gather_object(rows, bool(ok3))
    if rank == 0:
        print(f"soak: {'OK' if all(rows) else 'FAIL ' + str([i for i, r in enumerate(rows) if not r])} on {world} ranks, {SOAK} fused-exchange iterations, "
              f"{checked} bit-identity checks against the unsharded handle, {time.perf_counter() - t_s:.1f} s, protocol {os.environ.get('SSPPCU_XCHG_PROTOCOL', '1')}", flush=True)
t0 = torch.cuda.Event(enable_timing=True); t1 = torch.cuda.Event(enable_timing=True)
with torch.cuda.stream(px.stream):
    t0.record()
    for it in range(3, 13): full2.iterate(SEED, it, 1)
    t1.record()
torch.cuda.synchronize()
if rank == 0: print(f"unsharded iteration on one GPU {t0.elapsed_time(t1) / 10:.3f} ms for {N} candidates x 64 points", flush=True)
dist.barrier()
dist.destroy_process_group()
