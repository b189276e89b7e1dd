"""torchrun check on N GPUs: a population sharded over the ranks (NCCL all-gather of keys) reaches the
state of the unsharded population computed on rank 0's GPU.  Prints OK/FAIL per rank."""
import os, sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import numpy as np, torch, torch.distributed as dist
from sspp_b200 import capi
from sspp_b200.sharding import DeviceShard, PopulationSharded, block_range
from tests.helpers import task_case, SEED

rank, world, lr = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(lr)
dist.init_process_group("nccl", device_id=torch.device("cuda", lr))
N = 8192 * world
def mk(n, off=0, glob=0):
    c = task_case(n_samples=n, check_points=64, n_points=5); c.cfg.device = lr
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
print(f"rank {rank}: {'OK' if ok else 'FAIL'}  sharded iteration {ev[0].elapsed_time(ev[1]) / 50:.3f} ms for {N} candidates x 64 points", flush=True)
dist.destroy_process_group()
