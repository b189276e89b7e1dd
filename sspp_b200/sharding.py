"""Multi-GPU sharding of the scoring path: one process per GPU, torch.distributed for the plumbing.

Two ways the path shards; neither exists in the reference (OpenMP only, sspp.h:203-216):

* independent planning QUERIES (ICRA batch, pairwise cost matrix): `block_range` gives every rank a
  contiguous block, each rank plans its block with its own handle, `gather_rows` collects per-query
  results.  No collective inside the iteration loop.
* one large candidate POPULATION: `PopulationSharded` — rank r scores candidates
  [r*N/G, (r+1)*N/G) of the same query (Philox counters use the global candidate index), then ONE
  all-gather of the shards' sorted 8-byte (cost, index) keys per iteration, after which every rank
  applies the identical merge + CES update (elites' rows are regenerated from the counters, so the
  sampled control points never travel) and all ranks hold bit-identical state without a second collective.

`PopulationExchange` is the same iteration with the exchange fused into the kernel: the ranks map each
other's exchange blocks once (CUDA IPC handles passed through torch.distributed), after which an
iteration is ONE launch per rank that scores, ranks, stores its keys into every peer's block over NVLink,
waits for the peers' flags and applies the update - no collective call in the loop.

The collective code is backend-agnostic (nccl on GPUs; gloo in the CPU tests with a checker backend).
"""
from __future__ import annotations

import numpy as np
import torch
import torch.distributed as dist


def block_range(n: int, rank: int, world: int) -> tuple[int, int]:
    """contiguous block [lo, hi) of n items for `rank`; sizes differ by at most one"""
    base, rem = divmod(n, world)
    lo = rank * base + min(rank, rem)
    return lo, lo + base + (1 if rank < rem else 0)


def gather_rows(local: np.ndarray, n_total: int, device=None, group=None) -> np.ndarray:
    """all-gather per-query rows that were split with block_range; returns [n_total, ...] on every rank"""
    world = dist.get_world_size(group) if dist.is_initialized() else 1
    if world == 1:
        return np.asarray(local)
    per = (n_total + world - 1) // world
    local = np.ascontiguousarray(local)
    pad = np.zeros((per,) + local.shape[1:], local.dtype)
    pad[: len(local)] = local
    t = torch.from_numpy(pad.view(np.uint8).reshape(per, -1))
    if device is not None:
        t = t.to(device)
    out = [torch.empty_like(t) for _ in range(world)]
    dist.all_gather(out, t, group=group)
    rows = []
    for r, o in enumerate(out):
        lo, hi = block_range(n_total, r, world)
        a = o.cpu().numpy().reshape(-1).view(local.dtype).reshape((per,) + local.shape[1:])
        rows.append(a[: hi - lo])
    return np.concatenate(rows, 0)


class DeviceShard:
    """backend of PopulationSharded on a GPU: a capi.Handle created with shard_offset / shard_global"""

    def __init__(self, handle, stream: "torch.cuda.Stream | None" = None):
        self.h = handle
        # the legacy default stream cannot be handed to the C ABI (NULL means "the handle's own stream")
        self.stream = stream if stream is not None and stream.cuda_stream != 0 else torch.cuda.Stream()
        handle.set_stream(self.stream.cuda_stream)
        ptr, n = handle.shard_keys()
        self.n_local = n

        class _Keys:      # zero-copy torch view of the handle's sorted key buffer (uint64 carried as int64)
            __cuda_array_interface__ = dict(shape=(n,), typestr="<i8", data=(ptr, False), version=2)
        self.keys = torch.as_tensor(_Keys(), device=f"cuda:{torch.cuda.current_device()}")

    def score(self, seed: int, it: int) -> torch.Tensor:
        self.h.iterate(seed, it, 1)
        return self.keys

    def merge(self, gathered: torch.Tensor, n_shards: int, seed: int, it: int):
        self.h.merge_update(gathered.data_ptr(), gathered.numel(), n_shards, seed, it)


class PopulationSharded:
    """one CES iteration of ONE query over a population sharded across the ranks of `group`"""

    def __init__(self, backend, group=None):
        self.b = backend
        self.group = group
        self.world = dist.get_world_size(group) if dist.is_initialized() else 1
        self.gathered = None

    def iterate(self, seed: int, it: int):
        keys = self.b.score(seed, it)
        if self.gathered is None:
            self.gathered = torch.empty(self.world * keys.numel(), dtype=keys.dtype, device=keys.device)
        if self.world == 1:
            self.gathered.copy_(keys)
        else:
            if isinstance(self.b, DeviceShard):
                with torch.cuda.stream(self.b.stream):        # the collective is ordered after the scoring launch
                    dist.all_gather_into_tensor(self.gathered, keys, group=self.group)
            else:
                dist.all_gather_into_tensor(self.gathered, keys, group=self.group)
        self.b.merge(self.gathered, self.world, seed, it)


class PopulationExchange:
    """PopulationSharded with the key exchange inside the iteration kernel (peer memory over NVLink).

    `handle` is a capi.Handle created with shard_offset / shard_global (equal shards).  The constructor is
    collective: every rank of `group` must call it.  world == 1 needs no process group."""

    def __init__(self, handle, stream: "torch.cuda.Stream | None" = None, group=None):
        self.h = handle
        self.group = group
        self.world = dist.get_world_size(group) if dist.is_initialized() else 1
        self.rank = dist.get_rank(group) if dist.is_initialized() else 0
        self.stream = stream if stream is not None and stream.cuda_stream != 0 else torch.cuda.Stream()
        handle.set_stream(self.stream.cuda_stream)
        _, ipc = handle.exchange_export(self.world, self.rank)
        if self.world == 1:
            handle.exchange_attach_ipc([ipc])
        else:
            handles = [None] * self.world
            dist.all_gather_object(handles, ipc, group=group)     # every rank has exported once this returns
            handle.exchange_attach_ipc(handles)
            dist.barrier(group=group)                             # nobody publishes before everybody is attached

    def iterate(self, seed: int, it: int):
        self.h.iterate_exchange(seed, it)

    def check(self):
        """synchronise and raise if a peer failed to publish (device-side timeout)"""
        self.h.exchange_status()
