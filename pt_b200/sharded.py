"""Site-sharded operation over several GPUs (BASELINE config 3).

One process per GPU (torch.distributed, NCCL over NVLink): every rank owns an
Engine that holds a contiguous, tile-aligned range of the alignment's sites.
CLV updates, P matrices and sumtables never leave their GPU.  The only exchange
is the one the reference's own (unused) multi-partition hooks foresee
(pll-modules parallel_reduce_cb, PLLMOD/optimize/pll_optimize.c:1211-1212 and
1249-1255): a sum-allreduce of per-tree scalars -- 1 double for an edge lnL,
[2B] doubles (d1, d2) per Newton step of B branches, [B] doubles for the
candidates' final lnL -- issued by the engine through its allreduce hook on the
engine's own stream, so the Newton loop stays on the device between steps.

Two ways to do that exchange:

* `exchange="peer"` (default on CUDA): the engines' own exchange -- every rank exports a
  handle to its mailbox (ptgpu_shard_export), the handles go round once through
  torch.distributed, every rank maps the others' (ptgpu_shard_connect), and from then on the
  KERNELS sum the scalars over NVLink peer memory themselves, in rank order.  No call leaves
  the engine between Newton steps, and the whole-tree step program
  (ptgpu_optimize_all_branches) runs on a shard as one persistent kernel per GPU.
* `exchange="hook"`: a torch.distributed.all_reduce per reduction through the engine's
  allreduce hook (NCCL on GPUs; gloo with the CPU test-suite's stand-in engine).

This module is plumbing: ranges, the exchange set-up, and slicing of tip sequences.  All
arithmetic is in the CUDA engine; there is no CPU path (the `engine_factory`
argument exists so the CPU test-suite can exercise this plumbing over gloo with
a stand-in engine built from the test oracle).
"""
import ctypes as C

import numpy as np

from .engine import ALLREDUCE_FN, Engine

TILE_SITES = 256  # the engine's reduction tile (kernels.cuh kTileSites)


def shard_range(sites, world, rank, tile=TILE_SITES):
    """contiguous tile-aligned site range of `rank`: ceil(tiles/world) tiles each"""
    tiles = (sites + tile - 1) // tile
    per = (tiles + world - 1) // world
    begin = min(sites, rank * per * tile)
    end = min(sites, (rank + 1) * per * tile)
    return begin, end


class _DevBuf:
    """a raw device pointer as a __cuda_array_interface__ object (float64 vector)"""

    def __init__(self, ptr, count):
        self.__cuda_array_interface__ = dict(shape=(count,), typestr="<f8", data=(ptr, False),
                                             version=2)


class SiteShardedEngine:
    """Engine-compatible front-end whose reductions span all ranks of `group`."""

    def __init__(self, tips, clv_buffers, sites, rate_cats, prob_matrices, scale_buffers,
                 group=None, device=None, engine_factory=Engine, exchange=None):
        import torch
        import torch.distributed as dist
        self._torch, self._dist, self.group = torch, dist, group
        self.rank, self.world = dist.get_rank(group), dist.get_world_size(group)
        self.global_sites = sites
        self.begin, self.end = shard_range(sites, self.world, self.rank)
        if self.end <= self.begin:
            raise ValueError(f"rank {self.rank} of {self.world} has no sites "
                             f"({sites} sites): use fewer ranks")
        self.sites = self.end - self.begin
        self.device = device
        kw = {} if device is None else dict(device=device)
        self.engine = engine_factory(tips, clv_buffers, self.sites, rate_cats, prob_matrices,
                                     scale_buffers, **kw)
        self.on_cuda = bool(getattr(self.engine, "on_cuda", True))
        self.engine.set_shard(self.begin, sites)
        self.exchange = exchange or ("peer" if (self.on_cuda and engine_factory is Engine) else "hook")
        self._hook = None
        self.allreduce_calls = 0
        self.allreduce_doubles = 0
        self._views = {}
        if self.exchange == "peer":
            handles = [None] * self.world
            dist.all_gather_object(handles, self.engine.shard_export(), group=group)
            self.engine.shard_connect(self.rank, handles)
            dist.barrier(group=group)   # nobody computes before everybody is connected
        else:
            self._hook = ALLREDUCE_FN(self._allreduce)
            self.engine.set_allreduce(self._hook)

    def exchanges(self):
        """scalar exchanges between the shards so far (peer mode: counted on the device)"""
        if self.exchange == "peer":
            return self.engine.shard_status()[2]
        return self.allreduce_calls

    # the hook the engine calls between its local reduction and the scalar update
    def _allreduce(self, ctx, ptr, count, stream):
        try:
            torch, dist = self._torch, self._dist
            self.allreduce_calls += 1
            self.allreduce_doubles += int(count)
            if self.on_cuda:
                # the engine hands over the same few buffers on the same stream call after call
                # (several thousand times per OptimizeAllBranches): the zero-copy tensor view and
                # the stream wrapper are made once per (pointer, count, stream)
                key = (ptr, int(count), stream)
                hit = self._views.get(key)
                if hit is None:
                    dev = torch.device("cuda", self.device if self.device is not None else
                                       torch.cuda.current_device())
                    hit = (torch.as_tensor(_DevBuf(ptr, int(count)), device=dev),
                           torch.cuda.ExternalStream(stream, device=dev))
                    self._views[key] = hit
                t, ext = hit
                with torch.cuda.stream(ext):
                    dist.all_reduce(t, op=dist.ReduceOp.SUM, group=self.group)
            else:
                a = np.ctypeslib.as_array(C.cast(ptr, C.POINTER(C.c_double)), shape=(int(count),))
                t = torch.from_numpy(a)
                dist.all_reduce(t, op=dist.ReduceOp.SUM, group=self.group)
            return 0
        except Exception as exc:  # an exception must not unwind through the C frame
            self._hook_error = exc
            return 1

    # ---- sharded inputs
    def set_tip_states(self, tip, seq):
        if len(seq) != self.global_sites:
            raise ValueError("sequence length does not match the alignment")
        self.engine.set_tip_states(tip, seq[self.begin:self.end])

    # ---- everything else is the engine's own call; reductions come back global
    def __getattr__(self, name):
        return getattr(self.engine, name)

    def gather_scaler(self, idx):
        """all shards' scaler counts of buffer `idx`, concatenated (test hook)"""
        torch, dist = self._torch, self._dist
        sizes = [shard_range(self.global_sites, self.world, r) for r in range(self.world)]
        width = max(e - b for b, e in sizes)
        mine = torch.zeros(width, dtype=torch.int64)
        mine[: self.sites] = torch.from_numpy(self.engine.get_scaler(idx).astype(np.int64))
        dev = torch.device("cuda", self.device) if (self.on_cuda and self.device is not None) else None
        outs = [torch.empty(width, dtype=torch.int64, device=dev) for _ in sizes]
        dist.all_gather(outs, mine.to(dev) if dev else mine, group=self.group)
        return torch.cat([o.cpu()[: e - b] for o, (b, e) in zip(outs, sizes)]).numpy().astype(np.uint32)

    def close(self):
        if self.exchange == "peer":
            # a peer's mailbox must outlive every kernel that may still write into it
            self.engine.synchronize()
            self._dist.barrier(group=self.group)
            self.engine.shard_disconnect()
        else:
            self.engine.set_allreduce(None)
        self.engine.close()
