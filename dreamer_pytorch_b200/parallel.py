"""Data parallelism of the hot path.

The imagination start states (and the sequences they come from) are independent, so rank g of G owns
a contiguous shard of them and runs the unchanged single-GPU kernels; nothing crosses ranks on the data
path.  The only exchange is the gradient all-reduce of each optimizer group -- one flat fp32 bucket per
group (actor ~0.5 MB, value ~0.5 MB, world model 3-21 MB), averaged over ranks (equal shards: the mean of
the shard means is the global mean), over NCCL / NVLink on the GPU box and gloo in the CPU tests.
The reference has no distributed path (dreamer_algo.py:88-96 ignores world_size); this is new.
"""
from __future__ import annotations

import ctypes as C
from typing import Iterable, List, Optional, Sequence, Tuple

import torch
import torch.distributed as dist


def shard_bounds(n: int, rank: int, world: int) -> Tuple[int, int]:
    """Contiguous shard [lo, hi) of n units for `rank`; the first n % world ranks get one more unit."""
    if world < 1 or not 0 <= rank < world:
        raise ValueError(f"bad rank/world {rank}/{world}")
    base, extra = divmod(n, world)
    lo = rank * base + min(rank, extra)
    return lo, lo + base + (1 if rank < extra else 0)


def shard_state(state, rank: int, world: int):
    """Shard every field of an RSSMState (or any tuple of [N, .] tensors) along dim 0."""
    n = state[0].shape[0]
    lo, hi = shard_bounds(n, rank, world)
    return type(state)(*(f[lo:hi] for f in state))


def flatten_grads(params: Sequence[torch.Tensor]) -> torch.Tensor:
    """One flat fp32 bucket holding the gradients of `params` (zeros where a gradient is missing)."""
    parts = [(p.grad if p.grad is not None else torch.zeros_like(p)).reshape(-1).float() for p in params]
    return torch.cat(parts) if parts else torch.zeros(0)


def unflatten_into_grads(flat: torch.Tensor, params: Sequence[torch.Tensor]) -> None:
    off = 0
    dst, src = [], []
    for p in params:
        k = p.numel()
        g = flat[off:off + k].view_as(p)
        if p.grad is None:
            p.grad = g.to(p.dtype).clone()
        else:
            dst.append(p.grad)
            src.append(g)
        off += k
    if dst:
        torch._foreach_copy_(dst, src)  # one multi-tensor launch instead of one copy per parameter


def allreduce_grads(groups: Iterable[Sequence[torch.Tensor]], weight: float = 1.0, async_op: bool = False,
                    fuse: bool = True) -> List:
    """Average the gradients of every parameter group over all ranks.

    `weight` is this rank's share of the global batch times the world size (1.0 for equal shards), so that
    unequal shards still give the global mean.  With ``fuse`` (default) all groups travel in ONE flat fp32 bucket
    (the actor and value buckets are ~0.5 MB each: latency-class messages, one collective beats two); with
    ``fuse=False`` there is one collective per group.  With ``async_op`` the collectives are issued back to back
    and ``finish_allreduce`` completes them (lets them overlap the model backward).
    """
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size() == 1:
        return []
    world = dist.get_world_size()
    groups = [[p for p in params] for params in groups]
    if fuse and not all(_shared_bucket(params) is not None for params in groups):
        groups = [[p for params in groups for p in params]]
    pending = []
    for params in groups:
        base = _shared_bucket(params)
        if base is not None and weight == 1.0:
            # the backward kernels wrote these gradients into ONE flat bucket (functional.flat_like): average it in place
            work = dist.all_reduce(base, op=dist.ReduceOp.AVG if dist.get_backend() == "nccl" else dist.ReduceOp.SUM,
                                   async_op=async_op)
            if dist.get_backend() != "nccl":
                if work is not None:
                    work.wait()
                    work = None
                base.div_(world)
            pending.append((work, None, params))
            continue
        flat = flatten_grads(params)
        flat.mul_(weight / world)
        work = dist.all_reduce(flat, op=dist.ReduceOp.SUM, async_op=async_op)
        pending.append((work, flat, params))
    if not async_op:
        finish_allreduce(pending)
        return []
    return pending


def _shared_bucket(params: Sequence[torch.Tensor]):
    """The flat buffer all gradients of `params` are views of, or None."""
    base = None
    for p in params:
        g = p.grad
        if g is None or g._base is None or g.dtype != torch.float32:
            return None
        if base is None:
            base = g._base
        elif g._base is not base:
            return None
    return base


def finish_allreduce(pending) -> None:
    for work, flat, params in pending:
        if work is not None:
            work.wait()
        if flat is not None:
            unflatten_into_grads(flat, params)


# ------------------------------------------------------------------------------------------------------------------
# One-shot all-reduce over NVLink peer memory (csrc/drm_allreduce.cu)
# ------------------------------------------------------------------------------------------------------------------
_PAD_WORDS = 4 * 64 * 8  # signal pad in front of the bucket: uint32 [<= 3 phases][<= 64 CTAs][8 ranks] (+ spare)


def _agree(ok: bool, device) -> bool:
    """True only if every rank says so (a path taken by some ranks only would hang the others)."""
    flag = torch.tensor([1 if ok else 0], dtype=torch.int32, device=device)
    dist.all_reduce(flag, op=dist.ReduceOp.MIN)
    return bool(flag.item())


class PeerBucket:
    """The gradient buckets of several parameter groups as slices of ONE flat fp32 buffer in symmetric (peer-mapped)
    memory, averaged over the ranks of one NVLink / NVSwitch box by a single launch of this library's own kernel
    (``drm_allreduce``: two flag hand-shakes + one pass of P2P loads) instead of one NCCL call per group.

        bucket = PeerBucket.create([actor_weights, value_weights])     # collective; None when peer mapping is unavailable
        with functional.bucket_provider(bucket.provide): ... backward ...   # kernels write the gradients into the bucket
        bucket.allreduce()                                              # in place, recordable into a CUDA graph

    ``create`` is all-or-nothing across ranks (every step is agreed with an all-reduce(MIN)) and ends with a self-test
    against NCCL, so a box without working P2P mapping falls back to NCCL on every rank."""

    def __init__(self):
        self.mem = self.flat = self.staging = self.epoch = self.call = None
        self.slices = {}

    @staticmethod
    def _key(tensors):
        return tuple(tuple(t.shape) for t in tensors)

    @classmethod
    def create(cls, groups: Sequence[Sequence[torch.Tensor]], ctas: int = 16, algo: int = 0) -> Optional["PeerBucket"]:
        """algo: 0 auto, 1 one-shot, 2 two-shot (drm_allreduce_call.algo)."""
        from . import _lib

        if not (dist.is_available() and dist.is_initialized()) or dist.get_backend() != "nccl":
            return None
        world, rank = dist.get_world_size(), dist.get_rank()
        dev = groups[0][0].device
        if world < 2 or world > _lib.MAX_RANKS:
            return None
        keys = [cls._key(g) for g in groups]
        sizes = [sum((t.numel() + 3) // 4 * 4 for t in g) for g in groups]
        self = cls()
        ok, symm_mem = len(set(keys)) == len(keys), None
        try:
            import torch.distributed._symmetric_memory as symm_mem

            enable = getattr(symm_mem, "enable_symm_mem_for_group", None)
            if enable is not None:
                try:
                    enable(dist.group.WORLD.group_name)
                except Exception:
                    pass
            self.mem = symm_mem.empty(_PAD_WORDS + sum(sizes), dtype=torch.float32, device=dev)
            self.mem.zero_()
        except Exception:
            ok = False
        if not _agree(ok, dev):
            return None
        try:
            hdl = symm_mem.rendezvous(self.mem, dist.group.WORLD)
            ptrs = [int(p) for p in hdl.buffer_ptrs]
            ok = len(ptrs) == world and all(ptrs)
        except Exception:
            ok = False
        if not _agree(ok, dev):
            return None
        torch.cuda.synchronize(dev)
        dist.barrier()  # every rank has zero-filled its signal pad before anyone signals
        self.handle = hdl
        # independent tensors over the same storage, NOT views of `mem`: code that scales or reduces "the bucket of a
        # group" through grads[0]._base must find that group's slice there, never the whole allocation with its signal pad
        def window(first: int, count: int) -> torch.Tensor:
            return torch.empty(0, dtype=torch.float32, device=dev).set_(self.mem.untyped_storage(),
                                                                        self.mem.storage_offset() + first, (count,))

        self.flat = window(_PAD_WORDS, sum(sizes))
        off = _PAD_WORDS
        for k, n in zip(keys, sizes):
            self.slices[k] = window(off, n)
            off += n
        self.staging = torch.empty_like(self.flat)
        self.epoch = torch.zeros(2, dtype=torch.int64, device=dev)
        c = _lib.drm_allreduce_call()
        c.world, c.rank, c.ctas, c.scale, c.numel = world, rank, int(ctas), 1.0 / world, self.flat.numel()
        c.algo = int(algo)
        for r in range(world):
            c.signal[r] = ptrs[r]
            c.buf[r] = ptrs[r] + 4 * _PAD_WORDS
        c.staging, c.epoch = self.staging.data_ptr(), self.epoch.data_ptr()
        self.call = c
        # self-test against NCCL: rank-dependent data, twice (the second call reuses flags and epoch)
        ok = True
        try:
            for it in range(2):
                g = torch.Generator(device=dev).manual_seed(1000 * it + rank)
                self.flat.copy_(torch.randn(self.flat.numel(), generator=g, device=dev))
                want = self.flat.clone()
                dist.all_reduce(want, op=dist.ReduceOp.AVG)
                self.allreduce()
                torch.cuda.synchronize(dev)
                err = float((self.flat - want).abs().max())
                ok = ok and err <= 1e-5 * max(1.0, float(want.abs().max()))
        except Exception:
            ok = False
        if not _agree(ok, dev):
            return None
        return self

    def provide(self, tensors) -> Optional[torch.Tensor]:
        return self.slices.get(self._key(tensors))

    def owns(self, group: Sequence[torch.Tensor]) -> bool:
        """True when every tensor of `group` is a view into this bucket's slice for it."""
        sl = self.slices.get(self._key(group))
        if sl is None:
            return False
        lo, hi = sl.data_ptr(), sl.data_ptr() + 4 * sl.numel()
        return all(lo <= g.data_ptr() and g.data_ptr() + 4 * g.numel() <= hi for g in group)

    def allreduce(self) -> None:
        """Mean over ranks of the whole bucket, in place, on the current stream (no host synchronisation)."""
        from . import _lib

        _lib.check(_lib.load().drm_allreduce(C.byref(self.call), C.c_void_p(torch.cuda.current_stream().cuda_stream)),
                   "drm_allreduce")
