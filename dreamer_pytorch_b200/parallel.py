"""Data parallelism of the hot path.

The imagination start states (and the sequences they come from) are independent, so rank g of G owns
a contiguous shard of them and runs the unchanged single-GPU kernels; nothing crosses ranks on the data
path.  The only exchange is the gradient all-reduce of each optimizer group -- one flat fp32 bucket per
group (actor ~0.5 MB, value ~0.5 MB, world model 3-21 MB), averaged over ranks (equal shards: the mean of
the shard means is the global mean), over NCCL / NVLink on the GPU box and gloo in the CPU tests.
The reference has no distributed path (dreamer_algo.py:88-96 ignores world_size); this is new.
"""
from __future__ import annotations

from typing import Iterable, List, Sequence, Tuple

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
