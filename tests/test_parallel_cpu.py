"""Host-side logic of the data-parallel path on CPU: gloo, world_size 2."""
import os
import socket

import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from dreamer_pytorch_b200.parallel import allreduce_grads, shard_bounds


def test_shard_bounds_cover_everything_once():
    for n in (0, 1, 7, 2500, 65536):
        for world in (1, 2, 3, 8):
            spans = [shard_bounds(n, r, world) for r in range(world)]
            assert spans[0][0] == 0 and spans[-1][1] == n
            assert all(a[1] == b[0] for a, b in zip(spans, spans[1:]))
            sizes = [hi - lo for lo, hi in spans]
            assert max(sizes) - min(sizes) <= 1


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        torch.manual_seed(0)
        # a "loss" that is a mean over per-unit terms; each rank owns a shard of the units
        n = 10
        x = torch.randn(n, 4)
        w1 = torch.nn.Parameter(torch.randn(4, 3))
        w2 = torch.nn.Parameter(torch.randn(3))
        frozen = torch.nn.Parameter(torch.randn(2))          # never receives a gradient
        lo, hi = shard_bounds(n, rank, world)
        loss = ((x[lo:hi] @ w1) * w2).sum(-1).pow(2).mean()   # shard mean
        loss.backward()
        weight = (hi - lo) / n * world                        # unequal shards: weight by share of the batch
        allreduce_grads([[w1], [w2, frozen]], weight=weight)
        full = ((x @ w1.detach().requires_grad_(True)) * w2.detach()).sum(-1).pow(2).mean()
        if rank == 0:
            w1r = w1.detach().clone().requires_grad_(True)
            w2r = w2.detach().clone().requires_grad_(True)
            ((x @ w1r) * w2r).sum(-1).pow(2).mean().backward()
            out.put((float((w1.grad - w1r.grad).abs().max()), float((w2.grad - w2r.grad).abs().max()),
                     float(frozen.grad.abs().max())))
        del full
    finally:
        dist.destroy_process_group()


def test_gradient_allreduce_matches_single_process():
    world = 2
    ctx = mp.get_context("spawn")
    out = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, out)) for r in range(world)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(120)
        assert p.exitcode == 0
    e1, e2, ef = out.get(timeout=10)
    assert e1 < 1e-5 and e2 < 1e-5 and ef == 0.0
