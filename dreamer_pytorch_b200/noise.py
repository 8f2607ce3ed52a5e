"""Noise source of the drop-in modules.

In production the modules draw N(0,1) / Gumbel(0,1) tensors with ``torch.randn`` / ``torch.rand``
on the device (one launch per rollout instead of one per step).  For parity tests the same
tensors the reference was fed are injected by name, mirroring the reference's draw order
(SURVEY.md section 8(c)):

    with noise.injected(eps_prior=[...], eps_post=[...], noise_act=[...]): ...

Each name holds a list; every draw of that name pops the next tensor.
"""
from __future__ import annotations

import contextlib
from typing import Dict, List

import torch

_queues: Dict[str, List[torch.Tensor]] = {}


@contextlib.contextmanager
def injected(**named):
    global _queues
    old = _queues
    _queues = {k: (list(v) if isinstance(v, (list, tuple)) else [v]) for k, v in named.items()}
    try:
        yield
        left = {k: len(v) for k, v in _queues.items() if v}
        assert not left, f"unused injected noise: {left}"
    finally:
        _queues = old


def is_injected(*names) -> bool:
    """True if a tensor is queued for any of the names."""
    return any(_queues.get(n) for n in names)


def _pop(name, shape, device):
    q = _queues.get(name)
    if not q:
        return None
    t = q.pop(0)
    if tuple(t.shape) != tuple(shape):
        raise RuntimeError(f"injected noise '{name}' has shape {tuple(t.shape)}, expected {tuple(shape)}")
    return t.to(device=device, dtype=torch.float32).contiguous()


def normal(name, shape, device):
    t = _pop(name, shape, device)
    return t if t is not None else torch.randn(shape, dtype=torch.float32, device=device)


def gumbel(name, shape, device):
    t = _pop(name, shape, device)
    if t is not None:
        return t
    u = torch.rand(shape, dtype=torch.float32, device=device).clamp_(1e-20, 1.0)
    return -torch.log(-torch.log(u).clamp_min_(1e-20))
