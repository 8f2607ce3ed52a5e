"""Shared helpers of the parity tests: golden cases, oracle runs, error metric."""
import os

import torch

from oracle import dreamer_oracle as O
from oracle.make_golden import CASES, case_inputs, model_kwargs  # noqa: F401

GOLDEN_DIR = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")

FP32_TOL = 1e-4  # BASELINE.json north_star: max relative (normwise) error on the fp32 path


def load_golden(name):
    return torch.load(os.path.join(GOLDEN_DIR, f"{name}.pt"), map_location="cpu", weights_only=False)


def rel_err(x, ref):
    return O.normwise_err(x, ref)


def oracle_loss(c, dtype=torch.float32, device="cpu"):
    """Oracle ``dreamer_loss`` on the case's inputs; returns (out dict, grads dict keyed by canonical names)."""
    w, obs, act, rew, done, noise = case_inputs(c)
    w = {k: v.to(device=device, dtype=dtype).requires_grad_(True) for k, v in w.items()}
    noise = {k: v.to(device=device, dtype=dtype) for k, v in noise.items()}
    out = O.dreamer_loss(w, obs.to(device), act.to(device=device, dtype=dtype), rew.to(device=device, dtype=dtype),
                         done.to(device), noise, horizon=c["H"], dist=c["dist"], use_pcont=c["use_pcont"])
    total = out["model_loss"] + out["actor_loss"] + out["value_loss"]
    names = [k for k in w if not k.startswith("observation_")]
    grads = torch.autograd.grad(total, [w[k] for k in names], allow_unused=True)
    return out, {k: g for k, g in zip(names, grads) if g is not None}, w, noise
