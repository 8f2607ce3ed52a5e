"""Optimizer tail of ``Dreamer.optimize_agent`` (dreamer_algo.py:182-194), SURVEY.md 8(f) N1.

The reference does, per optimizer group (world model, actor, value):

    grad_norm = torch.nn.utils.clip_grad_norm_(params, self.grad_clip)
    optimizer.step()                                   # torch.optim.Adam(params, lr=...)

``ClipAdam`` is that pair as ONE C-ABI call (two launches: sum of squares, clip + Adam update; no host sync).
Same arithmetic as torch's (clip_grad.py / adam.py: amsgrad=False, weight_decay=0, maximize=False), same in-place scaling
of ``.grad``, same ``state_dict`` layout as ``torch.optim.Adam`` (so the reference's checkpoints --
``actor_optimizer_dict`` etc., dreamer_algo.py:139-152 -- load), a ``param_groups`` list for schedulers.  There is no
CPU fallback: parameters must live on a CUDA device.

Duplicated parameters: the reference's world-model group lists the transition parameters twice (the
transition module is also a sub-module of the representation module; torch warns "duplicate parameters").  torch's
sequential semantics (CPU, ``foreach=False``) then count such a gradient twice in the norm, scale it by the clip
coefficient twice and run two Adam updates per step on it, its per-parameter ``step`` advancing by two -- that is what
this class does, in the same launch (torch's CUDA foreach path races on the duplicated entries and is not a definition).
At most 64 distinct tensors per group (the world-model group has 48).
"""
from __future__ import annotations

import ctypes as C
import math
from typing import Iterable, Optional

import torch

from . import _lib

MAX_TENSORS = 64


class ClipAdam:
    def __init__(self, params: Iterable[torch.nn.Parameter], lr: float = 1e-3, betas=(0.9, 0.999), eps: float = 1e-8,
                 grad_clip: Optional[float] = None, capturable: bool = False):
        """capturable: keep the update counter on the device (like torch.optim.Adam(capturable=True)) so that step() can be
        recorded into a CUDA graph; ``step_count`` then costs a device read."""
        self.listed = list(params)                      # as given (duplicates kept: they count in the norm)
        seen, uniq = {}, []
        for p in self.listed:
            if id(p) not in seen:
                seen[id(p)] = 0
                uniq.append(p)
            seen[id(p)] += 1
        self.params = uniq
        self.mult = [seen[id(p)] for p in uniq]
        if not self.params:
            raise ValueError("optimizer got an empty parameter list")
        if len(self.params) > MAX_TENSORS:
            raise NotImplementedError(f"more than {MAX_TENSORS} distinct tensors in one group: split the group")
        for p in self.params:
            if not p.is_cuda or p.dtype != torch.float32:
                raise RuntimeError("ClipAdam needs fp32 CUDA parameters (there is no CPU fallback)")
        self.grad_clip = grad_clip
        self.defaults = dict(lr=float(lr), betas=(float(betas[0]), float(betas[1])), eps=float(eps), weight_decay=0,
                             amsgrad=False, maximize=False, foreach=None, capturable=bool(capturable),
                             differentiable=False, fused=None)
        self.param_groups = [dict(self.defaults, params=self.listed)]
        self._steps = 0
        self.state = {p: dict(exp_avg=torch.zeros_like(p), exp_avg_sq=torch.zeros_like(p)) for p in self.params}
        dev = self.params[0].device
        self._ws = torch.zeros(4, dtype=torch.float32, device=dev)
        self.grad_norm = torch.zeros((), dtype=torch.float32, device=dev)
        self.step_dev = torch.zeros(1, dtype=torch.int32, device=dev) if capturable else None

    # lr / betas / eps live in param_groups[0] (what LR schedulers edit), like torch
    @property
    def lr(self) -> float:
        return float(self.param_groups[0]["lr"])

    @property
    def betas(self):
        return self.param_groups[0]["betas"]

    @property
    def eps(self) -> float:
        return float(self.param_groups[0]["eps"])

    @property
    def step_count(self) -> int:
        """Number of step() calls that updated parameters."""
        return int(self.step_dev.item()) if self.step_dev is not None else self._steps

    @step_count.setter
    def step_count(self, v: int):
        if self.step_dev is not None:
            self.step_dev.fill_(int(v))
        else:
            self._steps = int(v)

    def zero_grad(self, set_to_none: bool = True):
        for p in self.params:
            if set_to_none:
                p.grad = None
            elif p.grad is not None:
                p.grad.zero_()

    @torch.no_grad()
    def step(self) -> torch.Tensor:
        """clip_grad_norm_ (when grad_clip is set) + Adam.step.  Returns the total gradient norm before clipping as a
        device scalar (what clip_grad_norm_ returns); no host synchronisation."""
        lib = _lib.load()
        live = [(p, k) for p, k in zip(self.params, self.mult) if p.grad is not None]
        if not live:
            return self.grad_norm.zero_()
        for p, _ in live:  # validate before touching any state
            g = p.grad
            if g.dtype != torch.float32:
                raise RuntimeError("ClipAdam needs fp32 gradients")
            # the kernel walks raw storage: parameter, gradient and both moments must share one dense layout (contiguous,
            # or channels-last for the conv weights of precision.conv_mode "channels_last" / "bf16")
            if g.stride() != p.stride():
                p.grad = torch.empty_like(p).copy_(g)
            if not (p.is_contiguous() or (p.dim() == 4 and p.is_contiguous(memory_format=torch.channels_last))):
                raise RuntimeError("ClipAdam needs dense (contiguous or channels-last) parameters")
            st = self.state[p]
            if st["exp_avg"].stride() != p.stride():  # the parameter changed its memory format after construction
                st["exp_avg"] = torch.empty_like(p).copy_(st["exp_avg"])
                st["exp_avg_sq"] = torch.empty_like(p).copy_(st["exp_avg_sq"])
        b1, b2 = self.betas
        c = _lib.drm_clip_adam_call()
        c.n_tensors = len(live)
        c.max_norm = float(self.grad_clip) if self.grad_clip else 0.0
        c.beta1, c.beta2, c.eps = b1, b2, self.eps
        c.one_minus_beta1, c.one_minus_beta2 = 1.0 - b1, 1.0 - b2
        c.lr, c.beta1_d, c.beta2_d = self.lr, b1, b2
        if self.step_dev is not None:
            c.step_dev = self.step_dev.data_ptr()
        else:
            self._steps += 1
            c.update = self._steps
            c.step_size = self.lr / (1.0 - b1 ** self._steps)
            c.bias2_sqrt = math.sqrt(1.0 - b2 ** self._steps)
        for i, (p, k) in enumerate(live):
            st = self.state[p]
            c.param[i], c.grad[i] = p.data_ptr(), p.grad.data_ptr()
            c.exp_avg[i], c.exp_avg_sq[i] = st["exp_avg"].data_ptr(), st["exp_avg_sq"].data_ptr()
            c.numel[i] = p.numel()
            c.mult[i] = k
        c.norm_out = self.grad_norm.data_ptr()
        c.workspace, c.workspace_bytes = self._ws.data_ptr(), self._ws.numel() * 4
        stream = C.c_void_p(torch.cuda.current_stream().cuda_stream)
        _lib.check(lib.drm_clip_adam(C.byref(c), stream), "drm_clip_adam")
        return self.grad_norm

    # ---- torch.optim.Adam's state_dict layout (torch/optim/optimizer.py state_dict(): a parameter listed several times
    # is numbered by its LAST position -- the dict comprehension there lets later duplicates overwrite earlier ones)
    def _index(self):
        idx = {id(p): i for i, p in enumerate(self.listed)}
        return idx, [idx[id(p)] for p in self.listed]

    def state_dict(self):
        idx, order = self._index()
        n = self.step_count
        state = {}
        for p, k in zip(self.params, self.mult):
            st = self.state[p]
            state[idx[id(p)]] = dict(step=torch.tensor(float(n * k)), exp_avg=st["exp_avg"], exp_avg_sq=st["exp_avg_sq"])
        group = {k: v for k, v in self.param_groups[0].items() if k != "params"}
        group["params"] = order
        return dict(state=state if n > 0 else {}, param_groups=[group], grad_clip=self.grad_clip)

    def load_state_dict(self, sd):
        """Accepts a ``torch.optim.Adam`` state_dict over the same parameter list (or one written by this class)."""
        idx, _ = self._index()
        groups = sd["param_groups"]
        if len(groups) != 1 or len(groups[0]["params"]) != len(self.listed):
            raise ValueError("loaded state dict has a different parameter layout")
        for k in ("lr", "betas", "eps"):
            if k in groups[0]:
                self.param_groups[0][k] = tuple(groups[0][k]) if k == "betas" else groups[0][k]
        steps = 0
        for p, k in zip(self.params, self.mult):
            st = sd["state"].get(idx[id(p)])
            if st is None:
                st = sd["state"].get(str(idx[id(p)]))
            if st is None:
                self.state[p]["exp_avg"].zero_(); self.state[p]["exp_avg_sq"].zero_()
                continue
            self.state[p]["exp_avg"].copy_(st["exp_avg"])
            self.state[p]["exp_avg_sq"].copy_(st["exp_avg_sq"])
            steps = max(steps, int(round(float(st["step"]) / k)))
        self.step_count = steps
        if "grad_clip" in sd:
            self.grad_clip = sd["grad_clip"]
