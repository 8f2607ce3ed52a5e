"""Optimizer tail of ``Dreamer.optimize_agent`` (dreamer_algo.py:182-194), SURVEY.md 8(f) N1.

The reference does, per optimizer group (world model, actor, value):

    grad_norm = torch.nn.utils.clip_grad_norm_(params, self.grad_clip)
    optimizer.step()                                   # torch.optim.Adam(params, lr=...)

``ClipAdam`` is that pair as ONE C-ABI call (two launches: sum of squares, clip + Adam update; no host sync).
Same arithmetic as torch's (clip_grad.py / adam.py foreach path: amsgrad=False, weight_decay=0, maximize=False),
same in-place scaling of ``.grad``, same state names (``exp_avg``, ``exp_avg_sq``, ``step``).  There is no CPU
fallback: parameters must live on a CUDA device.

Scope: groups without duplicated parameters and with at most 64 tensors -- the actor and value groups (8 tensors
each).  The reference's world-model group lists the transition parameters twice, which makes torch's
own foreach update order-dependent; that group raises NotImplementedError here and stays on torch.optim.Adam.
"""
from __future__ import annotations

import ctypes as C
import math
from typing import Iterable, Optional

import torch

from . import _lib


class ClipAdam:
    def __init__(self, params: Iterable[torch.nn.Parameter], lr: float = 1e-3, betas=(0.9, 0.999), eps: float = 1e-8,
                 grad_clip: Optional[float] = None, capturable: bool = False):
        """capturable: keep the step counter on the device (like torch.optim.Adam(capturable=True)) so that step() can be
        recorded into a CUDA graph; ``step_count`` then costs a device read."""
        self.listed = list(params)                      # as given (duplicates kept: they count in the norm)
        seen, uniq = set(), []
        for p in self.listed:
            if id(p) not in seen:
                seen.add(id(p))
                uniq.append(p)
        self.params = uniq
        if not self.params:
            raise ValueError("optimizer got an empty parameter list")
        for p in self.params:
            if not p.is_cuda or p.dtype != torch.float32:
                raise RuntimeError("ClipAdam needs fp32 CUDA parameters (there is no CPU fallback)")
        self.lr, self.betas, self.eps, self.grad_clip = float(lr), (float(betas[0]), float(betas[1])), float(eps), grad_clip
        self._steps = 0
        self.state = {p: dict(exp_avg=torch.zeros_like(p), exp_avg_sq=torch.zeros_like(p)) for p in self.params}
        dev = self.params[0].device
        self._ws = torch.zeros(4, dtype=torch.float32, device=dev)
        self.grad_norm = torch.zeros((), dtype=torch.float32, device=dev)
        self.step_dev = torch.zeros(1, dtype=torch.int32, device=dev) if capturable else None

    @property
    def step_count(self) -> int:
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
        live = [p for p in self.params if p.grad is not None]
        if not live:
            return self.grad_norm.zero_()
        b1, b2 = self.betas
        step_size = bias2_sqrt = 0.0
        if self.step_dev is None:
            self._steps += 1
            step_size = self.lr / (1.0 - b1 ** self._steps)
            bias2_sqrt = math.sqrt(1.0 - b2 ** self._steps)
        counts = {}
        for p in self.listed:
            counts[id(p)] = counts.get(id(p), 0) + 1
        dup = [p for p in live if counts[id(p)] > 1]
        stream = C.c_void_p(torch.cuda.current_stream().cuda_stream)
        for lo in range(0, len(live), 64):
            chunk = live[lo:lo + 64]
            c = _lib.drm_clip_adam_call()
            c.n_tensors = len(chunk)
            c.max_norm = float(self.grad_clip) if self.grad_clip else 0.0
            c.beta1, c.beta2, c.eps = b1, b2, self.eps
            c.one_minus_beta1, c.one_minus_beta2 = 1.0 - b1, 1.0 - b2
            if self.step_dev is not None:
                c.lr, c.beta1_d, c.beta2_d, c.step_dev = self.lr, b1, b2, self.step_dev.data_ptr()
            c.step_size, c.bias2_sqrt = step_size, bias2_sqrt
            for i, p in enumerate(chunk):
                g = p.grad
                if not g.is_contiguous() or g.dtype != torch.float32:
                    raise RuntimeError("ClipAdam needs contiguous fp32 gradients")
                st = self.state[p]
                c.param[i], c.grad[i] = p.data_ptr(), g.data_ptr()
                c.exp_avg[i], c.exp_avg_sq[i] = st["exp_avg"].data_ptr(), st["exp_avg_sq"].data_ptr()
                c.numel[i] = p.numel()
            c.norm_out = self.grad_norm.data_ptr()
            c.workspace, c.workspace_bytes = self._ws.data_ptr(), self._ws.numel() * 4
            if len(live) > 64 or dup:
                raise NotImplementedError("more than 64 tensors or duplicated parameters in one group: split the group "
                                          "(the actor and value groups of Dreamer have 8 tensors each)")
            _lib.check(lib.drm_clip_adam(C.byref(c), stream), "drm_clip_adam")
        return self.grad_norm

    def state_dict(self):
        return dict(step=self.step_count, lr=self.lr, betas=self.betas, eps=self.eps, grad_clip=self.grad_clip,
                    exp_avg=[self.state[p]["exp_avg"] for p in self.params],
                    exp_avg_sq=[self.state[p]["exp_avg_sq"] for p in self.params])

    def load_state_dict(self, sd):
        self.step_count = int(sd["step"])
        for p, m, v in zip(self.params, sd["exp_avg"], sd["exp_avg_sq"]):
            self.state[p]["exp_avg"].copy_(m)
            self.state[p]["exp_avg_sq"].copy_(v)
