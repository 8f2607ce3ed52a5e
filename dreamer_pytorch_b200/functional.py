"""torch.autograd bridge over the C ABI (``include/dreamer_b200.h``).

Each Function unpacks tensors into raw device pointers, calls one ``drm_*`` entry point on the
current CUDA stream and wraps the outputs; ``backward`` calls the matching ``*_bwd`` entry point
and honours ``ctx.needs_input_grad`` -- that is how the reference's ``FreezeParameters``
(dreamer/utils/module.py:17-38, which flips ``requires_grad`` *before* the forward call) is
respected.  No PyTorch math happens here: if the CUDA library is missing the call raises.
"""
from __future__ import annotations

import ctypes as C
import math
from typing import List, Optional, Sequence

import torch
from torch.autograd.function import once_differentiable

from . import _lib, noise
from .precision import get_precision
from ._lib import MAXL

RSSM_KEYS = _lib._RSSM_W  # order of the RSSM weight tensors in every Function signature
RAW_INIT_STD = math.log(math.exp(5.0) - 1.0)


def _stream():
    _lib.ensure_init()
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


def _buf(nbytes: int, device) -> torch.Tensor:
    return torch.empty((max(int(nbytes), 16) + 3) // 4, dtype=torch.float32, device=device)


def _f32c(t: Optional[torch.Tensor]) -> Optional[torch.Tensor]:
    if t is None:
        return None
    if not t.is_cuda:
        raise RuntimeError("dreamer_pytorch_b200 ops need CUDA tensors (there is no CPU fallback)")
    if t.dtype != torch.float32:
        t = t.float()
    return t.contiguous()


_bucket_provider = None  # callable(tensors) -> flat fp32 buffer of sum(rup4(numel)) floats, or None


class bucket_provider:
    """Context manager: gradient buckets handed out by flat_like come from `fn` (parallel.PeerBucket.provide) instead of
    torch.empty -- the backward kernels then write weight gradients straight into peer-mapped memory."""

    def __init__(self, fn):
        self.fn = fn

    def __enter__(self):
        global _bucket_provider
        self.old, _bucket_provider = _bucket_provider, self.fn

    def __exit__(self, *exc):
        global _bucket_provider
        _bucket_provider = self.old


def flat_like(tensors: Sequence[torch.Tensor]) -> List[torch.Tensor]:
    """Fresh fp32 tensors shaped like `tensors`, carved as 16-byte aligned views out of ONE flat buffer (``out[0]._base``):
    the gradient bucket of an optimizer group.  The <= 3 padding floats between pieces are never read."""
    sizes = [(w.numel() + 3) // 4 * 4 for w in tensors]
    flat = _bucket_provider(tensors) if _bucket_provider is not None else None  # e.g. a slice of peer-mapped memory
    if flat is None:
        flat = torch.empty(sum(sizes), dtype=torch.float32, device=tensors[0].device)
    out, o = [], 0
    for w, n in zip(tensors, sizes):
        out.append(flat[o:o + w.numel()].view(w.shape))
        o += n
    return out


def _prec(mode):
    return _lib.BF16 if (mode or get_precision()) == "bf16" else _lib.FP32


def _count(n=1):
    _lib.launch_count += n


# ------------------------------------------------------------------------------------------------
# struct builders
# ------------------------------------------------------------------------------------------------
def _fill_mlp(m: _lib.drm_mlp, weights: Sequence[torch.Tensor], in_dim: int):
    n_linear = len(weights) // 2
    if n_linear < 1 or n_linear > MAXL:
        raise RuntimeError(f"MLP with {n_linear} Linear layers is not supported (1..{MAXL})")
    m.n_linear = n_linear
    m.in_dim = in_dim
    m.hidden = weights[0].shape[0] if n_linear > 1 else 0
    m.out_dim = weights[2 * (n_linear - 1)].shape[0]
    for l in range(n_linear):
        w, b = weights[2 * l], weights[2 * l + 1]
        exp_in = in_dim if l == 0 else m.hidden
        exp_out = m.out_dim if l == n_linear - 1 else m.hidden
        if tuple(w.shape) != (exp_out, exp_in) or tuple(b.shape) != (exp_out,):
            raise RuntimeError(f"MLP layer {l}: unexpected shapes {tuple(w.shape)} / {tuple(b.shape)}")
        m.W[l] = _lib.ptr(w)
        m.b[l] = _lib.ptr(b)


def _fill_rssm(s: _lib.drm_rssm, dims, rssm: Sequence[Optional[torch.Tensor]]):
    s.A, s.S, s.D, s.Hd, s.E = dims
    for k, t in zip(RSSM_KEYS, rssm):
        setattr(s, k, _lib.ptr(t))


class RolloutCfg:
    """Static description of a rollout (not a tensor)."""

    def __init__(self, mode, steps, n, dims, act_dist="tanh_normal", n_actor=0, mean_scale=5.0,
                 raw_init_std=RAW_INIT_STD, act_min_std=1e-4, temperature=0.1, min_std=0.1):
        self.mode, self.steps, self.n, self.dims = mode, int(steps), int(n), tuple(int(d) for d in dims)
        self.act_dist, self.n_actor = act_dist, n_actor
        self.mean_scale, self.raw_init_std, self.act_min_std = mean_scale, raw_init_std, act_min_std
        self.temperature, self.min_std = temperature, min_std
        self.precision = None  # resolved (and frozen) by the forward call


def _build_rollout(cfg: RolloutCfg, t: dict, rssm, actor, training: bool) -> _lib.drm_rollout:
    r = _lib.drm_rollout()
    r.mode, r.steps, r.n = cfg.mode, cfg.steps, cfg.n
    r.act_dist = _lib.DIST[cfg.act_dist]
    r.precision = _prec(cfg.precision)
    r.training = 1 if training else 0
    r.act_mean_scale, r.act_raw_init_std, r.act_min_std = cfg.mean_scale, cfg.raw_init_std, cfg.act_min_std
    r.act_temperature, r.min_std = cfg.temperature, cfg.min_std
    _fill_rssm(r.rssm, cfg.dims, rssm)
    if cfg.mode == _lib.MODE_POLICY:
        _fill_mlp(r.actor, actor, cfg.dims[1] + cfg.dims[2])
    for k in ("start_stoch", "start_deter", "action_in", "obs_embed", "eps_prior", "eps_post", "noise_act",
              "prior_mean", "prior_std", "prior_stoch", "deter", "post_mean", "post_std", "post_stoch", "action_out"):
        x = t.get(k)
        if noise.is_pair(x):
            # {seed, offset} instead of a tensor: the fp32 kernels generate this stream themselves (drm_rollout.noise_pair)
            if r.noise_pair not in (None, 0) and r.noise_pair != x.data_ptr():
                raise RuntimeError("one rollout takes one noise pair: pass the same pair for every generated stream")
            r.noise_pair = x.data_ptr()
            continue
        setattr(r, k, _lib.ptr(x))
    return r


def _noise_arg(x):
    return x if noise.is_pair(x) else _f32c(x)


def _resolve_noise(cfg: RolloutCfg, t: dict):
    """Pair tensors stay pairs where the fp32 kernels will run; the bf16 tensor-core rollout (tanh-normal policy rollouts
    under precision("bf16")) takes tensors, so its pairs are materialised here -- with the values the kernels would have
    generated -- in one launch.  The dict keeps the tensors: the backward pass sees the same ones."""
    names = [k for k in ("eps_prior", "eps_post", "noise_act") if noise.is_pair(t.get(k))]
    if not names:
        return
    pairs = {t[k].data_ptr() for k in names}
    if len(pairs) > 1:
        raise RuntimeError("one rollout takes one noise pair: pass the same pair for every generated stream")
    if not (cfg.precision == "bf16" and cfg.mode == _lib.MODE_POLICY and cfg.act_dist == "tanh_normal"):
        return
    A, S = cfg.dims[0], cfg.dims[1]
    width = dict(eps_prior=S, eps_post=S, noise_act=A)
    outs = noise.fill(t[names[0]].device, [(k, (cfg.steps, cfg.n, width[k]), "normal") for k in names], pair=t[names[0]])
    for k, x in zip(names, outs):
        t[k] = x


def _check_shape(name, t, shape):
    if t is None or tuple(t.shape) != tuple(shape):
        raise RuntimeError(f"{name}: expected shape {tuple(shape)}, got {None if t is None else tuple(t.shape)}")


IO_KEYS = ("start_stoch", "start_deter", "action_in", "obs_embed", "eps_prior", "eps_post", "noise_act",
           "prior_mean", "prior_std", "prior_stoch", "deter", "post_mean", "post_std", "post_stoch", "action_out")
OUT_KEYS = {
    _lib.MODE_POLICY: ("prior_mean", "prior_std", "prior_stoch", "deter", "action_out"),
    _lib.MODE_TRANSITION: ("prior_mean", "prior_std", "prior_stoch", "deter"),
    _lib.MODE_REPRESENTATION: ("prior_mean", "prior_std", "prior_stoch", "deter", "post_mean", "post_std",
                               "post_stoch"),
}


def rollout_forward_raw(cfg: RolloutCfg, t: dict, rssm, actor, training: bool, step_flags=None, tiles_out=None):
    """Validates shapes, allocates the outputs into ``t`` and runs drm_rollout_fwd.
    Returns the save buffer (None unless training)."""
    lib = _lib.load()
    if cfg.precision is None:
        cfg.precision = get_precision()
    A, S, D, Hd, E = cfg.dims
    T, n = cfg.steps, cfg.n
    dev = t["start_stoch"].device
    _check_shape("start_stoch", t["start_stoch"], (n, S))
    _check_shape("start_deter", t["start_deter"], (n, D))
    _resolve_noise(cfg, t)
    if not noise.is_pair(t["eps_prior"]):
        _check_shape("eps_prior", t["eps_prior"], (T, n, S))
    if cfg.mode == _lib.MODE_POLICY:
        if not noise.is_pair(t["noise_act"]):
            _check_shape("noise_act", t["noise_act"], (T, n, A))
    else:
        _check_shape("action", t["action_in"], (T, n, A))
    if cfg.mode == _lib.MODE_REPRESENTATION:
        _check_shape("obs_embed", t["obs_embed"], (T, n, E))
        if not noise.is_pair(t["eps_post"]):
            _check_shape("eps_post", t["eps_post"], (T, n, S))
    new = lambda w: torch.empty(T, n, w, dtype=torch.float32, device=dev)
    t.update(prior_mean=new(S), prior_std=new(S), prior_stoch=new(S), deter=new(D))
    if cfg.mode == _lib.MODE_REPRESENTATION:
        t.update(post_mean=new(S), post_std=new(S), post_stoch=new(S))
    if cfg.mode == _lib.MODE_POLICY:
        t.update(action_out=new(A))
    r = _build_rollout(cfg, t, rssm, actor, training)
    ws = _buf(lib.drm_rollout_workspace_bytes(C.byref(r)), dev)
    save = _buf(lib.drm_rollout_save_bytes(C.byref(r)), dev) if training else None
    r.workspace, r.workspace_bytes = ws.data_ptr(), ws.numel() * 4
    if save is not None:
        r.save, r.save_bytes = save.data_ptr(), save.numel() * 4
    if step_flags is not None:
        # per-step completion counters for a consumer running beside the rollout (drm_rollout.step_flags)
        r.step_flags = step_flags.data_ptr()
    if tiles_out is not None:  # SMs the rollout kernels of this call occupy (0: fp32 kernels, no gating)
        tiles_out.append(int(lib.drm_rollout_tiles(C.byref(r))))
    _lib.check(lib.drm_rollout_fwd(C.byref(r), _stream()), "drm_rollout_fwd")
    _count()
    return save


def rollout_backward_raw(cfg: RolloutCfg, t: dict, rssm, actor, save, gouts: dict, want_inputs=(False,) * 4,
                         want_rssm=False, want_actor=False, phase: int = 0, plan: Optional[dict] = None, g_wait=None,
                         wgrad_stream=None):
    """Runs drm_rollout_bwd.  gouts: name ('g_prior_mean', ...) -> tensor or None.
    want_inputs: (start_stoch, start_deter, action_in, obs_embed).
    Returns (input grads [4], rssm grads list or None, actor grads list or None).

    phase / plan: a PHASE_PREPARE call (weight packing + zero-filled sinks, may be issued early on another stream)
    stores the call description, workspace and sinks in the dict `plan`; the later PHASE_RUN call with the same dict
    reuses them (include/dreamer_b200.h, DRM_PHASE_*)."""
    lib = _lib.load()
    dev = t["start_stoch"].device
    if save is None:
        raise RuntimeError("rollout was run without grad; cannot backpropagate")
    if plan is not None and "r" in plan:
        r, g, ws, din, grssm, gactor = (plan[k] for k in ("r", "g", "ws", "din", "grssm", "gactor"))
    else:
        r = _build_rollout(cfg, t, rssm, actor, True)
        ws = _buf(lib.drm_rollout_workspace_bytes(C.byref(r)), dev)
        r.workspace, r.workspace_bytes = ws.data_ptr(), ws.numel() * 4
        r.save, r.save_bytes = save.data_ptr(), save.numel() * 4
        g = _lib.drm_rollout_grad()
        policy = cfg.mode == _lib.MODE_POLICY
        din: List[Optional[torch.Tensor]] = [None] * 4
        if not policy:
            if want_inputs[0]:
                din[0] = torch.empty_like(t["start_stoch"]); g.d_start_stoch = din[0].data_ptr()
            if want_inputs[1]:
                din[1] = torch.empty_like(t["start_deter"]); g.d_start_deter = din[1].data_ptr()
            if want_inputs[2]:
                din[2] = torch.empty_like(t["action_in"]); g.d_action_in = din[2].data_ptr()
            if want_inputs[3] and cfg.mode == _lib.MODE_REPRESENTATION:
                din[3] = torch.empty_like(t["obs_embed"]); g.d_obs_embed = din[3].data_ptr()
        n_rssm = 14 if cfg.mode == _lib.MODE_REPRESENTATION else 10
        g.want_rssm_grad = 1 if want_rssm else 0
        grssm = None
        if want_rssm:
            grssm = [torch.empty_like(rssm[i]) for i in range(n_rssm)] + [None] * (14 - n_rssm)
            for i in range(n_rssm):
                setattr(g.rssm, _lib._RSSM_G[i], grssm[i].data_ptr())
        want_actor = policy and want_actor
        g.want_actor_grad = 1 if want_actor else 0
        gactor = None
        if want_actor:
            gactor = flat_like(actor)  # views of ONE flat bucket (gactor[0]._base): a single collective averages them
            for l in range(len(actor) // 2):
                g.actor.dW[l], g.actor.db[l] = gactor[2 * l].data_ptr(), gactor[2 * l + 1].data_ptr()
        if plan is not None:
            plan.update(r=r, g=g, ws=ws, din=din, grssm=grssm, gactor=gactor)
    keep = []
    for k, x in gouts.items():
        x = _f32c(x)
        keep.append(x)
        setattr(g, k, _lib.ptr(x))
    g.phase = phase
    if g_wait is not None:  # (int32 counters [steps], target): upstream gradients produced beside this kernel
        g.g_wait_flags, g.g_wait_target = g_wait[0].data_ptr(), int(g_wait[1])
    if wgrad_stream is not None:  # actor weight-gradient GEMMs step by step behind the rollout (drm_rollout_grad.wgrad_stream)
        g.wgrad_stream = wgrad_stream.cuda_stream
    _lib.check(lib.drm_rollout_bwd(C.byref(r), C.byref(g), _stream()), "drm_rollout_bwd")
    _count()
    return din, grssm, gactor


class RolloutFn(torch.autograd.Function):
    """One persistent-kernel rollout.

    apply(cfg, start_stoch, start_deter, action_in, obs_embed, eps_prior, eps_post, noise_act,
          *rssm(14), *actor(2*n_linear)) ->
        POLICY:          (mean, std, stoch, deter, action)
        TRANSITION:      (mean, std, stoch, deter)
        REPRESENTATION:  (prior_mean, prior_std, prior_stoch, deter, post_mean, post_std, post_stoch)
    """

    @staticmethod
    def forward(ctx, cfg: RolloutCfg, start_stoch, start_deter, action_in, obs_embed, eps_prior, eps_post, noise_act,
                *weights):
        rssm = [_f32c(w) for w in weights[:14]]
        actor = [_f32c(w) for w in weights[14:]]
        t = dict(start_stoch=_f32c(start_stoch), start_deter=_f32c(start_deter), action_in=_f32c(action_in),
                 obs_embed=_f32c(obs_embed), eps_prior=_noise_arg(eps_prior), eps_post=_noise_arg(eps_post),
                 noise_act=_noise_arg(noise_act))
        training = any(ctx.needs_input_grad)
        n_rssm = 14 if cfg.mode == _lib.MODE_REPRESENTATION else 10
        if get_precision() == "bf16" and any(ctx.needs_input_grad[8 + i] for i in range(n_rssm)):
            # the tensor-core rollout produces actor gradients only (the hot path freezes the world model,
            # dreamer_algo.py:293); transition / posterior weight gradients come from the fp32 kernels
            cfg.precision = "fp32"
        save = rollout_forward_raw(cfg, t, rssm, actor, training)
        ctx.cfg = cfg
        ctx.n_w = len(weights)
        ctx.save_for_backward(*[t.get(k) for k in IO_KEYS], save, *rssm, *actor)
        return tuple(t[k] for k in OUT_KEYS[cfg.mode])

    @staticmethod
    @once_differentiable
    def backward(ctx, *gouts):
        cfg: RolloutCfg = ctx.cfg
        saved = ctx.saved_tensors
        nk = len(IO_KEYS)
        t = dict(zip(IO_KEYS, saved[:nk]))
        save = saved[nk]
        rssm = list(saved[nk + 1: nk + 15])
        actor = list(saved[nk + 15:])
        need = ctx.needs_input_grad  # index 0 is cfg
        names = ["g_" + k if k != "action_out" else "g_action_out" for k in OUT_KEYS[cfg.mode]]
        n_rssm = 14 if cfg.mode == _lib.MODE_REPRESENTATION else 10
        want_rssm = any(need[8 + i] for i in range(n_rssm))
        want_actor = any(need[8 + 14 + i] for i in range(len(actor)))
        din, grssm, gactor = rollout_backward_raw(cfg, t, rssm, actor, save, dict(zip(names, gouts)),
                                                  (need[1], need[2], need[3], need[4]), want_rssm, want_actor)
        res: List[Optional[torch.Tensor]] = [None] + din + [None, None, None]
        res += grssm if grssm is not None else [None] * 14
        res += gactor if gactor is not None else [None] * len(actor)
        for i in range(len(res)):
            if not need[i]:
                res[i] = None
        return tuple(res)


# ------------------------------------------------------------------------------------------------
# batched MLP
# ------------------------------------------------------------------------------------------------
def _build_mlp_call(x0, x1, weights, M, training, precision=None) -> _lib.drm_mlp_call:
    c = _lib.drm_mlp_call()
    k0 = x0.shape[1]
    k1 = x1.shape[1] if x1 is not None else 0
    _fill_mlp(c.mlp, weights, k0 + k1)
    c.M, c.k0, c.k1 = M, k0, k1
    c.precision = _prec(precision)
    c.training = 1 if training else 0
    c.x0, c.ld0 = _lib.ptr(x0), k0
    c.x1, c.ld1 = (_lib.ptr(x1), k1) if x1 is not None else (None, 0)
    return c


def mlp_forward_raw(x0, x1, weights, training, precision=None, wait=None, max_ctas=0):
    """Runs drm_mlp_fwd; returns (y [M,out], save tensor or None).
    wait = (flags int32 [steps], target, rows_per_step) gates the input rows on a producer kernel running beside this
    one (drm_mlp_call.wait_flags); max_ctas bounds the grid so that producer and consumer stay resident together."""
    lib = _lib.load()
    M = x0.shape[0]
    c = _build_mlp_call(x0, x1, weights, M, training, precision)
    dev = x0.device
    y = torch.empty(M, c.mlp.out_dim, dtype=torch.float32, device=dev)
    c.y = y.data_ptr()
    ws = _buf(lib.drm_mlp_workspace_bytes(C.byref(c)), dev)
    c.workspace, c.workspace_bytes = ws.data_ptr(), ws.numel() * 4
    save = None
    if training:
        save = _buf(lib.drm_mlp_save_bytes(C.byref(c)), dev)
        c.save, c.save_bytes = save.data_ptr(), save.numel() * 4
    if wait is not None:
        flags, target, rows = wait
        c.wait_flags, c.wait_target, c.wait_rows, c.wait_steps = flags.data_ptr(), int(target), int(rows), int(flags.numel())
    c.max_ctas = int(max_ctas)
    _lib.check(lib.drm_mlp_fwd(C.byref(c), _stream()), "drm_mlp_fwd")
    _count()
    return y, save


def mlp_backward_raw(x0, x1, weights, save, dy, dy_scale=None, dx0=None, dx1=None, accumulate_dx=False,
                     want_wgrad=False, precision="fp32", phase: int = 0, plan: Optional[dict] = None, gate: Optional[dict] = None):
    """Runs drm_mlp_bwd; returns the list of weight grads [dW0, db0, dW1, ...] (or None).
    phase / plan: as in rollout_backward_raw (dy / dx may be None in the PHASE_PREPARE call).
    gate: hand-over between kernels on different streams (drm_mlp_bwd_call): dict(max_ctas=, reverse=, tile_done=,
    tile_wait=, step_done=(counters, rows_per_step, steps))."""
    lib = _lib.load()
    M = x0.shape[0]
    if plan is not None and "b" in plan:
        b, ws, grads = plan["b"], plan["ws"], plan["grads"]
    else:
        b = _lib.drm_mlp_bwd_call()
        c = _build_mlp_call(x0, x1, weights, M, True, precision)
        ws = _buf(lib.drm_mlp_workspace_bytes(C.byref(c)), x0.device)
        c.workspace, c.workspace_bytes = ws.data_ptr(), ws.numel() * 4
        if save is not None:
            c.save, c.save_bytes = save.data_ptr(), save.numel() * 4
        b.fwd = c
        b.want_wgrad = 1 if want_wgrad else 0
        grads = None
        if want_wgrad:
            # views of one flat buffer (16-byte aligned pieces): a caller can rescale / all-reduce all of them at once
            grads = flat_like(weights)
            for l in range(len(weights) // 2):
                b.grad.dW[l], b.grad.db[l] = grads[2 * l].data_ptr(), grads[2 * l + 1].data_ptr()
        if plan is not None:
            plan.update(b=b, ws=ws, grads=grads, keep=(x0, x1, weights, save))
    b.dy = _lib.ptr(dy)
    b.dy_scale = _lib.ptr(dy_scale)
    b.dx0 = _lib.ptr(dx0)
    b.dx1 = _lib.ptr(dx1)
    b.accumulate_dx = 1 if accumulate_dx else 0
    b.phase = phase
    if gate:
        b.fwd.max_ctas = int(gate.get("max_ctas", 0))
        b.reverse_tiles = 1 if gate.get("reverse") else 0
        b.tile_done = gate["tile_done"].data_ptr() if gate.get("tile_done") is not None else None
        b.tile_wait = gate["tile_wait"].data_ptr() if gate.get("tile_wait") is not None else None
        if gate.get("step_done") is not None:
            cnt, rows, steps = gate["step_done"]
            b.step_rows_done, b.step_rows, b.step_count = cnt.data_ptr(), int(rows), int(steps)
    _lib.check(lib.drm_mlp_bwd(C.byref(b), _stream()), "drm_mlp_bwd")
    _count()
    return grads


class MlpFn(torch.autograd.Function):
    """y = MLP([x0 | x1]) with ELU between the Linear layers.  apply(x0, x1_or_None, *weights)."""

    @staticmethod
    def forward(ctx, x0, x1, *weights):
        x0 = _f32c(x0)
        x1 = _f32c(x1)
        weights = [_f32c(w) for w in weights]
        training = any(ctx.needs_input_grad)
        ctx.precision = get_precision()
        y, save = mlp_forward_raw(x0, x1, weights, training, ctx.precision)
        ctx.save_for_backward(x0, x1, save, *weights)
        return y

    @staticmethod
    @once_differentiable
    def backward(ctx, dy):
        x0, x1, save, *weights = ctx.saved_tensors
        need = ctx.needs_input_grad
        dy = _f32c(dy)
        dx0 = torch.empty_like(x0) if need[0] else None
        dx1 = torch.empty_like(x1) if (x1 is not None and need[1]) else None
        want_w = any(need[2:])
        grads = mlp_backward_raw(x0, x1, weights, save, dy, None, dx0, dx1, False, want_w, ctx.precision)
        res = [dx0, dx1] + (grads if grads is not None else [None] * len(weights))
        for i in range(len(res)):
            if not need[i]:
                res[i] = None
        return tuple(res)


# ------------------------------------------------------------------------------------------------
# KL / entropies
# ------------------------------------------------------------------------------------------------
class KLFn(torch.autograd.Function):
    """(div_clamped, div_raw, prior_entropy, post_entropy) = KLFn.apply(post_mean, post_std, prior_mean,
    prior_std, free_nats); only div_clamped is differentiable."""

    @staticmethod
    def forward(ctx, post_mean, post_std, prior_mean, prior_std, free_nats: float):
        lib = _lib.load()
        ts = [_f32c(x) for x in (post_mean, post_std, prior_mean, prior_std)]
        S = ts[0].shape[-1]
        M = ts[0].numel() // S
        c = _lib.drm_kl_call()
        c.M, c.S, c.free_nats = M, S, float(free_nats)
        c.post_mean, c.post_std, c.prior_mean, c.prior_std = [_lib.ptr(x) for x in ts]
        out = torch.zeros(4, dtype=torch.float32, device=ts[0].device)
        c.out = out.data_ptr()
        _lib.check(lib.drm_kl_fwd(C.byref(c), _stream()), "drm_kl_fwd")
        _count()
        ctx.save_for_backward(out, *ts)
        ctx.free_nats = float(free_nats)
        res = [out[i].clone() for i in range(4)]
        ctx.mark_non_differentiable(res[1], res[2], res[3])
        return tuple(res)

    @staticmethod
    @once_differentiable
    def backward(ctx, g_div, *_unused):
        lib = _lib.load()
        out, *ts = ctx.saved_tensors
        S = ts[0].shape[-1]
        c = _lib.drm_kl_call()
        c.M, c.S, c.free_nats = ts[0].numel() // S, S, ctx.free_nats
        c.post_mean, c.post_std, c.prior_mean, c.prior_std = [_lib.ptr(x) for x in ts]
        c.out = out.data_ptr()
        need = ctx.needs_input_grad
        grads = [torch.empty_like(x) if need[i] else None for i, x in enumerate(ts)]
        c.d_post_mean, c.d_post_std, c.d_prior_mean, c.d_prior_std = [_lib.ptr(x) for x in grads]
        _lib.check(lib.drm_kl_bwd(C.byref(c), _stream()), "drm_kl_bwd")
        _count()
        grads = [None if x is None else x * g_div for x in grads]
        return (*grads, None)


# ------------------------------------------------------------------------------------------------
# lambda-returns + actor / value losses (forward and analytic gradients in one launch)
# ------------------------------------------------------------------------------------------------
def returns_loss_raw(reward, value, pcont, discount, lambda_):
    """reward/value/pcont: [H,N] contiguous.  Returns dict(losses[2], returns, disc, d_reward, d_value,
    d_pcont, d_value_pred)."""
    lib = _lib.load()
    H, N = reward.shape
    dev = reward.device
    c = _lib.drm_returns_loss_call()
    c.H, c.N, c.discount, c.lambda_ = H, N, float(discount), float(lambda_)
    out = dict(
        losses=torch.zeros(2, dtype=torch.float32, device=dev),
        returns=torch.empty(H - 1, N, dtype=torch.float32, device=dev),
        disc=torch.empty(H - 1, N, dtype=torch.float32, device=dev),
        d_reward=torch.empty(H, N, dtype=torch.float32, device=dev),
        d_value=torch.empty(H, N, dtype=torch.float32, device=dev),
        d_value_pred=torch.empty(H, N, dtype=torch.float32, device=dev),
        d_pcont=torch.empty(H, N, dtype=torch.float32, device=dev) if pcont is not None else None,
    )
    c.reward, c.value, c.pcont = _lib.ptr(reward), _lib.ptr(value), _lib.ptr(pcont)
    for k in ("returns", "disc", "losses", "d_reward", "d_value", "d_pcont", "d_value_pred"):
        setattr(c, k, _lib.ptr(out[k]))
    _lib.check(lib.drm_returns_loss(C.byref(c), _stream()), "drm_returns_loss")
    _count()
    return out


# ------------------------------------------------------------------------------------------------
# image log-likelihood of the observation decoder
# ------------------------------------------------------------------------------------------------
class ImageLogProbFn(torch.autograd.Function):
    """logp [rows] = Independent(Normal(mean, 1), k).log_prob(obs) for mean / obs of shape [rows, D]: one fused
    reduction launch (drm_image_logprob_fwd) instead of a chain of pointwise kernels; mean may be fp32 or bf16."""

    @staticmethod
    def forward(ctx, mean, obs):
        lib = _lib.load()
        if not (mean.is_cuda and obs.is_cuda):
            raise RuntimeError("dreamer_pytorch_b200 ops need CUDA tensors (there is no CPU fallback)")
        if mean.dtype not in (torch.float32, torch.bfloat16):
            mean = mean.float()
        mean = mean.contiguous()
        obs = obs.float().contiguous()
        rows, D = mean.shape
        c = _lib.drm_image_logprob_call()
        c.rows, c.D, c.mean_bf16 = rows, D, 1 if mean.dtype == torch.bfloat16 else 0
        logp = torch.empty(rows, dtype=torch.float32, device=mean.device)
        c.mean, c.obs, c.logp = mean.data_ptr(), obs.data_ptr(), logp.data_ptr()
        _lib.check(lib.drm_image_logprob_fwd(C.byref(c), _stream()), "drm_image_logprob_fwd")
        _count()
        ctx.save_for_backward(mean, obs)
        return logp

    @staticmethod
    @once_differentiable
    def backward(ctx, g):
        lib = _lib.load()
        mean, obs = ctx.saved_tensors
        if ctx.needs_input_grad[1]:
            raise RuntimeError("the fused image log-prob is differentiable w.r.t. the decoder mean only")
        g = _f32c(g)
        c = _lib.drm_image_logprob_call()
        c.rows, c.D, c.mean_bf16 = mean.shape[0], mean.shape[1], 1 if mean.dtype == torch.bfloat16 else 0
        dmean = torch.empty_like(mean)
        c.mean, c.obs, c.g, c.dmean = mean.data_ptr(), obs.data_ptr(), g.data_ptr(), dmean.data_ptr()
        _lib.check(lib.drm_image_logprob_bwd(C.byref(c), _stream()), "drm_image_logprob_bwd")
        _count()
        return dmean, None


# ------------------------------------------------------------------------------------------------
# environment-time policy head
# ------------------------------------------------------------------------------------------------
class BiasActFn(torch.autograd.Function):
    """``act(x + bias[c])`` IN PLACE on a channels-last fp32 conv output ``x`` [N,C,H,W] that was computed without bias
    (observation.py:18-25,83-93).  One pass instead of torch's broadcast add + ReLU; the backward is one
    pass too: ``g * (y > 0)`` with the bias gradient reduced on the way (drm_bias_act_fwd / _bwd)."""

    @staticmethod
    def eligible(x: torch.Tensor, bias) -> bool:
        if bias is None or not x.is_cuda or x.dtype != torch.float32 or x.dim() != 4 or bias.dtype != torch.float32:
            return False
        C = x.shape[1]
        return (C % 4 == 0 and C <= 1024 and 256 % (C // 4) == 0 and x.numel() > 0
                and x.is_contiguous(memory_format=torch.channels_last))

    @staticmethod
    def forward(ctx, x, bias, relu: bool):
        lib = _lib.load()
        c = _lib.drm_bias_act_call()
        c.rows, c.C, c.relu = x.numel() // x.shape[1], x.shape[1], 1 if relu else 0
        c.x, c.bias = x.data_ptr(), _f32c(bias).data_ptr()
        _lib.check(lib.drm_bias_act_fwd(C.byref(c), _stream()), "drm_bias_act_fwd")
        _count()
        ctx.mark_dirty(x)
        ctx.relu = bool(relu)
        if relu:
            ctx.save_for_backward(x)
        return x

    @staticmethod
    @once_differentiable
    def backward(ctx, g):
        lib = _lib.load()
        g = g.contiguous(memory_format=torch.channels_last)
        c = _lib.drm_bias_act_call()
        c.rows, c.C, c.relu = g.numel() // g.shape[1], g.shape[1], 1 if ctx.relu else 0
        c.g = g.data_ptr()
        out = g
        if ctx.relu:
            (y,) = ctx.saved_tensors
            out = torch.empty_like(g, memory_format=torch.channels_last)
            c.y, c.g_pre = y.data_ptr(), out.data_ptr()
        db = None
        if ctx.needs_input_grad[1]:
            db = torch.empty(g.shape[1], dtype=torch.float32, device=g.device)
            c.dbias = db.data_ptr()
        if ctx.relu or db is not None:
            _lib.check(lib.drm_bias_act_bwd(C.byref(c), _stream()), "drm_bias_act_bwd")
            _count()
        return (out if ctx.needs_input_grad[0] else None), db, None


def policy_mode(raw: torch.Tensor, samples: int, mean_scale: float, raw_init_std: float, min_std: float,
                return_logp: bool = False):
    """SampleDist.mode() of the tanh-Normal action distribution (distribution.py:24-35) from the actor MLP's output
    ``raw`` [..., 2A]: ``samples`` draws, their log-probabilities, arg-max, gather -- one launch (drm_policy_mode).
    No gradient (the reference only calls it under ``torch.no_grad`` at environment-step time)."""
    from . import noise

    lib = _lib.load()
    lead = raw.shape[:-1]
    A = raw.shape[-1] // 2
    flat = _f32c(raw.reshape(-1, 2 * A))
    n = flat.shape[0]
    eps = noise.normal("mode_noise", (samples, n, A), flat.device)
    action = torch.empty(n, A, dtype=torch.float32, device=flat.device)
    c = _lib.drm_policy_mode_call()
    c.n, c.A, c.samples = n, A, int(samples)
    c.mean_scale, c.raw_init_std, c.min_std = float(mean_scale), float(raw_init_std), float(min_std)
    c.raw, c.noise, c.action = flat.data_ptr(), eps.data_ptr(), action.data_ptr()
    logp = None
    if return_logp:
        logp = torch.empty(samples, n, dtype=torch.float32, device=flat.device)
        c.logp = logp.data_ptr()
    _lib.check(lib.drm_policy_mode(C.byref(c), _stream()), "drm_policy_mode")
    _count()
    action = action.reshape(*lead, A)
    return (action, logp) if return_logp else action
