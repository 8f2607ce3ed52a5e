"""Fused loss section of ``Dreamer.loss`` (dreamer/algos/dreamer_algo.py:270-337,398-422).

``imagination_losses`` is the B200-native replacement of the actor / value block: one persistent
imagination kernel (actor MLP -> action sample -> RSSM transition, H steps), three fused head MLP
kernels over all H*N imagined states, one kernel for lambda-returns + both losses + their analytic
gradients, and hand-written backward kernels.  The math is the reference's; the freezing pattern of
``FreezeParameters`` (world model frozen for the actor loss, everything but the value model frozen
for the value loss) is built in.

``dreamer_loss`` mirrors the whole of ``Dreamer.loss`` on top of it for callers that do not go
through the unchanged ``dreamer_algo.py`` (which also works, via the drop-in modules).
"""
from __future__ import annotations

from collections import namedtuple
from typing import Optional

import torch

from . import _lib, noise
from .dropin.dense import mlp_weights
from .precision import get_precision
from .dropin.rnns import RSSMState, get_feat
from .functional import (KLFn, RolloutCfg, _f32c, mlp_backward_raw, mlp_forward_raw, returns_loss_raw,
                         rollout_backward_raw, rollout_forward_raw)

_SIDE = {}


# Scheduling switches (results do not depend on them, tests/test_graph_gpu.py).  DRM_B200_SCHEDULE=plain in the environment
# turns all of them off: every kernel in stream order; a comma-separated subset of "heads,prepare,backward" keeps those.
import os as _os

_PIPELINE_HEADS = _PREPARE_BACKWARD = _PIPELINE_BACKWARD = _FOLLOW_WGRAD = True
# SMs per head kernel below which the heads do NOT run beside the rollout (plain stream order instead).  With few SMs the
# gated head kernels cannot keep up with the rollout and finish long after it: measured with 128-row tiles, ms per step
# gated / plain at 34 SMs per head 2.35 / 2.48, at 26 SMs 3.48 / 2.67, at 18 SMs 5.43 / 2.91 (profiles/r2_experiments.md).
_MIN_FREE = int(_os.environ.get("DRM_B200_MIN_FREE", "32"))


def set_schedule(spec: str = "heads,prepare,backward") -> str:
    """Selects which kernels run beside each other; returns the previous setting.  "plain": every kernel in stream
    order (what bench.py uses to time the dominant kernel in isolation)."""
    global _PIPELINE_HEADS, _PREPARE_BACKWARD, _PIPELINE_BACKWARD, _FOLLOW_WGRAD, _sched_spec
    prev = _sched_spec
    parts = spec.split(",")
    _PIPELINE_HEADS = "heads" in parts  # heads beside the rollout when SMs are idle (see _ActorLossFn.forward)
    _PREPARE_BACKWARD = "prepare" in parts  # weight packing / zero-filled sinks of the backward pass beside the forward pass
    _PIPELINE_BACKWARD = "backward" in parts  # head data-gradient kernels beside the reverse-time rollout (_ActorLossFn.backward)
    _FOLLOW_WGRAD = "nofollow" not in parts  # actor weight gradients step by step behind the rollout backward
    _sched_spec = spec
    return prev


_sched_spec = ""
set_schedule(_os.environ.get("DRM_B200_SCHEDULE", "heads,prepare,backward"))


# CUDA loads a kernel on its first launch (lazy module loading), and that load may have to wait for the device to go idle.
# A kernel that spins on counters must therefore never be on the device while another kernel is launched for the first
# time.  This library's own kernels are all loaded up front (drm_init, _lib.ensure_init); for everything else a step
# launches (torch's fill / random / copy kernels) the first full forward + backward pass of every distinct configuration
# on a device runs in plain stream order, the overlapped schedule starts with the second.
_WARM = {}


def _warm_key(device, prec, cfg, pcont) -> tuple:
    return (device.type, device.index, prec, cfg.act_dist, cfg.dims, cfg.steps, bool(pcont), cfg.n >= 128 * _sm_count(device))


def _overlap_ok(key) -> bool:
    return _WARM.get(key, 0) >= 1


# Single tenant: kernels that run beside each other count on the SMs their producers leave idle, and the side streams are
# per device.  A pass issued on ANOTHER stream while the previous pass on this device may still be running must not
# make the same assumption (two sets of spinning consumers could keep each other's producers off the SMs until the
# bounded spin traps): it is demoted to plain stream order, which never waits on anything but its own stream.
# Passes on one stream (every training loop, every graph) are ordered by the stream and never take this branch.
_LAST_PASS = {}     # device index -> (cuda_stream handle, end-of-pass event)
demoted_passes = 0  # forward / backward halves demoted to plain order because of a pass in flight on another stream


def _exclusive(device) -> bool:
    global demoted_passes
    if torch.cuda.is_current_stream_capturing():
        return True  # a capture owns its streams; replays of one graph are ordered by the stream they are launched on
    last = _LAST_PASS.get(device.index)
    if last is None or last[0] == torch.cuda.current_stream(device).cuda_stream or last[1].query():
        return True
    demoted_passes += 1
    return False


def _mark_pass(device):
    if torch.cuda.is_current_stream_capturing():
        return
    cur = torch.cuda.current_stream(device)
    ev = torch.cuda.Event()
    ev.record(cur)
    _LAST_PASS[device.index] = (cur.cuda_stream, ev)


def _side_stream(device, k: int = 0) -> "torch.cuda.Stream":
    key = (device.type, device.index, k)
    if key not in _SIDE:
        _SIDE[key] = torch.cuda.Stream(device=device)
    return _SIDE[key]


def _sm_count(device) -> int:
    return torch.cuda.get_device_properties(device).multi_processor_count


def stoch_dev(t):
    return t["start_stoch"].device


ImagInfo = namedtuple("ImagInfo", ["imag", "actions", "reward", "value", "returns", "discount"])


class _ActorLossFn(torch.autograd.Function):
    """actor_loss = -mean(discount * lambda_returns) over an H-step imagined rollout.

    Differentiable w.r.t. the actor weights only (dreamer_algo.py:293-324: the world model, the
    value model and pcont are frozen; the start state is detached, rnns.py:236).
    Also returns, as non-differentiable by-products, everything the value loss needs.
    """

    @staticmethod
    def forward(ctx, pack, *actor_w):
        (cfg, t, rssm, reward_w, value_w, pcont_w, discount, lambda_, shared) = pack
        actor = [_f32c(w) for w in actor_w]
        training = any(ctx.needs_input_grad)
        H, N = cfg.steps, cfg.n
        S, D = cfg.dims[1], cfg.dims[2]
        prec = get_precision()
        heads = [(reward_w, training), (value_w, True)] + ([(pcont_w, training)] if pcont_w else [])
        # The rollout occupies one SM per 64 (or 128) start states.  When that leaves SMs idle, the heads run BESIDE it
        # on side streams, gated per time step by the completion counters the rollout kernel publishes: the heads of
        # step t overlap the rollout of step t+1 (drm_rollout.step_flags / drm_mlp_call.wait_flags).
        ctx.set_materialize_grads(False)  # 12 by-products: no zero tensors for their (unused) gradients
        flags = tiles = fork = None
        wkey = _warm_key(stoch_dev(t), prec, cfg, pcont_w)
        warm = _overlap_ok(wkey) and _exclusive(stoch_dev(t))
        pipe_heads, prepare = _PIPELINE_HEADS and warm, _PREPARE_BACKWARD and warm
        if prec == "bf16" and pipe_heads:
            # zero-filled on the main stream BEFORE the fork: the head kernels on the side streams poll these counters
            # and must never see the previous pass' values (the block is reused across iterations / graph replays)
            flags = torch.zeros(H, dtype=torch.int32, device=stoch_dev(t))
        if prec == "bf16" and (pipe_heads or prepare):
            fork = torch.cuda.Event()
            fork.record(torch.cuda.current_stream())  # the side streams branch off BEFORE the rollout is enqueued
        if prec == "bf16":
            tiles = []
        save = rollout_forward_raw(cfg, t, rssm, actor, training, step_flags=flags, tiles_out=tiles)
        x0 = t["prior_stoch"].view(H * N, S)
        x1 = t["deter"].view(H * N, D)
        # heads over all imagined states; hidden activations are kept: the actor path needs their
        # data-gradients, the value loss the value head's weight-gradients
        free = (_sm_count(x0.device) - tiles[0]) // len(heads) if tiles and tiles[0] > 0 else 0
        if flags is not None and free >= _MIN_FREE:
            main = torch.cuda.current_stream()
            outs = []
            for i, (w, keep) in enumerate(heads):
                side = _side_stream(x0.device, 1 + i)
                side.wait_event(fork)  # grids are capped (max_ctas): rollout and heads are resident together
                with torch.cuda.stream(side):
                    outs.append(mlp_forward_raw(x0, x1, w, keep, prec, wait=(flags, tiles[0], N), max_ctas=free))
                    for o in outs[-1]:
                        if o is not None:
                            o.record_stream(main)
            for i in range(len(heads)):
                main.wait_stream(_side_stream(x0.device, 1 + i))
        else:
            outs = [mlp_forward_raw(x0, x1, w, keep, prec) for (w, keep) in heads]
        (reward, rsave), (value, vsave) = outs[0], outs[1]
        pc, psave = outs[2] if pcont_w else (None, None)
        # The backward pass starts by re-packing the weights for the transposed products and zero-filling the gradient
        # sinks.  None of that depends on the forward results, so it is issued now, on its own stream, beside the rollout
        # (DRM_PHASE_PREPARE); backward() then only launches the compute kernels (DRM_PHASE_RUN).
        plans = None
        if training and fork is not None and prepare:
            main = torch.cuda.current_stream()
            prep = _side_stream(x0.device, 0)
            prep.wait_event(fork)
            with torch.cuda.stream(prep):  # buffers are allocated on `prep` too: it runs ahead of the main stream
                plans = dict(g_stoch=torch.empty_like(t["prior_stoch"]), g_deter=torch.empty_like(t["deter"]),
                             reward={}, value={}, pcont={}, vw={}, roll={})
                gs, gd = plans["g_stoch"].view(H * N, -1), plans["g_deter"].view(H * N, -1)
                P = _lib.PHASE_PREPARE
                mlp_backward_raw(x0, x1, reward_w, rsave, None, None, gs, gd, False, False, prec, P, plans["reward"])
                mlp_backward_raw(x0, x1, value_w, vsave, None, None, gs, gd, True, False, prec, P, plans["value"])
                if pcont_w:
                    mlp_backward_raw(x0, x1, pcont_w, psave, None, None, gs, gd, True, False, prec, P, plans["pcont"])
                mlp_backward_raw(x0, x1, value_w, vsave, None, None, None, None, False, True, prec, P, plans["vw"])
                rollout_backward_raw(cfg, t, rssm, actor, save, {}, want_actor=True, phase=P, plan=plans["roll"])
                plans["ev"] = torch.cuda.Event()
                plans["ev"].record(prep)
            held = [plans["g_stoch"], plans["g_deter"], plans["roll"]["ws"]] + plans["roll"]["gactor"] + [plans["vw"]["grads"][0]._base]
            held += [plans[k]["ws"] for k in ("reward", "value", "vw")] + ([plans["pcont"]["ws"]] if pcont_w else [])
            for x in held:
                x.record_stream(main)
        out = returns_loss_raw(reward.view(H, N), value.view(H, N), None if pc is None else pc.view(H, N), discount,
                               lambda_)
        ctx.pack = (cfg, t, rssm, actor, save, reward_w, value_w, pcont_w, rsave, vsave, psave, out, x0, x1, prec, shared,
                    plans, tiles[0] if tiles else 0, wkey)
        _mark_pass(x0.device)
        aux = (t["prior_mean"], t["prior_std"], t["prior_stoch"], t["deter"], t["action_out"], reward.view(H, N, 1),
               value.view(H, N, 1), out["returns"].view(H - 1, N, 1), out["disc"].view(H - 1, N, 1),
               out["losses"][1].clone(), out["d_value_pred"], vsave if vsave is not None else torch.empty(0))
        ctx.mark_non_differentiable(*aux)
        return (out["losses"][0].clone(), *aux)

    @staticmethod
    def backward(ctx, g_loss, *_):
        (cfg, t, rssm, actor, save, reward_w, value_w, pcont_w, rsave, vsave, psave, out, x0, x1, prec, shared,
         plans, roll_tiles, wkey) = ctx.pack
        ctx.pack = None
        if g_loss is None:
            return (None,) * (1 + len(actor))
        H, N = cfg.steps, cfg.n
        g_loss = _f32c(g_loss)
        main = torch.cuda.current_stream()
        if plans is not None:
            main.wait_event(plans["ev"])
            g_stoch, g_deter = plans["g_stoch"], plans["g_deter"]
            ph = _lib.PHASE_RUN
            pr, pv, pp, pvw, proll = (plans[k] for k in ("reward", "value", "pcont", "vw", "roll"))
        else:
            g_stoch = torch.empty_like(t["prior_stoch"])
            g_deter = torch.empty_like(t["deter"])
            ph, pr, pv, pp, pvw, proll = _lib.PHASE_ALL, None, None, None, None, None
        gs, gd = g_stoch.view(H * N, -1), g_deter.view(H * N, -1)
        heads = [(reward_w, rsave, out["d_reward"], pr), (value_w, vsave, out["d_value"], pv)]
        if pcont_w:
            heads.append((pcont_w, psave, out["d_pcont"], pp))
        # The reverse-time rollout needs the heads' data gradients of step t only when it gets to step t, and it leaves
        # most SMs idle: the head kernels run BESIDE it on side streams, last step first, each accumulating into the rows
        # the previous head has finished (tile_done / tile_wait) and the last one publishing per-step row counters the
        # rollout kernel waits on (step_rows_done / g_wait_flags).  Grids are capped so that all of them are resident.
        free = (_sm_count(g_loss.device) - roll_tiles) // len(heads) if roll_tiles > 0 else 0
        g_wait = None
        if prec == "bf16" and _PIPELINE_BACKWARD and free >= _MIN_FREE and _overlap_ok(wkey) and _exclusive(g_loss.device):
            mlp_tiles = (H * N + 127) // 128
            cnt = torch.zeros(len(heads) * mlp_tiles + H, dtype=torch.int32, device=g_loss.device)
            steps_done = cnt[len(heads) * mlp_tiles:]
            fork = torch.cuda.Event()
            fork.record(main)
            for i, (w, sv, dy, pl) in enumerate(heads):
                gate = dict(max_ctas=free, reverse=True, tile_done=cnt[i * mlp_tiles:(i + 1) * mlp_tiles],
                            tile_wait=cnt[(i - 1) * mlp_tiles:i * mlp_tiles] if i > 0 else None,
                            step_done=(steps_done, N, H) if i == len(heads) - 1 else None)
                hs = _side_stream(g_loss.device, 1 + i)
                hs.wait_event(fork)
                with torch.cuda.stream(hs):
                    mlp_backward_raw(x0, x1, w, sv, dy.view(H * N, 1), g_loss, gs, gd, i > 0, False, prec, ph, pl, gate)
            g_wait = (steps_done, N)
            joins = [_side_stream(g_loss.device, 1 + i) for i in range(len(heads))]
        else:
            for i, (w, sv, dy, pl) in enumerate(heads):
                mlp_backward_raw(x0, x1, w, sv, dy.view(H * N, 1), g_loss, gs, gd, i > 0, False, prec, ph, pl)
            joins = []
        # The reverse-time rollout kernel occupies one SM per 64 (or 128) start states; the value loss' weight-gradient
        # pass (dreamer_algo.py:331-337) depends on nothing computed here, so it runs beside it on a side stream and
        # _ValueLossFn.backward only has to scale the result (it recomputes if this backward never ran).
        side = _side_stream(g_loss.device)
        side.wait_stream(main)
        for hs in joins:  # not critical: starts once the head kernels have left, so that the rollout is never kept waiting
            side.wait_stream(hs)
        if shared is not None and not shared.get("value_done") and "value_grads" not in shared:
            with torch.cuda.stream(side):
                vg = mlp_backward_raw(x0, x1, value_w, vsave, out["d_value_pred"].view(H * N, 1), None, None, None, False,
                                      True, prec, ph, pvw)
                ev = torch.cuda.Event()
                ev.record(side)
            for g in vg:
                g.record_stream(main)
            shared["value_grads"] = (vg, ev)
        # the actor's weight-gradient GEMMs follow the rollout step by step on the side stream (behind the value pass)
        follow = side if (g_wait is not None and plans is not None and _FOLLOW_WGRAD) else None
        _, _, gactor = rollout_backward_raw(cfg, t, rssm, actor, save, dict(g_prior_stoch=g_stoch, g_deter=g_deter),
                                            want_actor=True, phase=ph, plan=proll, g_wait=g_wait, wgrad_stream=follow)
        for hs in joins:
            main.wait_stream(hs)
        # always joined: the side stream's value pass reads tensors owned by the main stream's allocator pool (x0 / x1 /
        # vsave / d_value_pred / the prepared workspace); after the join none of them can be recycled under it, whether or
        # not value_loss.backward() is ever called
        main.wait_stream(side)
        _WARM[wkey] = _WARM.get(wkey, 0) + 1
        _mark_pass(g_loss.device)
        return (None, *gactor)


class _ValueLossFn(torch.autograd.Function):
    """value_loss = -mean(discount * Normal(value_model(feat[:-1]), 1).log_prob(returns))
    (dreamer_algo.py:331-337).  The forward values were already produced by the imagination pass
    (same weights, same detached features), so only the weight-gradient pass runs here."""

    @staticmethod
    def forward(ctx, pack, *value_w):
        (x0, x1, vsave, d_value_pred, loss, prec, shared) = pack
        ctx.set_materialize_grads(False)
        ctx.pack = (x0, x1, vsave, d_value_pred, [_f32c(w) for w in value_w], prec, shared)
        return loss.clone()

    @staticmethod
    def backward(ctx, g_loss):
        x0, x1, vsave, d_value_pred, value_w, prec, shared = ctx.pack
        ctx.pack = None
        if g_loss is None:
            return (None,) * (1 + len(value_w))
        cached = shared.pop("value_grads", None) if shared is not None else None
        if shared is not None:
            shared["value_done"] = True
        if cached is not None:
            grads, ev = cached  # computed beside the rollout backward with unit upstream gradient
            torch.cuda.current_stream().wait_event(ev)
            base = grads[0]._base  # mlp_backward_raw hands out views of one flat buffer: one kernel rescales them all
            if base is not None:
                base.mul_(_f32c(g_loss))
            else:
                torch._foreach_mul_(grads, _f32c(g_loss))
            return (None, *grads)
        grads = mlp_backward_raw(x0, x1, value_w, vsave, d_value_pred.view(-1, 1), _f32c(g_loss), None, None, False,
                                 True, prec)
        return (None, *grads)


def imagination_losses(model, start_state: RSSMState, horizon: int, discount: float = 0.99,
                       discount_lambda: float = 0.95, use_pcont: bool = False):
    """Actor and value losses of ``Dreamer.loss`` from (detached) start states [N, .].

    Returns (actor_loss, value_loss, ImagInfo).  ``actor_loss.backward()`` fills the action
    decoder's gradients, ``value_loss.backward()`` the value model's -- exactly the pattern of
    dreamer_algo.py:174-180.
    """
    dec = model.action_decoder
    if not model.can_fuse_policy():
        raise RuntimeError("imagination_losses needs a policy the fused kernel covers (tanh_normal in train() mode, "
                           "one_hot or relaxed_one_hot)")
    stoch = _f32c(start_state.stoch.detach())
    deter = _f32c(start_state.deter.detach())
    N, dev = stoch.shape[0], stoch.device
    A, S = int(model.action_size), model.stochastic_size
    cfg = RolloutCfg(_lib.MODE_POLICY, horizon, N, model.transition.dims(), act_dist=model.action_dist,
                     mean_scale=float(dec.mean_scale), raw_init_std=float(dec.raw_init_std),
                     act_min_std=float(dec.min_std))
    if (model.action_dist == "tanh_normal" and get_precision() == "bf16"
            and not noise.is_injected("noise_act", "eps_prior")):
        # tensor-core rollout: both N(0,1) blocks from ONE launch of the Philox kernel (which also advances the offset)
        noise_act, eps = noise.fill(dev, [("noise_act", (horizon, N, A), "normal"), ("eps_prior", (horizon, N, S), "normal")])
    else:
        # fp32 kernels: a {seed, offset} pair; forward and backward generate the values where they would have loaded them
        draw = noise.normal if model.action_dist == "tanh_normal" else noise.gumbel
        noise_act = draw("noise_act", (horizon, N, A), dev, lazy=True)
        eps = noise.normal("eps_prior", (horizon, N, S), dev, lazy=True, pair=noise_act if noise.is_pair(noise_act) else None)
    t = dict(start_stoch=stoch, start_deter=deter, eps_prior=eps, noise_act=noise_act)
    frozen = lambda ws: [_f32c(w.detach()) for w in ws]
    rssm = frozen(model.transition.weights()) + [None] * 4
    reward_w = frozen(mlp_weights(model.reward_model.model))
    value_params = mlp_weights(model.value_model.model)
    pcont_w = frozen(mlp_weights(model.pcont.model)) if use_pcont else []
    shared = {}  # lets the two loss nodes overlap independent backward work (see _ActorLossFn.backward)
    pack = (cfg, t, rssm, reward_w, frozen(value_params), pcont_w, discount, discount_lambda, shared)
    (actor_loss, mean, std, st, dt, actions, reward, value, returns, disc, vloss, d_value_pred,
     vsave) = _ActorLossFn.apply(pack, *dec.weights())
    x0, x1 = st.view(horizon * N, -1), dt.view(horizon * N, -1)
    value_loss = _ValueLossFn.apply((x0, x1, vsave, d_value_pred, vloss, get_precision(), shared), *value_params)
    info = ImagInfo(RSSMState(mean, std, st, dt), actions, reward, value, returns, disc)
    return actor_loss, value_loss, info


LossInfo = namedtuple("LossInfo", ["model_loss", "actor_loss", "value_loss", "prior_entropy", "post_entropy", "divergence",
                                   "reward_loss", "image_loss", "pcont_loss"])


def dreamer_loss(model, observation, action, reward, done=None, horizon=15, discount=0.99, discount_lambda=0.95,
                 free_nats=3.0, kl_scale=1.0, use_pcont=False, pcont_scale=10.0, obs_embed: Optional[torch.Tensor] = None):
    """``Dreamer.loss`` (dreamer_algo.py:213-371) on the fused kernels.

    observation: float [T,B,C,H,W] already normalised (or None with ``obs_embed`` [T,B,E] given, the
    RSSM-only configuration); action [T,B,A]; reward [T,B,1]; done [T,B,1].
    Returns (model_loss, actor_loss, value_loss, LossInfo).
    """
    T, B = action.shape[:2]
    embed = model.observation_encoder(observation) if obs_embed is None else obs_embed
    prev = model.representation.initial_state(B, device=action.device, dtype=action.dtype)
    prior, post = model.rollout.rollout_representation(T, embed, action, prev)
    feat = get_feat(post)
    reward_pred = model.reward_model(feat)
    reward_loss = -torch.mean(reward_pred.log_prob(reward))
    if obs_embed is None:
        image_pred = model.observation_decoder(feat)
        image_loss = -torch.mean(image_pred.log_prob(observation))
    else:
        image_loss = reward_loss.new_zeros(())
    pcont_loss = reward_loss.new_zeros(())
    if use_pcont:
        pcont_pred = model.pcont(feat)
        pcont_loss = -torch.mean(pcont_pred.log_prob(discount * (1 - done.float())))
    div, _, prior_ent, post_ent = KLFn.apply(post.mean, post.std, prior.mean, prior.std, free_nats)
    model_loss = kl_scale * div + reward_loss + image_loss
    if use_pcont:
        model_loss = model_loss + pcont_scale * pcont_loss
    with torch.no_grad():
        src = post[:-1] if use_pcont else post
        flat = RSSMState(*(f.reshape(-1, f.shape[-1]) for f in src))
    actor_loss, value_loss, _ = imagination_losses(model, flat, horizon, discount, discount_lambda, use_pcont)
    info = LossInfo(model_loss, actor_loss, value_loss, prior_ent, post_ent, div, reward_loss, image_loss, pcont_loss)
    return model_loss, actor_loss, value_loss, info
