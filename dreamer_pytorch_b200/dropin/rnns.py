"""Drop-in for ``dreamer/models/rnns.py`` (reference lines cited per symbol).

Same names, constructor signatures, parameter names/shapes (so ``state_dict`` round-trips with the
reference) and return types -- but every step / rollout is ONE launch of the persistent sm_100a
kernel behind ``drm_rollout_fwd`` (include/dreamer_b200.h) instead of a Python loop over ~40 ATen
kernels per step.  CUDA fp32 tensors only; anything else raises (no CPU fallback).
"""
from __future__ import annotations

import torch
import torch.distributions as td
import torch.nn as nn

from .. import _lib, noise
from ..functional import RolloutCfg, RolloutFn
from ._compat import buffer_method, namedarraytuple

RSSMState = namedarraytuple("RSSMState", ["mean", "std", "stoch", "deter"])  # rnns.py:10


def stack_states(rssm_states: list, dim):
    """rnns.py:13-19"""
    return RSSMState(*(torch.stack([getattr(s, f) for s in rssm_states], dim=dim) for f in RSSMState._fields))


def get_feat(rssm_state: RSSMState):
    """rnns.py:22-23"""
    return torch.cat((rssm_state.stoch, rssm_state.deter), dim=-1)


def get_dist(rssm_state: RSSMState):
    """rnns.py:26-27"""
    return td.independent.Independent(td.Normal(rssm_state.mean, rssm_state.std), 1)


class TransitionBase(nn.Module):
    def forward(self, prev_action, prev_state):
        raise NotImplementedError


class RepresentationBase(nn.Module):
    def forward(self, obs_embed, prev_action, prev_state):
        raise NotImplementedError


class RollOutModule(nn.Module):
    def forward(self, steps, obs_embed, prev_action, prev_state):
        raise NotImplementedError


def _only_elu_normal(activation, distribution):
    if activation is not nn.ELU or distribution is not td.Normal:
        raise NotImplementedError("the fused RSSM kernels implement activation=nn.ELU and distribution=td.Normal")


def _zeros_state(batch_size, stoch, deter, **kwargs):
    return RSSMState(torch.zeros(batch_size, stoch, **kwargs), torch.zeros(batch_size, stoch, **kwargs),
                     torch.zeros(batch_size, stoch, **kwargs), torch.zeros(batch_size, deter, **kwargs))


class RSSMTransition(TransitionBase):
    """rnns.py:56-107.  Parameters live in the same sub-modules as the reference
    (``_cell``, ``_rnn_input_model.0``, ``_stochastic_prior_model.{0,2}``); the kernels read their storage."""

    def __init__(self, action_size, stochastic_size=30, deterministic_size=200, hidden_size=200,
                 activation=nn.ELU, distribution=td.Normal):
        super().__init__()
        _only_elu_normal(activation, distribution)
        self._action_size = int(action_size)
        self._stoch_size = stochastic_size
        self._deter_size = deterministic_size
        self._hidden_size = hidden_size
        self._activation = activation
        self._cell = nn.GRUCell(hidden_size, deterministic_size)
        self._rnn_input_model = nn.Sequential(nn.Linear(self._action_size + stochastic_size, hidden_size), activation())
        self._stochastic_prior_model = nn.Sequential(
            nn.Linear(deterministic_size, hidden_size), activation(), nn.Linear(hidden_size, 2 * stochastic_size))
        self._dist = distribution

    def initial_state(self, batch_size, **kwargs):
        return _zeros_state(batch_size, self._stoch_size, self._deter_size, **kwargs)

    def weights(self):
        """The 10 transition tensors in the C-ABI order (drm_rssm)."""
        c, i, p = self._cell, self._rnn_input_model[0], self._stochastic_prior_model
        return [i.weight, i.bias, c.weight_ih, c.weight_hh, c.bias_ih, c.bias_hh, p[0].weight, p[0].bias,
                p[2].weight, p[2].bias]

    def dims(self, embed=0):
        return (self._action_size, self._stoch_size, self._deter_size, self._hidden_size, embed)

    def rollout(self, steps, action, prev_state):
        """[steps,n,A] actions -> stacked prior states; one launch (rollout_transition, rnns.py:207-222)."""
        n = prev_state.stoch.shape[0]
        cfg = RolloutCfg(_lib.MODE_TRANSITION, steps, n, self.dims())
        eps = noise.normal("eps_prior", (steps, n, self._stoch_size), prev_state.stoch.device, lazy=True)
        mean, std, stoch, deter = RolloutFn.apply(cfg, prev_state.stoch, prev_state.deter, action, None, eps, None,
                                                  None, *self.weights(), None, None, None, None)
        return RSSMState(mean, std, stoch, deter)

    def forward(self, prev_action: torch.Tensor, prev_state: RSSMState):
        """rnns.py:98-107 (a one-step rollout)."""
        lead = prev_action.shape[:-1]
        flat = RSSMState(*(f.reshape(-1, f.shape[-1]) for f in prev_state))
        out = self.rollout(1, prev_action.reshape(1, -1, prev_action.shape[-1]), flat)
        return RSSMState(*(f.reshape(*lead, f.shape[-1]) for f in out))


class RSSMRepresentation(RepresentationBase):
    """rnns.py:110-159."""

    def __init__(self, transition_model: RSSMTransition, obs_embed_size, action_size, stochastic_size=30,
                 deterministic_size=200, hidden_size=200, activation=nn.ELU, distribution=td.Normal):
        super().__init__()
        _only_elu_normal(activation, distribution)
        self._transition_model = transition_model
        self._obs_embed_size = obs_embed_size
        self._action_size = int(action_size)
        self._stoch_size = stochastic_size
        self._deter_size = deterministic_size
        self._hidden_size = hidden_size
        self._activation = activation
        self._dist = distribution
        self._stochastic_posterior_model = nn.Sequential(
            nn.Linear(deterministic_size + obs_embed_size, hidden_size), activation(),
            nn.Linear(hidden_size, 2 * stochastic_size))

    def initial_state(self, batch_size, **kwargs):
        return _zeros_state(batch_size, self._stoch_size, self._deter_size, **kwargs)

    def weights(self):
        q = self._stochastic_posterior_model
        return self._transition_model.weights() + [q[0].weight, q[0].bias, q[2].weight, q[2].bias]

    def rollout(self, steps, obs_embed, action, prev_state):
        """One launch for all ``steps`` (rollout_representation, rnns.py:179-205)."""
        n = prev_state.stoch.shape[0]
        dev = prev_state.stoch.device
        cfg = RolloutCfg(_lib.MODE_REPRESENTATION, steps, n, self._transition_model.dims(self._obs_embed_size))
        S = self._stoch_size
        # the reference draws prior eps then posterior eps at every step
        # (production: one {seed, offset} pair, the kernel generates both streams itself, noise.py)
        eps_prior = noise.normal("eps_prior", (steps, n, S), dev, lazy=True)
        eps_post = noise.normal("eps_post", (steps, n, S), dev, lazy=True,
                                pair=eps_prior if noise.is_pair(eps_prior) else None)
        pm, ps, pst, deter, qm, qs, qst = RolloutFn.apply(cfg, prev_state.stoch, prev_state.deter, action, obs_embed,
                                                          eps_prior, eps_post, None, *self.weights())
        return RSSMState(pm, ps, pst, deter), RSSMState(qm, qs, qst, deter)

    def forward(self, obs_embed: torch.Tensor, prev_action: torch.Tensor, prev_state: RSSMState):
        """rnns.py:149-159 (a one-step rollout)."""
        lead = prev_action.shape[:-1]
        flat = RSSMState(*(f.reshape(-1, f.shape[-1]) for f in prev_state))
        prior, post = self.rollout(1, obs_embed.reshape(1, -1, obs_embed.shape[-1]),
                                   prev_action.reshape(1, -1, prev_action.shape[-1]), flat)
        shape = lambda st: RSSMState(*(f.reshape(*lead, f.shape[-1]) for f in st))
        return shape(prior), shape(post)


class RSSMRollout(RollOutModule):
    """rnns.py:162-244."""

    def __init__(self, representation_model: RSSMRepresentation, transition_model: RSSMTransition):
        super().__init__()
        self.representation_model = representation_model
        self.transition_model = transition_model

    def forward(self, steps: int, obs_embed: torch.Tensor, action: torch.Tensor, prev_state: RSSMState):
        return self.rollout_representation(steps, obs_embed, action, prev_state)

    def rollout_representation(self, steps: int, obs_embed: torch.Tensor, action: torch.Tensor, prev_state: RSSMState):
        """rnns.py:179-205: prior, posterior, each [steps, batch, .]"""
        return self.representation_model.rollout(steps, obs_embed[:steps], action[:steps], prev_state)

    def rollout_transition(self, steps: int, action: torch.Tensor, prev_state: RSSMState):
        """rnns.py:207-222"""
        return self.transition_model.rollout(steps, action[:steps], prev_state)

    def rollout_policy(self, steps: int, policy, prev_state: RSSMState):
        """rnns.py:224-244.  ``policy`` is a free callable in the reference API
        (tests/test_rnns.py:83-90).  When it is the bound ``AgentModel.policy`` of a model that owns
        this transition, the whole H-step imagination (actor MLP -> action sample -> transition) is
        one persistent kernel; any other callable is driven step by step through the fused
        one-step transition."""
        state = buffer_method(prev_state, "detach")
        owner = getattr(policy, "__self__", None)
        fused = getattr(owner, "fused_policy_rollout", None)
        if fused is not None and getattr(policy, "__func__", None) is getattr(type(owner), "policy", None) \
                and getattr(owner, "transition", None) is self.transition_model and owner.can_fuse_policy():
            return fused(steps, state)
        next_states, actions = [], []
        for _ in range(steps):
            action, _ = policy(buffer_method(state, "detach"))
            state = self.transition_model(action, state)
            next_states.append(state)
            actions.append(action)
        return stack_states(next_states, dim=0), torch.stack(actions, dim=0)
