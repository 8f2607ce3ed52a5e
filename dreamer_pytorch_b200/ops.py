"""``torch.ops.dreamer_b200.*`` -- the torch operator layer of SURVEY.md 8(b).

The survey asks for ops registered by a C++ ``TORCH_LIBRARY(dreamer_b200, m)`` shim.  The library's boundary is a torch-free C
ABI (``include/dreamer_b200.h``) bound with ctypes, so the ops are registered from Python with ``torch.library`` instead:
same namespace, same schemas a C++ shim would declare, dispatched as ``CompositeImplicitAutograd`` kernels that run the
``torch.autograd.Function`` bridges of ``functional.py`` -- every op is differentiable exactly like the module path and
honours ``requires_grad=False`` inputs (``FreezeParameters``, dreamer/utils/module.py:17-38).  Importing this module
registers them (``import dreamer_pytorch_b200.ops``).

    transition_step          RSSMTransition.forward                      rnns.py:98-107
    transition_rollout       RSSMRollout.rollout_transition              rnns.py:207-222
    representation_rollout   RSSMRollout.rollout_representation          rnns.py:179-205
    policy_rollout           RSSMRollout.rollout_policy + AgentModel.policy + ActionDecoder   rnns.py:224-244, agent.py:101-117
    mlp                      DenseModel / ActionDecoder feed-forward     dense.py:28-43, action.py:37-52
    kl_entropy               KL(post || prior), free nats, entropies     dreamer_algo.py:270-273,342-343
    heads_returns_loss       lambda-returns, actor / value loss + grads  dreamer_algo.py:314-337,398-422
    image_logprob            Independent(Normal(mean, 1), 3).log_prob    observation.py:117, dreamer_algo.py:264

Tensor lists follow the C-ABI order: ``rssm`` = the 10 transition tensors of ``RSSMTransition.weights()`` (+ the 4
posterior tensors for representation_rollout), ``actor`` / ``weights`` = [W0, b0, W1, b1, ...].  Noise tensors are the
injected-noise inputs of the kernels (N(0,1); Gumbel(0,1) for the categorical policies) or an int64 {seed, offset} pair
(``noise.next_pair``) to have the fp32 kernels generate them.
"""
from __future__ import annotations

from typing import List, Optional

import torch

from . import _lib
from .functional import ImageLogProbFn, KLFn, MlpFn, RolloutCfg, RolloutFn, returns_loss_raw

_LIB = torch.library.Library("dreamer_b200", "DEF")


def _dims(rssm: List[torch.Tensor], embed: int = 0):
    # W_in [Hd, A + S], W_hh [3 D, D], W_p2 [2 S, Hd]  (drm_rssm order: W_in b_in W_ih W_hh b_ih b_hh W_p1 b_p1 W_p2 b_p2)
    Hd, D, S = rssm[0].shape[0], rssm[3].shape[1], rssm[8].shape[0] // 2
    return (rssm[0].shape[1] - S, S, D, Hd, embed)


def _pad14(rssm: List[torch.Tensor]):
    return list(rssm) + [None] * (14 - len(rssm))


def transition_rollout(stoch, deter, action, eps_prior, rssm: List[torch.Tensor]) -> List[torch.Tensor]:
    steps, n = action.shape[0], action.shape[1]
    cfg = RolloutCfg(_lib.MODE_TRANSITION, steps, n, _dims(rssm))
    return list(RolloutFn.apply(cfg, stoch, deter, action, None, eps_prior, None, None, *_pad14(rssm[:10])))


def transition_step(prev_action, stoch, deter, eps, rssm: List[torch.Tensor]) -> List[torch.Tensor]:
    out = transition_rollout(stoch, deter, prev_action.unsqueeze(0), eps if eps.dtype == torch.int64 else eps.unsqueeze(0),
                             rssm)
    return [x.squeeze(0) for x in out]


def representation_rollout(stoch, deter, action, obs_embed, eps_prior, eps_post, rssm: List[torch.Tensor]) -> List[torch.Tensor]:
    steps, n = action.shape[0], action.shape[1]
    cfg = RolloutCfg(_lib.MODE_REPRESENTATION, steps, n, _dims(rssm, obs_embed.shape[-1]))
    return list(RolloutFn.apply(cfg, stoch, deter, action, obs_embed, eps_prior, eps_post, None, *rssm[:14]))


def policy_rollout(stoch, deter, noise_act, eps_prior, rssm: List[torch.Tensor], actor: List[torch.Tensor], act_dist: str,
                   steps: int, mean_scale: float = 5.0, raw_init_std: float = 4.993239, min_std: float = 0.0001) -> List[torch.Tensor]:
    # (defaults mirror the schema: the dispatcher drops trailing arguments that equal their schema default)
    cfg = RolloutCfg(_lib.MODE_POLICY, steps, stoch.shape[0], _dims(rssm), act_dist=act_dist, mean_scale=mean_scale,
                     raw_init_std=raw_init_std, act_min_std=min_std)
    return list(RolloutFn.apply(cfg, stoch.detach(), deter.detach(), None, None, eps_prior, None, noise_act,
                                *_pad14(rssm[:10]), *actor))


def mlp(x0, x1: Optional[torch.Tensor], weights: List[torch.Tensor]) -> torch.Tensor:
    return MlpFn.apply(x0, x1, *weights)


def kl_entropy(post_mean, post_std, prior_mean, prior_std, free_nats: float) -> List[torch.Tensor]:
    return list(KLFn.apply(post_mean, post_std, prior_mean, prior_std, free_nats))


def heads_returns_loss(reward, value, pcont: Optional[torch.Tensor], discount: float, lambda_: float) -> List[torch.Tensor]:
    """[losses (actor, value), returns, disc, d_reward, d_value, d_value_pred, d_pcont (empty without pcont)]: the loss
    kernel produces the analytic gradients w.r.t. the head outputs in the same launch; nothing here is differentiable."""
    o = returns_loss_raw(reward.detach().float().contiguous(), value.detach().float().contiguous(),
                         None if pcont is None else pcont.detach().float().contiguous(), discount, lambda_)
    d_pc = o["d_pcont"] if o["d_pcont"] is not None else reward.new_zeros(0)
    return [o["losses"], o["returns"], o["disc"], o["d_reward"], o["d_value"], o["d_value_pred"], d_pc]


def image_logprob(mean, obs) -> torch.Tensor:
    return ImageLogProbFn.apply(mean, obs)


_SCHEMAS = [
    ("transition_step(Tensor prev_action, Tensor stoch, Tensor deter, Tensor eps, Tensor[] rssm) -> Tensor[]", transition_step),
    ("transition_rollout(Tensor stoch, Tensor deter, Tensor action, Tensor eps_prior, Tensor[] rssm) -> Tensor[]",
     transition_rollout),
    ("representation_rollout(Tensor stoch, Tensor deter, Tensor action, Tensor obs_embed, Tensor eps_prior, Tensor eps_post, "
     "Tensor[] rssm) -> Tensor[]", representation_rollout),
    ("policy_rollout(Tensor stoch, Tensor deter, Tensor noise_act, Tensor eps_prior, Tensor[] rssm, Tensor[] actor, "
     "str act_dist, int steps, float mean_scale=5.0, float raw_init_std=4.993239, float min_std=0.0001) -> Tensor[]",
     policy_rollout),
    ("mlp(Tensor x0, Tensor? x1, Tensor[] weights) -> Tensor", mlp),
    ("kl_entropy(Tensor post_mean, Tensor post_std, Tensor prior_mean, Tensor prior_std, float free_nats) -> Tensor[]",
     kl_entropy),
    ("heads_returns_loss(Tensor reward, Tensor value, Tensor? pcont, float discount, float lambda_) -> Tensor[]",
     heads_returns_loss),
    ("image_logprob(Tensor mean, Tensor obs) -> Tensor", image_logprob),
]
for _schema, _fn in _SCHEMAS:
    _LIB.define(_schema)
    _LIB.impl(_schema.split("(")[0], _fn, "CompositeImplicitAutograd")

OP_NAMES = [s.split("(")[0] for s, _ in _SCHEMAS]
