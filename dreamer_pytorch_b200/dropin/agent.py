"""Drop-in for ``dreamer/models/agent.py`` (agent.py:25-191): same constructor, attribute names
and ``state_dict`` keys, so ``dreamer/algos/dreamer_algo.py`` and the rlpyt agents run unchanged.
"""
import numpy as np
import torch
import torch.nn as nn

from .. import _lib, noise
from ..functional import RolloutCfg, RolloutFn
from ._compat import buffer_func, infer_leading_dims, namedarraytuple, restore_leading_dims
from .action import ActionDecoder
from .dense import DenseModel
from .observation import ObservationDecoder, ObservationEncoder
from .rnns import RSSMRepresentation, RSSMRollout, RSSMState, RSSMTransition, get_feat


class AgentModel(nn.Module):
    def __init__(self, action_shape, stochastic_size=30, deterministic_size=200, hidden_size=200,
                 image_shape=(3, 64, 64), action_hidden_size=200, action_layers=3, action_dist="one_hot",
                 reward_shape=(1,), reward_layers=3, reward_hidden=300, value_shape=(1,), value_layers=3,
                 value_hidden=200, dtype=torch.float, use_pcont=False, pcont_layers=3, pcont_hidden=200, **kwargs):
        super().__init__()
        self.observation_encoder = ObservationEncoder(shape=image_shape)
        encoder_embed_size = self.observation_encoder.embed_size
        feature_size = stochastic_size + deterministic_size
        self.observation_decoder = ObservationDecoder(embed_size=feature_size, shape=image_shape)
        self.action_shape = action_shape
        output_size = np.prod(action_shape)  # int or tuple (agent.py:56)
        self.transition = RSSMTransition(output_size, stochastic_size, deterministic_size, hidden_size)
        self.representation = RSSMRepresentation(self.transition, encoder_embed_size, output_size, stochastic_size,
                                                 deterministic_size, hidden_size)
        self.rollout = RSSMRollout(self.representation, self.transition)
        self.action_size = output_size
        self.action_dist = action_dist
        self.action_decoder = ActionDecoder(output_size, feature_size, action_hidden_size, action_layers, action_dist)
        self.reward_model = DenseModel(feature_size, reward_shape, reward_layers, reward_hidden)
        self.value_model = DenseModel(feature_size, value_shape, value_layers, value_hidden)
        self.dtype = dtype
        self.stochastic_size = stochastic_size
        self.deterministic_size = deterministic_size
        if use_pcont:
            self.pcont = DenseModel(feature_size, (1,), pcont_layers, pcont_hidden, dist="binary")

    # ------------------------------------------------------------------ policy (agent.py:101-117)
    def policy(self, state: RSSMState):
        feat = get_feat(state)
        action_dist = self.action_decoder(feat)
        if self.action_dist == "tanh_normal":
            action = action_dist.rsample() if self.training else action_dist.mode()
        elif self.action_dist == "one_hot":
            action = action_dist.sample()
            action = action + action_dist.probs - action_dist.probs.detach()  # straight-through
        elif self.action_dist == "relaxed_one_hot":
            action = action_dist.rsample()
        else:
            action = action_dist.sample()
        return action, action_dist

    def can_fuse_policy(self) -> bool:
        """The imagination kernel covers: tanh_normal in training mode (rsample), one_hot
        (Gumbel-max sample + straight-through, both modes) and relaxed_one_hot."""
        if self.action_dist == "tanh_normal":
            return self.training
        return self.action_dist in ("one_hot", "relaxed_one_hot")

    def fused_policy_rollout(self, steps: int, state: RSSMState):
        """``RSSMRollout.rollout_policy(steps, self.policy, state)`` (rnns.py:224-244) as one launch.
        ``state`` is already detached by the caller."""
        dec = self.action_decoder
        n = state.stoch.shape[0]
        dev = state.stoch.device
        A = int(self.action_size)
        cfg = RolloutCfg(_lib.MODE_POLICY, steps, n, self.transition.dims(), act_dist=self.action_dist,
                         mean_scale=float(dec.mean_scale), raw_init_std=float(dec.raw_init_std),
                         act_min_std=float(dec.min_std))
        draw = noise.normal if self.action_dist == "tanh_normal" else noise.gumbel
        noise_act = draw("noise_act", (steps, n, A), dev, lazy=True)
        eps = noise.normal("eps_prior", (steps, n, self.stochastic_size), dev, lazy=True,
                           pair=noise_act if noise.is_pair(noise_act) else None)
        mean, std, stoch, deter, action = RolloutFn.apply(
            cfg, state.stoch, state.deter, None, None, eps, None, noise_act,
            *self.transition.weights(), None, None, None, None, *dec.weights())
        return RSSMState(mean, std, stoch, deter), action

    # ------------------------------------------------------------------ single steps (agent.py:119-155)
    def get_state_representation(self, observation: torch.Tensor, prev_action: torch.Tensor = None,
                                 prev_state: RSSMState = None):
        obs_embed = self.observation_encoder(observation)
        if prev_action is None:
            prev_action = torch.zeros(observation.size(0), int(self.action_size), device=observation.device,
                                      dtype=observation.dtype)
        if prev_state is None:
            prev_state = self.representation.initial_state(prev_action.size(0), device=prev_action.device,
                                                           dtype=prev_action.dtype)
        _, state = self.representation(obs_embed, prev_action, prev_state)
        return state

    def get_state_transition(self, prev_action: torch.Tensor, prev_state: RSSMState):
        return self.transition(prev_action, prev_state)

    def forward(self, observation: torch.Tensor, prev_action: torch.Tensor = None, prev_state: RSSMState = None):
        # agent.py:157-164: the effective base-class forward raises
        raise NotImplementedError()


class AtariDreamerModel(AgentModel):
    """agent.py:167-188: sampling-time single step."""

    def forward(self, observation: torch.Tensor, prev_action: torch.Tensor = None, prev_state: RSSMState = None):
        lead_dim, T, B, img_shape = infer_leading_dims(observation, 3)
        observation = observation.reshape(T * B, *img_shape).type(self.dtype) / 255.0 - 0.5
        prev_action = prev_action.reshape(T * B, -1).to(self.dtype)
        if prev_state is None:
            prev_state = self.representation.initial_state(prev_action.size(0), device=prev_action.device,
                                                           dtype=self.dtype)
        state = self.get_state_representation(observation, prev_action, prev_state)
        action, action_dist = self.policy(state)
        return_spec = ModelReturnSpec(action, state)
        return buffer_func(return_spec, restore_leading_dims, lead_dim, T, B)


ModelReturnSpec = namedarraytuple("ModelReturnSpec", ["action", "state"])
