"""Drop-in for ``dreamer/models/action.py`` (action.py:11-66).

Stand-alone calls run the feed-forward stack as one fused MLP kernel and wrap the result in the
same ``torch.distributions`` objects as the reference.  Inside ``rollout_policy`` the decoder is
fused into the imagination kernel instead (see ``AgentModel.fused_policy_rollout``).
"""
import numpy as np
import torch
import torch.distributions
import torch.nn as nn
import torch.nn.functional as F
from torch.distributions import TanhTransform as TanhBijector

from ..functional import MlpFn
from .dense import mlp_weights
from .distribution import SampleDist


class ActionDecoder(nn.Module):
    def __init__(self, action_size, feature_size, hidden_size, layers, dist="tanh_normal", activation=nn.ELU,
                 min_std=1e-4, init_std=5, mean_scale=5):
        super().__init__()
        if activation is not nn.ELU:
            raise NotImplementedError("the fused actor kernel implements activation=nn.ELU")
        self.action_size = int(action_size)
        self.feature_size = feature_size
        self.hidden_size = hidden_size
        self.layers = layers
        self.dist = dist
        self.activation = activation
        self.min_std = min_std
        self.init_std = init_std
        self.mean_scale = mean_scale
        self.feedforward_model = self.build_model()
        self.raw_init_std = np.log(np.exp(self.init_std) - 1)

    def build_model(self):
        mods = [nn.Linear(self.feature_size, self.hidden_size), self.activation()]
        for _ in range(1, self.layers):
            mods += [nn.Linear(self.hidden_size, self.hidden_size), self.activation()]
        if self.dist == "tanh_normal":
            mods += [nn.Linear(self.hidden_size, self.action_size * 2)]
        elif self.dist in ("one_hot", "relaxed_one_hot"):
            mods += [nn.Linear(self.hidden_size, self.action_size)]
        else:
            raise NotImplementedError(f"{self.dist} not implemented")
        return nn.Sequential(*mods)

    def weights(self):
        return mlp_weights(self.feedforward_model)

    def forward(self, state_features):
        lead = state_features.shape[:-1]
        flat = state_features.reshape(-1, state_features.shape[-1])
        x = MlpFn.apply(flat, None, *self.weights())
        x = x.reshape(*lead, x.shape[-1])
        if self.dist == "tanh_normal":
            mean, std = torch.chunk(x, 2, -1)
            mean = self.mean_scale * torch.tanh(mean / self.mean_scale)
            std = F.softplus(std + self.raw_init_std) + self.min_std
            # validate_args=False: torch's argument validation is a host synchronisation per distribution (SURVEY 8(b):
            # "no host sync inside"), and it cannot run under CUDA-graph capture (agent_step.EnvStepGraph)
            dist = torch.distributions.Normal(mean, std, validate_args=False)
            dist = torch.distributions.TransformedDistribution(dist, TanhBijector(), validate_args=False)
            dist = torch.distributions.Independent(dist, 1, validate_args=False)
            return SampleDist(dist, raw=x, params=(float(self.mean_scale), float(self.raw_init_std), float(self.min_std)))
        if self.dist == "one_hot":
            return torch.distributions.OneHotCategorical(logits=x, validate_args=False)
        if self.dist == "relaxed_one_hot":
            return torch.distributions.RelaxedOneHotCategorical(0.1, logits=x, validate_args=False)
        return None
