"""Drop-in for ``dreamer/models/dense.py`` (dense.py:7-50): reward / value / pcont heads.

The ``nn.Sequential`` of Linear+ELU layers keeps the reference's parameter names
(``model.{0,2,4,6}``) but runs as ONE fused CUDA kernel over all rows (``drm_mlp_fwd``), with a
hand-written backward (``drm_mlp_bwd``).  The returned object is a real ``torch.distributions``
distribution built from the kernel output, as ``dreamer_algo.py:262-264,306-311,335-336`` expects.
"""
import numpy as np
import torch
import torch.distributions as td
import torch.nn as nn

from ..functional import MlpFn


def mlp_weights(seq: nn.Sequential):
    out = []
    for m in seq:
        if isinstance(m, nn.Linear):
            out += [m.weight, m.bias]
    return out


class DenseModel(nn.Module):
    def __init__(self, feature_size: int, output_shape: tuple, layers: int, hidden_size: int, dist="normal",
                 activation=nn.ELU):
        super().__init__()
        if activation is not nn.ELU:
            raise NotImplementedError("the fused MLP kernel implements activation=nn.ELU")
        self._output_shape = output_shape
        self._layers = layers
        self._hidden_size = hidden_size
        self._dist = dist
        self.activation = activation
        self._feature_size = feature_size
        self.model = self.build_model()

    def build_model(self):
        mods = [nn.Linear(self._feature_size, self._hidden_size), self.activation()]
        for _ in range(self._layers - 1):
            mods += [nn.Linear(self._hidden_size, self._hidden_size), self.activation()]
        mods += [nn.Linear(self._hidden_size, int(np.prod(self._output_shape)))]
        return nn.Sequential(*mods)

    def forward(self, features):
        lead = features.shape[:-1]
        flat = features.reshape(-1, features.shape[-1])
        dist_inputs = MlpFn.apply(flat, None, *mlp_weights(self.model))
        reshaped = dist_inputs.reshape(lead + tuple(self._output_shape))
        if self._dist == "normal":
            # scale as a 0-dim device tensor: a Python scalar would be turned into a tensor through a host->device copy,
            # which cannot be captured into a CUDA graph (graph.UpdateGraph)
            return td.independent.Independent(td.Normal(reshaped, reshaped.new_ones(()), validate_args=False),
                                              len(self._output_shape), validate_args=False)
        if self._dist == "binary":
            # validate_args=False: the reference feeds the soft target discount*(1-done) to log_prob
            # (dreamer_algo.py:266-269), which torch >= 1.8 rejects under argument validation
            return td.independent.Independent(td.Bernoulli(logits=reshaped, validate_args=False),
                                              len(self._output_shape), validate_args=False)
        raise NotImplementedError(self._dist)
