"""Drop-in for ``dreamer/models/distribution.py`` (distribution.py:7-44).

``SampleDist`` wraps the tanh-Normal action distribution.  The training path only calls
``.rsample()`` (forwarded to the wrapped distribution; in the fused imagination kernel the same
formula ``tanh(mean + std * eps)`` runs on-chip).  ``mean/mode/entropy`` are the reference's
Monte-Carlo estimates over ``samples`` draws and are used at environment-step time only (B=1).
"""
import torch


class SampleDist:
    def __init__(self, dist: torch.distributions.Distribution, samples=100, raw=None, params=None):
        """raw / params (optional, set by ActionDecoder): the actor MLP's output [..., 2A] and (mean_scale, raw_init_std,
        min_std) the distribution was built from.  With them ``mode()`` on CUDA tensors outside autograd is ONE fused
        kernel (drm_policy_mode: draws, log-probs, arg-max, gather) instead of ~15 torch.distributions launches."""
        self._dist = dist
        self._samples = samples
        self._raw, self._params = raw, params

    @property
    def name(self):
        return "SampleDist"

    def __getattr__(self, name):
        return getattr(self._dist, name)

    def _draw(self):
        wide = self._dist.expand((self._samples, *self._dist.batch_shape))
        return wide, wide.rsample()

    def mean(self):
        _, sample = self._draw()
        return sample.mean(0)

    def mode(self):
        raw = self._raw
        if raw is not None and raw.is_cuda and not (torch.is_grad_enabled() and raw.requires_grad):
            from ..functional import policy_mode

            return policy_mode(raw.detach(), self._samples, *self._params)
        wide, sample = self._draw()
        logprob = wide.log_prob(sample)  # [samples, batch]
        best = logprob.argmax(dim=0)     # [batch]
        idx = best.reshape(1, -1, 1).expand(1, sample.size(1), sample.size(2))
        return sample.gather(0, idx).squeeze(0)

    def entropy(self):
        wide, sample = self._draw()
        return -wide.log_prob(sample).mean(0)

    def sample(self):
        return self._dist.sample()
