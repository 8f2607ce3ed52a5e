"""Conv encoder / decoder with the reference's parameter names and shapes
(``dreamer/models/observation.py:7-136``), SURVEY.md 8(f) N2.

The dense convolutions stay library (cuDNN) calls -- plain library GEMM-class work -- but are laid out for the B200
(``precision.get_conv_mode``): under ``precision("bf16")`` they run on channels-last tensors (tensor-core implicit
GEMMs, no NCHW<->NHWC conversion kernels around every layer), optionally under bf16 autocast (``conv_mode("bf16")``);
under ``precision("fp32")`` they are the reference's fp32 NCHW calls, bit for bit.  The decoder's distribution is ``Independent(Normal(mean, 1), 3)`` as in the reference, with ``log_prob`` routed to
ONE fused reduction kernel per call (``drm_image_logprob_fwd/bwd``) instead of ATen's chain of pointwise kernels over
[T*B, 3, 64, 64] (dreamer_algo.py:264)."""
import numpy as np
import torch
import torch.distributions as td
import torch.nn as nn

import contextlib

from ..precision import get_conv_mode


def _mode(x: torch.Tensor) -> str:
    return get_conv_mode() if x.is_cuda else "nchw"


def _cast(mode: str):
    return torch.autocast("cuda", dtype=torch.bfloat16) if mode == "bf16" else contextlib.nullcontext()


class ImageNormal(td.Independent):
    """``td.Independent(td.Normal(mean, 1), k)`` (observation.py:117) whose ``log_prob`` on CUDA tensors is the fused
    reduction kernel.  ``mean`` may be bf16 (autocast decoder): ``.mean`` widens it."""

    def __init__(self, mean: torch.Tensor, event_dims: int):
        super().__init__(td.Normal(mean, mean.new_ones(()), validate_args=False), event_dims, validate_args=False)
        self._loc, self._k = mean, event_dims

    @property
    def mean(self):
        return self._loc.float()

    def log_prob(self, value):
        loc = self._loc
        if loc.is_cuda and value.is_cuda and value.shape == loc.shape and loc.dim() > self._k:
            from ..functional import ImageLogProbFn

            lead = loc.shape[:loc.dim() - self._k]
            D = int(np.prod(loc.shape[loc.dim() - self._k:]))
            if D % 4 == 0:
                return ImageLogProbFn.apply(loc.reshape(-1, D), value.reshape(-1, D)).reshape(lead)
        return super().log_prob(value)


def _run_conv_stack(mods, x):
    """A Sequential of Conv2d / ConvTranspose2d (+ ReLU) on a channels-last fp32 tensor: the convolution is called without
    bias and bias + ReLU run as ONE in-place pass of this library (functional.BiasActFn), where the layer qualifies
    (fp32, channels a multiple of 4: every layer but the decoder's 3-channel output); anything else runs as the module."""
    import torch.nn.functional as F

    from ..functional import BiasActFn

    mods = list(mods)
    i = 0
    while i < len(mods):
        m = mods[i]
        relu = i + 1 < len(mods) and type(mods[i + 1]) is nn.ReLU
        y = None
        bias = getattr(m, "bias", None)
        if bias is not None and bias.shape[0] % 4 == 0 and x.dtype == torch.float32 and not torch.is_autocast_enabled():
            if type(m) is nn.Conv2d and m.padding_mode == "zeros":
                y = F.conv2d(x, m.weight, None, m.stride, m.padding, m.dilation, m.groups)
            elif type(m) is nn.ConvTranspose2d:
                y = F.conv_transpose2d(x, m.weight, None, m.stride, m.padding, m.output_padding, m.groups, m.dilation)
        if y is not None and BiasActFn.eligible(y, m.bias):
            x = BiasActFn.apply(y, m.bias, relu)
            i += 2 if relu else 1
        elif y is not None:
            x = y + m.bias.view(1, -1, 1, 1)
            i += 1
        else:
            x = m(x)
            i += 1
    return x


def conv_out(h_in, padding, kernel_size, stride):
    return int((h_in + 2.0 * padding - (kernel_size - 1.0) - 1.0) / stride + 1.0)


def output_padding(h_in, conv_out_, padding, kernel_size, stride):
    return h_in - (conv_out_ - 1) * stride + 2 * padding - (kernel_size - 1) - 1


def conv_out_shape(h_in, padding, kernel_size, stride):
    return tuple(conv_out(x, padding, kernel_size, stride) for x in h_in)


def output_padding_shape(h_in, conv_out_, padding, kernel_size, stride):
    return tuple(output_padding(h_in[i], conv_out_[i], padding, kernel_size, stride) for i in range(len(h_in)))


class ObservationEncoder(nn.Module):
    def __init__(self, depth=32, stride=2, shape=(3, 64, 64), activation=nn.ReLU):
        super().__init__()
        chans = [shape[0], depth, 2 * depth, 4 * depth, 8 * depth]
        mods = []
        for i in range(4):
            mods += [nn.Conv2d(chans[i], chans[i + 1], 4, stride), activation()]
        self.convolutions = nn.Sequential(*mods)
        self.shape, self.stride, self.depth = shape, stride, depth

    def forward(self, obs):
        lead = obs.shape[:-3]
        x = obs.reshape(-1, *obs.shape[-3:])
        mode = _mode(x)
        if mode != "nchw":
            if not getattr(self, "_cl", False):
                self.convolutions.to(memory_format=torch.channels_last)  # in place: same Parameter objects
                self._cl = True
            with _cast(mode):
                x = _run_conv_stack(self.convolutions, x.contiguous(memory_format=torch.channels_last))
            return x.float().reshape(*lead, -1)  # logical NCHW order, like the reference's flatten
        x = self.convolutions(x)
        return x.reshape(*lead, -1)

    @property
    def embed_size(self):
        hw = self.shape[1:]
        for _ in range(4):
            hw = conv_out_shape(hw, 0, 4, self.stride)
        return 8 * self.depth * int(np.prod(hw))


class ObservationDecoder(nn.Module):
    def __init__(self, depth=32, stride=2, activation=nn.ReLU, embed_size=1024, shape=(3, 64, 64)):
        super().__init__()
        self.depth, self.shape = depth, shape
        c, h, w = shape
        kernels = [6, 6, 5, 5]  # conv1..conv4 as seen from the image side
        shapes, pads, cur = [], [], (h, w)
        for k in kernels:
            nxt = conv_out_shape(cur, 0, k, stride)
            pads.append(output_padding_shape(cur, nxt, 0, k, stride))
            shapes.append(nxt)
            cur = nxt
        self.conv_shape = (32 * depth, *shapes[3])
        self.linear = nn.Linear(embed_size, 32 * depth * int(np.prod(shapes[3])))
        chans = [32 * depth, 4 * depth, 2 * depth, depth, c]
        mods = []
        for i, j in enumerate((3, 2, 1, 0)):
            mods.append(nn.ConvTranspose2d(chans[i], chans[i + 1], kernels[j], stride, output_padding=pads[j]))
            if i < 3:
                mods.append(activation())
        self.decoder = nn.Sequential(*mods)

    def forward(self, x):
        lead = x.shape[:-1]
        flat = x.reshape(int(np.prod(lead)), x.shape[-1])
        mode = _mode(flat)
        if mode != "nchw":
            if not getattr(self, "_cl", False):
                self.decoder.to(memory_format=torch.channels_last)
                self._cl = True
            with _cast(mode):
                y = self.linear(flat).reshape(flat.shape[0], *self.conv_shape)
                mean = _run_conv_stack(self.decoder, y.contiguous(memory_format=torch.channels_last))
            mean = mean.contiguous().reshape(*lead, *self.shape)  # NCHW order (bf16 or fp32): what the log-prob kernel reads
        else:
            y = self.linear(flat).reshape(flat.shape[0], *self.conv_shape)
            mean = self.decoder(y).reshape(*lead, *self.shape)
        return ImageNormal(mean, len(self.shape))
