"""Conv encoder / decoder with the reference's parameter names and shapes
(``dreamer/models/observation.py:7-136``).  OUT OF SCOPE as kernels:
dense convolutions stay library (cuDNN) calls; they only feed ``obs_embed`` into the RSSM path and
consume ``get_feat(post)``.  Present so that ``AgentModel`` is self-contained on a box without the
reference checkout."""
import numpy as np
import torch
import torch.distributions as td
import torch.nn as nn


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
        x = self.convolutions(obs.reshape(-1, *obs.shape[-3:]))
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
        y = self.linear(flat).reshape(flat.shape[0], *self.conv_shape)
        mean = self.decoder(y).reshape(*lead, *self.shape)
        return td.Independent(td.Normal(mean, 1), len(self.shape))
