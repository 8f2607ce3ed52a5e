"""Small driver for ncu captures of the bf16 tensor-core kernels: one batched-MLP forward/backward over
148 row tiles and one imagination rollout (fwd + bwd) over 2 500 start states x 15 steps."""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from dreamer_pytorch_b200 import _lib  # noqa: E402
from dreamer_pytorch_b200.functional import RolloutCfg, RolloutFn, MlpFn  # noqa: E402
from dreamer_pytorch_b200.precision import set_precision  # noqa: E402

dev = torch.device("cuda:0")
set_precision("bf16")
torch.manual_seed(0)
reps = int(sys.argv[1]) if len(sys.argv) > 1 else 2

# ---- MLP (value head shape) over 148 tiles
M = 148 * 128
ws = []
d = 230
for _ in range(3):
    ws += [torch.randn(200, d, device=dev) / d ** 0.5, torch.zeros(200, device=dev)]
    d = 200
ws += [torch.randn(1, d, device=dev) / d ** 0.5, torch.zeros(1, device=dev)]
ws = [w.requires_grad_(True) for w in ws]
x0 = torch.randn(M, 30, device=dev, requires_grad=True)
x1 = torch.randn(M, 200, device=dev, requires_grad=True)
for _ in range(reps):
    y = MlpFn.apply(x0, x1, *ws)
    y.sum().backward()

# ---- imagination rollout, walker sizes
A, S, D, Hd, n, H = 6, 30, 200, 200, 2500, 15


def lin(o, i):
    return [torch.randn(o, i, device=dev) / i ** 0.5, torch.zeros(o, device=dev)]


w_in = lin(Hd, A + S)
rssm = [w_in[0], w_in[1], torch.randn(3 * D, Hd, device=dev) / Hd ** 0.5, torch.randn(3 * D, D, device=dev) / D ** 0.5,
        torch.zeros(3 * D, device=dev), torch.zeros(3 * D, device=dev)] + lin(Hd, D) + lin(2 * S, Hd)
actor = lin(200, S + D) + lin(200, 200) + lin(200, 200) + lin(2 * A, 200)
actor = [w.requires_grad_(True) for w in actor]
cfg = RolloutCfg(_lib.MODE_POLICY, H, n, (A, S, D, Hd, 0), act_dist="tanh_normal")
ss, sd = torch.randn(n, S, device=dev), torch.tanh(torch.randn(n, D, device=dev))
for _ in range(reps):
    outs = RolloutFn.apply(cfg, ss, sd, None, None, torch.randn(H, n, S, device=dev), None, torch.randn(H, n, A, device=dev),
                           *rssm, None, None, None, None, *actor)
    (outs[2].sum() + outs[3].sum()).backward()
torch.cuda.synchronize()
print("ok")
