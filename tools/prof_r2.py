"""Driver for the round-2 ncu captures: `reps` whole RSSM updates (dreamer_loss on a synthetic obs_embed: observation
rollout fwd/bwd on the fp32 kernels, imagination rollout fwd/bwd + heads on the bf16 tcgen05 kernels), walker sizes."""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from bench import WORKLOADS, build_model  # noqa: E402
from dreamer_pytorch_b200.algo import dreamer_loss  # noqa: E402
from dreamer_pytorch_b200.precision import set_precision  # noqa: E402

reps = int(sys.argv[1]) if len(sys.argv) > 1 else 2
dev = torch.device("cuda:0")
set_precision("bf16")
w = dict(WORKLOADS["walker"])
model = build_model(w, dev)
g = torch.Generator().manual_seed(5)
T = B = 50
embed = torch.randn(T, B, 1024, generator=g).to(dev)
act = (torch.rand(T, B, 6, generator=g) * 2 - 1).to(dev)
rew = torch.randn(T, B, 1, generator=g).to(dev)
for _ in range(reps):
    model.zero_grad(set_to_none=True)
    ml, al, vl, _ = dreamer_loss(model, None, act, rew, None, horizon=15, obs_embed=embed)
    ml.backward(); al.backward(); vl.backward()
torch.cuda.synchronize()
print("ok")
