"""Driver for the ncu capture of the HBM-bound kernels: one whole update (replay gather, dreamer_loss with the conv stack,
3 backwards, clip + Adam of the three groups), walker sizes, eager launches."""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from bench import WORKLOADS, build_model  # noqa: E402
from dreamer_pytorch_b200.algo import dreamer_loss  # noqa: E402
from dreamer_pytorch_b200.dropin.module import get_parameters  # noqa: E402
from dreamer_pytorch_b200.optim import ClipAdam  # noqa: E402
from dreamer_pytorch_b200.precision import set_precision  # noqa: E402

reps = int(sys.argv[1]) if len(sys.argv) > 1 else 2
dev = torch.device("cuda:0")
set_precision("bf16")
w = dict(WORKLOADS["walker"])
model = build_model(w, dev)
g = torch.Generator().manual_seed(5)
T = B = 50
obs = torch.randint(0, 256, (T, B, 3, 64, 64), generator=g, dtype=torch.uint8).to(dev).float() / 255.0 - 0.5
act = (torch.rand(T, B, 6, generator=g) * 2 - 1).to(dev)
rew = torch.randn(T, B, 1, generator=g).to(dev)
mods = [model.observation_encoder, model.observation_decoder, model.reward_model, model.representation, model.transition]
opts = [ClipAdam(get_parameters(mods), lr=6e-4, grad_clip=100.0), ClipAdam(list(model.action_decoder.parameters()), lr=8e-5, grad_clip=100.0),
        ClipAdam(list(model.value_model.parameters()), lr=8e-5, grad_clip=100.0)]
for _ in range(reps):
    model.zero_grad(set_to_none=True)
    ml, al, vl, _ = dreamer_loss(model, obs, act, rew, None, horizon=15)
    ml.backward(); al.backward(); vl.backward()
    for o in opts:
        o.step()
torch.cuda.synchronize()
print("ok")
