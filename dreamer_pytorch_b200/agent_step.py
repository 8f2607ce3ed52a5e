"""Environment-time agent step as ONE CUDA-graph replay.

``AtariDreamerModel.forward`` (agent.py:167-188) is called once per environment step (B = 1, no grad, 1000x between
two training phases, dreamer_agent.py:44-61): normalise the frame, conv encoder, one posterior step of the RSSM, actor
MLP, and -- in eval mode -- ``SampleDist.mode()`` (100 draws, log-probs, arg-max).  In the reference that is ~60 small
launches plus distribution-validation host syncs; here it is the drop-in model's own forward (one fused RSSM step kernel,
one MLP kernel, one ``drm_policy_mode`` kernel, cuDNN for the four convolutions) captured once and replayed:

    g = EnvStepGraph(model)                       # model.eval() or .train(): the mode is captured
    action, state = g.step(observation_u8, prev_action)      # recurrent state carried on the device
    g.reset()                                     # new episode

Pure latency work: nothing here is timed against a roofline.
"""
from __future__ import annotations

import torch

from .dropin.rnns import RSSMState


class EnvStepGraph:
    def __init__(self, model, batch: int = 1, warmup: int = 3):
        p0 = next(model.parameters())
        if not p0.is_cuda:
            raise RuntimeError("EnvStepGraph needs a CUDA model (there is no CPU fallback)")
        dev = p0.device
        self.model, self.batch = model, int(batch)
        shape = tuple(model.observation_encoder.shape)
        A = int(model.action_size)
        self.obs = torch.zeros(self.batch, *shape, dtype=torch.uint8, device=dev)
        self.prev_action = torch.zeros(self.batch, A, device=dev)
        self.prev = model.representation.initial_state(self.batch, device=dev, dtype=torch.float32)
        cur = torch.cuda.current_stream(dev)
        side = torch.cuda.Stream(device=dev)
        side.wait_stream(cur)
        with torch.cuda.stream(side), torch.no_grad():
            for _ in range(max(1, warmup)):
                self._eager()
        cur.wait_stream(side)
        torch.cuda.synchronize(dev)
        self.reset()
        self.graph = torch.cuda.CUDAGraph()
        with torch.no_grad(), torch.cuda.graph(self.graph):
            self.action, self.state = self._eager()
            # carry the recurrent state on the device: the next replay starts from it
            for dst, src in zip(self.prev, self.state):
                dst.copy_(src)
        self.reset()

    def _eager(self):
        out = self.model(self.obs, self.prev_action, self.prev)
        return out.action, RSSMState(*out.state)

    def reset(self):
        """Zero recurrent state and previous action (start of an episode, agents/base.py reset())."""
        for f in self.prev:
            f.zero_()
        self.prev_action.zero_()
        self._fresh = True

    @torch.no_grad()
    def step(self, observation: torch.Tensor, prev_action: torch.Tensor = None, prev_state: RSSMState = None):
        """observation: uint8 [batch, C, H, W] (any device; pinned host memory copies asynchronously).  prev_action: the
        action actually taken (after exploration noise), default: the graph's own last action.  prev_state: default the
        state carried on the device.  Returns (action [batch, A], RSSMState) -- static tensors, overwritten by the next step."""
        self.obs.copy_(observation.reshape(self.obs.shape), non_blocking=True)
        if prev_action is not None:
            self.prev_action.copy_(prev_action.reshape(self.prev_action.shape), non_blocking=True)
        elif not self._fresh:
            self.prev_action.copy_(self.action.reshape(self.prev_action.shape))
        self._fresh = False
        if prev_state is not None:
            for dst, src in zip(self.prev, prev_state):
                dst.copy_(src.reshape(dst.shape), non_blocking=True)
        self.graph.replay()
        return self.action, self.state
