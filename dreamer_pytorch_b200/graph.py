"""One CUDA graph for the whole imagination step (actor + value loss, forward and backward).

The fused loss section launches ~35 small kernels around the two persistent rollout kernels (weight packing,
head MLPs, weight-gradient GEMMs, memsets).  At the reference batch (2 500 start states) the host cannot issue
them as fast as the GPU retires them, so the step is captured once -- kernels, the side-stream fork/join of
``algo._ActorLossFn.backward`` and the noise generation included -- and replayed: one launch per update, no
Python on the critical path, gradients land in static ``.grad`` tensors (the usual whole-step capture pattern).

    g = ImaginationGraph(model, n_states=2500, horizon=15)
    actor_loss, value_loss = g.step(stoch, deter)      # tensors [n_states, S] / [n_states, D], any device
    # model.action_decoder / model.value_model parameters now hold .grad; run the optimizers as usual

Replaces the Python-level sequence of dreamer_algo.py:283-337 + :174-180 (actor_loss.backward(),
value_loss.backward()) for callers that own the training loop; ``imagination_losses`` stays the eager API.
"""
from __future__ import annotations

import torch

from . import _lib
from .algo import imagination_losses
from .dropin.dense import mlp_weights
from .dropin.rnns import RSSMState


class ImaginationGraph:
    def __init__(self, model, n_states: int, horizon: int, discount: float = 0.99, discount_lambda: float = 0.95,
                 use_pcont: bool = False, warmup: int = 3, actor_opt=None, value_opt=None, allreduce: bool = False,
                 peer_allreduce: bool = True):
        """actor_opt / value_opt: optional ``optim.ClipAdam(..., capturable=True)`` over the action decoder's / the value
        model's parameters.  Their step() (clip_grad_norm_ + Adam, dreamer_algo.py:182-194) is then recorded behind the
        backward pass, so that one replay is one whole actor + value update.  Parameters and optimizer state are restored
        after the warm-up passes; single-process only (with several ranks the gradient all-reduce sits between the
        backward pass and the optimizers).
        allreduce: with torch.distributed initialised, average the two gradient buckets (actor, value: one flat fp32 buffer
        each, the tensors the backward kernels write into) over all ranks INSIDE the graph -- two NCCL all-reduces recorded
        behind the backward pass, no Python between the backward kernels and the collective; the optimizers, when given,
        are recorded behind them, so one replay is one whole data-parallel update.  ``allreduce_in_graph`` tells whether the
        collectives were captured (False: capture of NCCL failed on this build, the caller all-reduces after ``step``)."""
        p0 = next(model.parameters())
        if not p0.is_cuda:
            raise RuntimeError("ImaginationGraph needs a CUDA model (there is no CPU fallback)")
        self.model, self.n, self.horizon = model, int(n_states), int(horizon)
        self._kw = dict(discount=discount, discount_lambda=discount_lambda, use_pcont=use_pcont)
        dev = p0.device
        self.stoch = torch.zeros(self.n, model.stochastic_size, device=dev)
        self.deter = torch.zeros(self.n, model.deterministic_size, device=dev)
        self.actor_params = list(model.action_decoder.parameters())
        self.value_params = list(model.value_model.parameters())
        self.params = self.actor_params + self.value_params
        import torch.distributed as dist

        self._dist = dist if (allreduce and dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1) else None
        self.allreduce_in_graph = self._dist is not None
        self._coalesce = True
        self._peer, self.allreduce_kind = None, None
        if self._dist is not None and peer_allreduce:
            from .parallel import PeerBucket

            self._peer = PeerBucket.create([list(model.action_decoder.weights()), list(mlp_weights(model.value_model.model))])
        self.opts = [(o, ps) for o, ps in ((actor_opt, self.actor_params), (value_opt, self.value_params)) if o is not None]
        for o, ps in self.opts:
            if getattr(o, "step_dev", None) is None:
                raise ValueError("optimizers recorded into the graph must be ClipAdam(..., capturable=True)")
            if [id(p) for p in o.params] != [id(p) for p in ps]:
                raise ValueError("the optimizer must cover exactly the parameters of its module, in order")
        keep = [(p, p.detach().clone()) for p in self.params] if self.opts else []
        for o, _ in self.opts:  # warm-up passes step the optimizers: remember their state
            keep += [(t, t.clone()) for st in o.state.values() for t in st.values()] + [(o.step_dev, o.step_dev.clone())]
        # eager warm-up on a side stream (lazy initialisation, allocator high-water mark), as capture requires
        cur = torch.cuda.current_stream(dev)
        s = torch.cuda.Stream(device=dev)
        s.wait_stream(cur)
        with torch.cuda.stream(s):
            for _ in range(max(1, warmup)):
                self._eager()
        cur.wait_stream(s)
        torch.cuda.synchronize(dev)
        lib = _lib.load()
        try:
            self._capture(lib)
        except Exception:
            if self._dist is None:
                raise
            # this torch / NCCL build cannot record the collective: capture the step without it
            self._dist, self.allreduce_in_graph = None, False
            torch.cuda.synchronize(dev)
            self._capture(lib)
        for p, g in zip(self.params, self.grads):
            p.grad = g  # static tensors from the graph's pool, refreshed by every replay
        with torch.no_grad():
            for t, saved in keep:
                t.copy_(saved)

    def _capture(self, lib):
        self.graph = torch.cuda.CUDAGraph()
        n0 = lib.drm_launch_count()
        with torch.cuda.graph(self.graph):
            self.actor_loss, self.value_loss, self.grads = self._eager()
            self.losses = torch.stack([self.actor_loss, self.value_loss])  # [2], one device->host read per step
        self.launches_per_replay = int(lib.drm_launch_count() - n0)  # kernels of this library inside the graph

    def _eager(self):
        start = RSSMState(self.stoch, self.stoch, self.stoch, self.deter)  # mean / std of a start state are never read
        from .functional import bucket_provider

        # with a peer bucket, the backward kernels write both groups' weight gradients straight into peer-mapped memory
        # (the sinks are allocated as early as the forward pass: the PREPARE half of the backward runs beside it)
        with bucket_provider(self._peer.provide if self._peer is not None else None):
            actor_loss, value_loss, _ = imagination_losses(self.model, start, self.horizon, **self._kw)
            # autograd.grad rather than .backward(): no AccumulateGrad nodes (they remember the stream they were created
            # on, which breaks capture when the model has been used on the default stream before)
            ga = torch.autograd.grad(actor_loss, self.actor_params)
            gv = torch.autograd.grad(value_loss, self.value_params)
        grads = list(ga) + list(gv)
        if self._peer is not None and self._peer.owns(ga) and self._peer.owns(gv):
            # ... and one launch of this library's one-shot kernel averages it over NVLink (parallel.PeerBucket)
            self._peer.allreduce()
            self.allreduce_kind = "peer_oneshot"
        elif self._dist is not None:
            self.allreduce_kind = "nccl"
            # equal shards: the mean of the shard means is the global mean.  Each group's gradients are
            # views of one flat bucket written by the backward kernels themselves (functional.flat_like).
            # Both buckets go out as ONE grouped NCCL launch (ncclGroupStart / End through torch's coalescing manager):
            # at ~0.5 MB each the cost is launch + handshake latency, not bandwidth.
            bufs = []
            for group in (ga, gv):
                base = group[0]._base
                bufs += [base] if base is not None and all(g._base is base for g in group) else list(group)
            try:
                if self._coalesce:
                    with self._dist._coalescing_manager(device=bufs[0].device):
                        for b in bufs:
                            self._dist.all_reduce(b, op=self._dist.ReduceOp.AVG)
            except Exception:  # pragma: no cover  (a backend without the grouped fast path)
                if torch.cuda.is_current_stream_capturing():
                    raise
                self._coalesce = False
            if not self._coalesce:
                for b in bufs:
                    self._dist.all_reduce(b, op=self._dist.ReduceOp.AVG)
        if self.opts:
            for p, g in zip(self.params, grads):
                p.grad = g
            self.grad_norms = [o.step() for o, _ in self.opts]  # device scalars (norm before clipping), static
        return actor_loss.detach(), value_loss.detach(), grads

    def step(self, stoch: torch.Tensor, deter: torch.Tensor):
        """Copies the start states into the captured buffers (a host->device copy when they live in pinned host
        memory), replays the step and returns the (static) loss tensors; parameter gradients are refreshed in place."""
        if stoch.shape != self.stoch.shape or deter.shape != self.deter.shape:
            raise ValueError(f"start states must be {tuple(self.stoch.shape)} / {tuple(self.deter.shape)}")
        if stoch.data_ptr() != self.stoch.data_ptr():
            self.stoch.copy_(stoch, non_blocking=True)
        if deter.data_ptr() != self.deter.data_ptr():
            self.deter.copy_(deter, non_blocking=True)
        for p, g in zip(self.params, self.grads):
            if p.grad is not g:
                p.grad = g  # an optimizer's zero_grad(set_to_none=True) dropped it; the graph writes into g
        self.graph.replay()
        return self.actor_loss, self.value_loss


class UpdateGraph:
    """One whole Dreamer update -- ``Dreamer.loss`` + the three backwards + clip_grad_norm_ + Adam for the three optimizer
    groups (dreamer_algo.py:174-194,213-371) -- as ONE CUDA graph replay.

        g = UpdateGraph(model, batch_size=50, batch_length=50, horizon=15)
        model_loss, actor_loss, value_loss = g.step(samples)   # SamplesFromReplay-like: all_observation u8 [T+1,B,C,H,W],
                                                               # all_action [T+1,B,A], all_reward [T+1,B], done [T,B]

    The batch is copied into static buffers (device tensors from ``replay.DeviceSequenceReplay`` cost a device copy, pinned
    host tensors an asynchronous H2D copy), the graph is replayed, parameters are updated in place.  The frame
    normalisation (uint8 -> float / 255 - 0.5, dreamer_algo.py:241), the conv encoder / decoder (cuDNN), every fused kernel
    of this library, the noise generation and the optimizers are inside the graph; nothing synchronises with the host.
    Optimizers: ``optim.ClipAdam(capturable=True)`` with the reference's hyper-parameters (dreamer_algo.py:28-30,111-131:
    model 6e-4, actor / value 8e-5, grad_clip 100), created here unless passed in.
    """

    def __init__(self, model, batch_size: int, batch_length: int, horizon: int = 15, discount: float = 0.99,
                 discount_lambda: float = 0.95, free_nats: float = 3.0, kl_scale: float = 1.0, use_pcont: bool = False,
                 pcont_scale: float = 10.0, model_lr: float = 6e-4, actor_lr: float = 8e-5, value_lr: float = 8e-5,
                 grad_clip: float = 100.0, optimizers=None, warmup: int = 3):
        from .algo import dreamer_loss
        from .dropin.module import get_parameters
        from .optim import ClipAdam

        p0 = next(model.parameters())
        if not p0.is_cuda:
            raise RuntimeError("UpdateGraph needs a CUDA model (there is no CPU fallback)")
        dev = p0.device
        self.model, self._loss_fn = model, dreamer_loss
        T, B, A = int(batch_length), int(batch_size), int(model.action_size)
        self._kw = dict(horizon=horizon, discount=discount, discount_lambda=discount_lambda, free_nats=free_nats,
                        kl_scale=kl_scale, use_pcont=use_pcont, pcont_scale=pcont_scale)
        shape = tuple(model.observation_encoder.shape)
        self.all_observation = torch.zeros(T + 1, B, *shape, dtype=torch.uint8, device=dev)
        self.all_action = torch.zeros(T + 1, B, A, device=dev)
        self.all_reward = torch.zeros(T + 1, B, device=dev)
        self.done = torch.zeros(T, B, dtype=torch.bool, device=dev)
        # the reference's three groups (dreamer_algo.py:111-122); the world-model list keeps its duplicated transition tensors
        model_modules = [model.observation_encoder, model.observation_decoder, model.reward_model, model.representation,
                         model.transition] + ([model.pcont] if use_pcont else [])
        listed = get_parameters(model_modules)
        self.model_params = list({id(p): p for p in listed}.values())
        self.actor_params = list(model.action_decoder.parameters())
        self.value_params = list(model.value_model.parameters())
        if optimizers is None:
            optimizers = (ClipAdam(listed, lr=model_lr, grad_clip=grad_clip, capturable=True),
                          ClipAdam(self.actor_params, lr=actor_lr, grad_clip=grad_clip, capturable=True),
                          ClipAdam(self.value_params, lr=value_lr, grad_clip=grad_clip, capturable=True))
        self.model_opt, self.actor_opt, self.value_opt = optimizers
        for o in optimizers:
            if getattr(o, "step_dev", None) is None:
                raise ValueError("optimizers recorded into the graph must be ClipAdam(..., capturable=True)")
        params = self.model_params + self.actor_params + self.value_params
        keep = [(p, p.detach().clone()) for p in params]
        for o in optimizers:
            keep += [(t, t.clone()) for st in o.state.values() for t in st.values()] + [(o.step_dev, o.step_dev.clone())]
        cur = torch.cuda.current_stream(dev)
        side = torch.cuda.Stream(device=dev)
        side.wait_stream(cur)
        with torch.cuda.stream(side):
            for _ in range(max(1, warmup)):
                self._eager()
        cur.wait_stream(side)
        torch.cuda.synchronize(dev)
        self.graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(self.graph):
            self.losses, self.grad_norms = self._eager()
        with torch.no_grad():  # the warm-up passes stepped the optimizers: restore parameters and optimizer state
            for t, saved in keep:
                t.copy_(saved)

    def _eager(self):
        observation = self.all_observation[:-1].float() / 255.0 - 0.5  # dreamer_algo.py:225-241
        ml, al, vl, info = self._loss_fn(self.model, observation, self.all_action[1:], self.all_reward[1:].unsqueeze(2),
                                         self.done.unsqueeze(2), **self._kw)
        # autograd.grad rather than .backward(): no AccumulateGrad nodes (they pin the stream they were created on)
        gm = torch.autograd.grad(ml, self.model_params, allow_unused=True)
        ga = torch.autograd.grad(al, self.actor_params)
        gv = torch.autograd.grad(vl, self.value_params)
        for ps, gs in ((self.model_params, gm), (self.actor_params, ga), (self.value_params, gv)):
            for p, g in zip(ps, gs):
                p.grad = g
        norms = [o.step() for o in (self.model_opt, self.actor_opt, self.value_opt)]
        return torch.stack([ml.detach(), al.detach(), vl.detach()]), torch.stack([n.detach() for n in norms])

    def step(self, samples):
        """samples: any object with all_observation / all_action / all_reward / done (rlpyt's SamplesFromReplay or
        replay.SamplesFromReplay).  Returns the static tensor [model_loss, actor_loss, value_loss]."""
        self.all_observation.copy_(samples.all_observation, non_blocking=True)
        self.all_action.copy_(samples.all_action, non_blocking=True)
        self.all_reward.copy_(samples.all_reward, non_blocking=True)
        self.done.copy_(samples.done, non_blocking=True)
        self.graph.replay()
        return self.losses
