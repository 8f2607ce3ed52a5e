"""Drives the UNMODIFIED reference -- TEST / BASELINE INFRASTRUCTURE ONLY.

The reference is imported from ``/root/reference`` (build container) or, where that does not exist (the
GPU box), from the git-ignored copy ``oracle/_ref/`` that ``oracle/fetch_ref.py`` makes and that travels
with the tree.  Used to (a) validate ``oracle/dreamer_oracle.py``, (b) generate the golden fixtures under
``tests/golden/`` (``oracle/make_golden.py``), (c) time the reference itself on the host cores and on
CUDA eager (``bench.py``: ``--impl reference``, ``cpu_baseline``, ``reference_cuda``) and (d) run the
reference's own ``Dreamer.loss`` on the drop-in modules (``tests/test_reference_algo_gpu.py``).
The product never imports this module.

Noise injection follows SURVEY.md section 8(c): ``torch.distributions.normal._standard_normal`` is
patched to pop pre-generated tensors in the reference's own draw order; the categorical samplers
are patched to the Gumbel-max formulation on injected Gumbel noise.
"""
from __future__ import annotations

import contextlib
import os
import sys
import types

import torch

_LOCAL_COPY = os.path.join(os.path.dirname(os.path.abspath(__file__)), "_ref")


def _find_root() -> str:
    env = os.environ.get("DREAMER_REFERENCE_ROOT")
    if env:
        return env
    if os.path.isdir("/root/reference/dreamer"):
        return "/root/reference"
    return _LOCAL_COPY


REF_ROOT = _find_root()


def reference_available() -> bool:
    return os.path.isdir(os.path.join(REF_ROOT, "dreamer")) and os.path.isdir(os.path.join(REF_ROOT, "rlpyt"))


def import_reference():
    """Put the reference on sys.path (read-only mount: no bytecode) and import its modules."""
    if not reference_available():
        raise RuntimeError(f"reference checkout not found at {REF_ROOT}")
    sys.dont_write_bytecode = True
    for p in (REF_ROOT, os.path.join(REF_ROOT, "rlpyt")):
        if p not in sys.path:
            sys.path.insert(0, p)
    for name in list(sys.modules):  # drop a drop-in alias of `dreamer.*` if one is installed
        if name == "dreamer" or name.startswith("dreamer."):
            mod = sys.modules[name]
            if not str(getattr(mod, "__file__", "")).startswith(REF_ROOT):
                del sys.modules[name]
    import dreamer.models.agent as agent  # noqa
    import dreamer.algos.dreamer_algo as algo  # noqa

    torch.autograd.set_detect_anomaly(False)  # dreamer_algo.py:16 turns it on at import
    torch.distributions.Distribution.set_default_validate_args(False)  # pcont path
    return agent, algo


@contextlib.contextmanager
def injected_noise(normal_queue, gumbel_queue=None):
    """normal_queue: list of tensors popped by every ``Normal.rsample``;
    gumbel_queue: list of Gumbel(0,1) tensors popped by the categorical samplers."""
    import torch.distributions.normal as tdn
    import torch.distributions.one_hot_categorical as tdo
    import torch.distributions.relaxed_categorical as tdr

    nq = list(normal_queue)
    gq = list(gumbel_queue or [])
    orig_sn = tdn._standard_normal
    orig_oh = tdo.OneHotCategorical.sample
    orig_rx = tdr.ExpRelaxedCategorical.rsample

    def fake_standard_normal(shape, dtype, device):
        t = nq.pop(0)
        assert tuple(t.shape) == tuple(shape), (tuple(t.shape), tuple(shape))
        return t.to(dtype=dtype, device=device)

    def fake_onehot_sample(self, sample_shape=torch.Size()):
        g = gq.pop(0)
        logits = self._categorical.logits
        idx = torch.argmax(logits + g.to(logits), -1)
        return torch.nn.functional.one_hot(idx, logits.shape[-1]).to(self._categorical.probs)

    def fake_relaxed_rsample(self, sample_shape=torch.Size()):
        g = gq.pop(0)
        scores = (self.logits + g.to(self.logits)) / self.temperature
        return scores - scores.logsumexp(dim=-1, keepdim=True)

    tdn._standard_normal = fake_standard_normal
    tdo.OneHotCategorical.sample = fake_onehot_sample
    tdr.ExpRelaxedCategorical.rsample = fake_relaxed_rsample
    try:
        yield
    finally:
        tdn._standard_normal = orig_sn
        tdo.OneHotCategorical.sample = orig_oh
        tdr.ExpRelaxedCategorical.rsample = orig_rx
    assert not nq and not gq, f"unused injected noise: {len(nq)} normal, {len(gq)} gumbel"


def build_reference_model(weights, A, dist, use_pcont=False, dtype=torch.float32, **model_kwargs):
    """Reference ``AtariDreamerModel`` loaded with ``weights`` (oracle.expand_aliases form not
    required: missing conv tensors keep their default init)."""
    agent, _ = import_reference()
    from oracle.dreamer_oracle import expand_aliases

    model = agent.AtariDreamerModel(action_shape=(A,), action_dist=dist, use_pcont=use_pcont, **model_kwargs)
    sd = model.state_dict()
    full = expand_aliases(weights)
    for k in sd:
        if k in full:
            sd[k] = full[k].clone()
    model.load_state_dict(sd)
    model = model.to(dtype)
    model.train()
    return model


def queues_for_loss(noise, T, H, dist):
    """The reference's draw order inside ``Dreamer.loss``."""
    nq, gq = [], []
    for t in range(T):
        nq.append(noise["eps_prior_obs"][t])
        nq.append(noise["eps_post_obs"][t])
    for h in range(H):
        if dist == "tanh_normal":
            nq.append(noise["noise_act"][h])
        else:
            gq.append(noise["noise_act"][h])
        nq.append(noise["eps_prior_img"][h])
    return nq, gq


def reference_loss(model, all_observation, all_action, all_reward, done, noise, horizon, dist,
                   use_pcont=False, kl_scale=1.0, **algo_kwargs):
    """Run the reference's ``Dreamer.loss`` on a synthetic ``SamplesFromReplay``."""
    _, algo_mod = import_reference()
    from rlpyt.replays.sequence.n_step import SamplesFromReplay

    T = all_action.shape[0] - 1
    B = all_action.shape[1]
    algo = algo_mod.Dreamer(batch_size=B, batch_length=T, horizon=horizon, use_pcont=use_pcont,
                            kl_scale=kl_scale, log_video=False, type=all_action.dtype, **algo_kwargs)
    algo.agent = types.SimpleNamespace(model=model, device=all_action.device)
    import warnings
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")  # duplicate transition params in the model optimizer (Q3)
        algo.optim_initialize()
    samples = SamplesFromReplay(all_observation=all_observation, all_action=all_action,
                                all_reward=all_reward, return_=None, done=done, done_n=None,
                                init_rnn_state=None)
    nq, gq = queues_for_loss(noise, T, horizon, dist)
    with injected_noise(nq, gq):
        model_loss, actor_loss, value_loss, info = algo.loss(samples, 0, 0)
    return algo, model_loss, actor_loss, value_loss, info


def reference_rollouts(model, obs_embed, action, noise, horizon, dist, use_pcont=False):
    """Reference ``rollout_representation`` + ``rollout_policy`` on a synthetic ``obs_embed``
    (RSSM-only configuration: no conv)."""
    import_reference()
    from rlpyt.utils.buffer import buffer_method

    T, B = action.shape[:2]
    prev = model.representation.initial_state(B, device=action.device, dtype=action.dtype)
    nq = []
    for t in range(T):
        nq += [noise["eps_prior_obs"][t], noise["eps_post_obs"][t]]
    with injected_noise(nq):
        prior, post = model.rollout.rollout_representation(T, obs_embed, action, prev)
    with torch.no_grad():
        if use_pcont:
            flat = buffer_method(post[:-1, :], "reshape", (T - 1) * B, -1)
        else:
            flat = buffer_method(post, "reshape", T * B, -1)
    nq, gq = [], []
    for h in range(horizon):
        (nq if dist == "tanh_normal" else gq).append(noise["noise_act"][h])
        nq.append(noise["eps_prior_img"][h])
    with injected_noise(nq, gq):
        imag, actions = model.rollout.rollout_policy(horizon, model.policy, flat)
    return prior, post, imag, actions


# ------------------------------------------------------------------------------------------------
# timing harnesses (bench.py): the reference's own modules, on the host cores or on CUDA eager
# ------------------------------------------------------------------------------------------------
def set_as_shipped(on: bool):
    """as shipped: autograd anomaly mode on (dreamer_algo.py:16) and distribution argument validation on (torch
    default); best case: both off."""
    torch.autograd.set_detect_anomaly(bool(on))
    torch.distributions.Distribution.set_default_validate_args(bool(on))


def default_reference_model(A, dist, use_pcont=False, device="cpu", seed=0, **model_kwargs):
    """Reference ``AtariDreamerModel`` with its own default init (SURVEY 8(d) config recipes)."""
    agent, _ = import_reference()
    torch.manual_seed(seed)
    model = agent.AtariDreamerModel(action_shape=(A,), action_dist=dist, use_pcont=use_pcont, **model_kwargs)
    return model.to(device).train()


def reference_algo(model, B, T, horizon, use_pcont=False, device="cpu", **algo_kwargs):
    """A reference ``Dreamer`` wired to ``model`` without sampler / runner."""
    _, algo_mod = import_reference()
    algo = algo_mod.Dreamer(batch_size=B, batch_length=T, horizon=horizon, use_pcont=use_pcont, log_video=False,
                            **algo_kwargs)
    algo.agent = types.SimpleNamespace(model=model, device=torch.device(device))
    import warnings
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        algo.optim_initialize()
    torch.autograd.set_detect_anomaly(False)
    return algo


def imagination_step_fn(model, algo, flat_post):
    """The actor / value block of ``Dreamer.loss`` (dreamer_algo.py:283-337) on the reference's OWN modules, then
    ``actor_loss.backward()`` and ``value_loss.backward()`` (dreamer_algo.py:176-180) -- the metric's "step".
    flat_post: reference RSSMState [N, .] (detached start states)."""
    import_reference()
    from dreamer.models.rnns import get_feat
    from dreamer.utils.module import FreezeParameters

    params = list(model.action_decoder.parameters()) + list(model.value_model.parameters())

    def step():
        for p in params:
            p.grad = None
        with FreezeParameters(algo.model_modules):  # :293-296
            imag_dist, _ = model.rollout.rollout_policy(algo.horizon, model.policy, flat_post)
        imag_feat = get_feat(imag_dist)  # :299
        with FreezeParameters(algo.model_modules + algo.value_modules):  # :305-307
            imag_reward = model.reward_model(imag_feat).mean
            value = model.value_model(imag_feat).mean
        if algo.use_pcont:  # :309-313
            with FreezeParameters([model.pcont]):
                discount_arr = model.pcont(imag_feat).mean
        else:
            discount_arr = algo.discount * torch.ones_like(imag_reward)
        returns = algo.compute_return(imag_reward[:-1], value[:-1], discount_arr[:-1], bootstrap=value[-1],
                                      lambda_=algo.discount_lambda)  # :314-320
        discount_arr = torch.cat([torch.ones_like(discount_arr[:1]), discount_arr[1:]])  # :322
        discount = torch.cumprod(discount_arr[:-1], 0)
        actor_loss = -torch.mean(discount * returns)  # :324
        with torch.no_grad():  # :331-334
            value_feat = imag_feat[:-1].detach()
            value_discount = discount.detach()
            value_target = returns.detach()
        value_pred = model.value_model(value_feat)
        log_prob = value_pred.log_prob(value_target)
        value_loss = -torch.mean(value_discount * log_prob.unsqueeze(2))  # :337
        actor_loss.backward()
        value_loss.backward()
        return actor_loss.detach(), value_loss.detach()

    return step


def synthetic_samples(A, dist, B, T, device="cpu", seed=1):
    """SamplesFromReplay of SURVEY 8(d) config 2 / 3 shapes."""
    import_reference()
    from rlpyt.replays.sequence.n_step import SamplesFromReplay

    g = torch.Generator().manual_seed(seed)
    obs = torch.randint(0, 256, (T + 1, B, 3, 64, 64), generator=g, dtype=torch.uint8)
    if dist == "tanh_normal":
        act = torch.rand(T + 1, B, A, generator=g) * 2 - 1
    else:
        act = torch.nn.functional.one_hot(torch.randint(0, A, (T + 1, B), generator=g), A).float()
    rew = torch.randn(T + 1, B, generator=g)
    done = torch.rand(T, B, generator=g) < 0.01
    return SamplesFromReplay(all_observation=obs.to(device), all_action=act.to(device), all_reward=rew.to(device),
                             return_=None, done=done.to(device), done_n=None, init_rnn_state=None)


def full_loss_step_fn(model, algo, samples, optimize=False):
    """The reference's whole ``Dreamer.loss`` + the three backwards of ``optimize_agent`` (dreamer_algo.py:174-180);
    with ``optimize`` also its clip_grad_norm_ x 3 + Adam.step x 3 + the 13 ``.item()`` read-backs (dreamer_algo.py:182-209),
    i.e. the body of one iteration of the reference's training loop."""
    import_reference()
    from dreamer.utils.module import get_parameters

    def step():
        if optimize:
            algo.model_optimizer.zero_grad(); algo.actor_optimizer.zero_grad(); algo.value_optimizer.zero_grad()
        else:
            model.zero_grad(set_to_none=True)
        model_loss, actor_loss, value_loss, info = algo.loss(samples, 0, 0)
        model_loss.backward()
        actor_loss.backward()
        value_loss.backward()
        if optimize:
            import warnings
            with warnings.catch_warnings():
                warnings.simplefilter("ignore")
                norms = [torch.nn.utils.clip_grad_norm_(get_parameters(ms), algo.grad_clip)
                         for ms in (algo.model_modules, algo.actor_modules, algo.value_modules)]
            algo.model_optimizer.step(); algo.actor_optimizer.step(); algo.value_optimizer.step()
            out = [(model_loss + actor_loss + value_loss).item()] + [n.item() for n in norms]
            out += [getattr(info, f).item() for f in info._fields]
            return out
        return model_loss.detach(), actor_loss.detach(), value_loss.detach()

    return step
