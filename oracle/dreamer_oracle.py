"""CPU restatement of the Dreamer RSSM hot path -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.

Only ``tests/``, ``__graft_entry__.smoke()`` and the ``cpu_baseline`` / ``--impl reference``
legs of ``bench.py`` may import this module.  The product path (``dreamer_pytorch_b200``) never
does: it runs hand-written CUDA and fails loudly when the extension is missing.

What this is
------------
A plain PyTorch (fp32 or fp64, any device, autograd for the backward) restatement of the
reference algorithm for the one hot path named by BASELINE.json: observation rollout, imagination
rollout, heads, lambda-returns and the actor / value / KL loss section.  Every function cites the
reference ``file:line`` it follows (paths relative to the reference checkout).  Weights are a flat
``dict`` keyed by the reference ``AgentModel.state_dict()`` names, so a reference checkpoint can
be fed in unchanged.  Noise is *injected* (explicit ``eps`` tensors) so that the CUDA path, this
oracle and the real reference can be compared on identical draws.

Pinning
-------
The reference's own tests hold no golden vectors for this path (all 19 are shape tests), so this
oracle is pinned against outputs of the *reference itself*: ``oracle/make_golden.py`` imports the
unmodified reference from ``/root/reference``, runs it on seeded inputs with the same injected
noise and commits the results under ``tests/golden/``; ``tests/test_oracle_golden.py`` checks this
restatement against those fixtures (and, when ``/root/reference`` is present, against the live
reference on further cases).
"""
from __future__ import annotations

import math
from collections import namedtuple
from typing import Dict, Optional

import torch
import torch.nn.functional as F

State = namedtuple("State", ["mean", "std", "stoch", "deter"])

TR = "transition."
POST = "representation._stochastic_posterior_model."
ACT = "action_decoder.feedforward_model."
RAW_INIT_STD = math.log(math.exp(5.0) - 1.0)  # action.py:35
LOG_2PI = math.log(2.0 * math.pi)


# --------------------------------------------------------------------------------------------
# helpers
# --------------------------------------------------------------------------------------------
def _linear(w: Dict[str, torch.Tensor], key: str, x: torch.Tensor) -> torch.Tensor:
    return F.linear(x, w[key + ".weight"], w[key + ".bias"])


def mlp(w, prefix: str, x: torch.Tensor, layers: int) -> torch.Tensor:
    """``layers`` x (Linear + ELU) followed by one Linear.

    Follows ``DenseModel.build_model`` (dense.py:28-35) and ``ActionDecoder.build_model``
    (action.py:37-49): nn.Sequential indices 0,2,4,... are the Linear modules.
    """
    for i in range(layers):
        x = F.elu(_linear(w, f"{prefix}{2 * i}", x))
    return _linear(w, f"{prefix}{2 * layers}", x)


def get_feat(state: State) -> torch.Tensor:
    """rnns.py:22-23"""
    return torch.cat((state.stoch, state.deter), dim=-1)


def stack_states(states, dim=0) -> State:
    """rnns.py:13-19"""
    return State(*(torch.stack([getattr(s, f) for s in states], dim=dim) for f in State._fields))


def gru_cell(w, x, h):
    """torch ``nn.GRUCell`` (rnns.py:72,102): gate order r,z,n; reset gate applied to W_hn h + b_hn."""
    gi = F.linear(x, w[TR + "_cell.weight_ih"], w[TR + "_cell.bias_ih"])
    gh = F.linear(h, w[TR + "_cell.weight_hh"], w[TR + "_cell.bias_hh"])
    i_r, i_z, i_n = gi.chunk(3, -1)
    h_r, h_z, h_n = gh.chunk(3, -1)
    r = torch.sigmoid(i_r + h_r)
    z = torch.sigmoid(i_z + h_z)
    n = torch.tanh(i_n + r * h_n)
    return (1.0 - z) * n + z * h


# --------------------------------------------------------------------------------------------
# RSSM steps and rollouts
# --------------------------------------------------------------------------------------------
def transition_step(w, prev_action, prev_stoch, prev_deter, eps) -> State:
    """``RSSMTransition.forward`` (rnns.py:98-107)."""
    x = F.elu(_linear(w, TR + "_rnn_input_model.0", torch.cat([prev_action, prev_stoch], -1)))
    deter = gru_cell(w, x, prev_deter)
    hid = F.elu(_linear(w, TR + "_stochastic_prior_model.0", deter))
    mean, std = _linear(w, TR + "_stochastic_prior_model.2", hid).chunk(2, -1)
    std = F.softplus(std) + 0.1
    stoch = mean + std * eps  # Normal.rsample, torch/distributions/normal.py:82-85
    return State(mean, std, stoch, deter)


def representation_step(w, obs_embed, prev_action, prev_stoch, prev_deter, eps_prior, eps_post):
    """``RSSMRepresentation.forward`` (rnns.py:149-159)."""
    prior = transition_step(w, prev_action, prev_stoch, prev_deter, eps_prior)
    x = torch.cat([prior.deter, obs_embed], -1)
    hid = F.elu(_linear(w, POST + "0", x))
    mean, std = _linear(w, POST + "2", hid).chunk(2, -1)
    std = F.softplus(std) + 0.1
    stoch = mean + std * eps_post
    return prior, State(mean, std, stoch, prior.deter)


def rollout_representation(w, steps, obs_embed, action, prev_stoch, prev_deter, eps_prior, eps_post):
    """``RSSMRollout.rollout_representation`` (rnns.py:179-205)."""
    priors, posts = [], []
    s, h = prev_stoch, prev_deter
    for t in range(steps):
        prior, post = representation_step(w, obs_embed[t], action[t], s, h, eps_prior[t], eps_post[t])
        priors.append(prior)
        posts.append(post)
        s, h = post.stoch, post.deter
    return stack_states(priors), stack_states(posts)


def rollout_transition(w, steps, action, prev_stoch, prev_deter, eps):
    """``RSSMRollout.rollout_transition`` (rnns.py:207-222)."""
    out = []
    s, h = prev_stoch, prev_deter
    for t in range(steps):
        st = transition_step(w, action[t], s, h, eps[t])
        out.append(st)
        s, h = st.stoch, st.deter
    return stack_states(out)


def actor_raw(w, feat, layers=3):
    return mlp(w, ACT, feat, layers)


def policy_action(w, feat, dist: str, noise, layers=3, temperature=0.1):
    """``AgentModel.policy`` in training mode (agent.py:101-117) on top of
    ``ActionDecoder.forward`` (action.py:51-66).

    noise: N(0,1) draws for ``tanh_normal``; Gumbel(0,1) draws for the two categorical variants
    (the reference draws them with ``torch.multinomial`` / ``torch.rand``; the injected-noise
    formulation is the Gumbel-max one).
    """
    x = actor_raw(w, feat, layers)
    if dist == "tanh_normal":
        mean, std = x.chunk(2, -1)
        mean = 5.0 * torch.tanh(mean / 5.0)  # mean_scale=5, action.py:56
        std = F.softplus(std + RAW_INIT_STD) + 1e-4  # action.py:57
        return torch.tanh(mean + std * noise)  # TransformedDistribution.rsample + TanhTransform
    if dist == "one_hot":
        probs = torch.softmax(x, -1)
        idx = torch.argmax(x + noise, -1)
        sample = F.one_hot(idx, x.shape[-1]).to(x.dtype)
        return sample + probs - probs.detach()  # straight-through, agent.py:110-112
    if dist == "relaxed_one_hot":
        # RelaxedOneHotCategorical(0.1).rsample: softmax((logits + g) / T)
        return torch.softmax((x + noise) / temperature, -1)
    raise NotImplementedError(dist)


def rollout_policy(w, steps, prev_stoch, prev_deter, noise_act, eps_prior, dist="tanh_normal", layers=3):
    """``RSSMRollout.rollout_policy`` (rnns.py:224-244) with ``policy = AgentModel.policy``.

    The start state is detached and so is the state handed to the policy at every step
    (rnns.py:236-238): the action has no gradient path back into the state that produced it.
    """
    s, h = prev_stoch.detach(), prev_deter.detach()
    states, actions = [], []
    for t in range(steps):
        feat = torch.cat([s.detach(), h.detach()], -1)
        a = policy_action(w, feat, dist, noise_act[t], layers)
        st = transition_step(w, a, s, h, eps_prior[t])
        states.append(st)
        actions.append(a)
        s, h = st.stoch, st.deter
    return stack_states(states), torch.stack(actions, 0)


# --------------------------------------------------------------------------------------------
# heads, returns, losses
# --------------------------------------------------------------------------------------------
def dense_mean(w, prefix, feat, layers=3):
    """``DenseModel.forward`` (dense.py:37-50): the location of the Normal / the Bernoulli logits."""
    return mlp(w, prefix + ".model.", feat, layers)


def normal_unit_logprob(pred, target):
    """Independent(Normal(pred, 1), 1).log_prob(target) with a trailing event dim of size 1."""
    return (-0.5 * (target - pred) ** 2 - 0.5 * LOG_2PI).sum(-1)


def bernoulli_logprob(logits, target):
    """Independent(Bernoulli(logits), 1).log_prob(target) = -BCEWithLogits (soft targets allowed)."""
    return (target * logits - F.softplus(logits)).sum(-1)


def kl_normal(mean_q, std_q, mean_p, std_p):
    """KL(N(q) || N(p)) summed over the last dim (torch/distributions/kl.py ``_kl_normal_normal``)."""
    var_ratio = (std_q / std_p) ** 2
    t1 = ((mean_q - mean_p) / std_p) ** 2
    return (0.5 * (var_ratio + t1 - 1.0 - var_ratio.log())).sum(-1)


def normal_entropy(std):
    return (0.5 + 0.5 * LOG_2PI + std.log()).sum(-1)


def compute_return(reward, value, discount, bootstrap, lambda_):
    """``Dreamer.compute_return`` (dreamer_algo.py:398-422)."""
    next_values = torch.cat([value[1:], bootstrap[None]], 0)
    target = reward + discount * next_values * (1 - lambda_)
    outputs = []
    acc = bootstrap
    for t in range(reward.shape[0] - 1, -1, -1):
        acc = target[t] + discount[t] * lambda_ * acc
        outputs.append(acc)
    return torch.flip(torch.stack(outputs), [0])


def imagination_losses(
    w,
    start_stoch,
    start_deter,
    noise_act,
    eps_prior,
    horizon,
    dist="tanh_normal",
    discount=0.99,
    lambda_=0.95,
    use_pcont=False,
    w_actor: Optional[dict] = None,
    w_value: Optional[dict] = None,
):
    """Actor / value block of ``Dreamer.loss`` (dreamer_algo.py:283-337).

    ``w_actor`` / ``w_value`` optionally supply the leaf tensors that should receive gradients
    (mirrors ``FreezeParameters``: the world model is frozen for the actor loss, world model and
    value model for the imagined heads, everything but the value model for the value loss).
    Returns ``(actor_loss, value_loss, aux)``.
    """
    frozen = {k: v.detach() for k, v in w.items()}
    wa = dict(frozen)
    if w_actor is not None:
        wa.update(w_actor)
    imag, actions = rollout_policy(wa, horizon, start_stoch, start_deter, noise_act, eps_prior, dist)
    imag_feat = get_feat(imag)
    imag_reward = dense_mean(frozen, "reward_model", imag_feat)
    value = dense_mean(frozen, "value_model", imag_feat)
    if use_pcont:
        discount_arr = torch.sigmoid(dense_mean(frozen, "pcont", imag_feat))  # Bernoulli.mean
    else:
        discount_arr = discount * torch.ones_like(imag_reward)
    returns = compute_return(imag_reward[:-1], value[:-1], discount_arr[:-1], value[-1], lambda_)
    discount_arr = torch.cat([torch.ones_like(discount_arr[:1]), discount_arr[1:]])
    disc = torch.cumprod(discount_arr[:-1], 0)
    actor_loss = -torch.mean(disc * returns)

    wv = dict(frozen)
    if w_value is not None:
        wv.update(w_value)
    value_pred = dense_mean(wv, "value_model", imag_feat[:-1].detach())
    log_prob = normal_unit_logprob(value_pred, returns.detach())
    value_loss = -torch.mean(disc.detach() * log_prob.unsqueeze(2))
    aux = dict(imag=imag, actions=actions, reward=imag_reward, value=value, returns=returns, discount=disc)
    return actor_loss, value_loss, aux


# --------------------------------------------------------------------------------------------
# conv encoder / decoder (observation.py) -- needed only for the full ``Dreamer.loss`` restatement
# --------------------------------------------------------------------------------------------
def observation_encoder(w, obs):
    """``ObservationEncoder.forward`` (observation.py:24-29): 4 x (Conv2d k4 s2 + ReLU)."""
    batch_shape = obs.shape[:-3]
    x = obs.reshape(-1, *obs.shape[-3:])
    for i in (0, 2, 4, 6):
        p = f"observation_encoder.convolutions.{i}"
        x = F.relu(F.conv2d(x, w[p + ".weight"], w[p + ".bias"], stride=2))
    return x.reshape(*batch_shape, -1)


def observation_decoder(w, feat, shape=(3, 64, 64)):
    """``ObservationDecoder.forward`` (observation.py:103-117) for a 64x64 image (output_padding 0)."""
    batch_shape = feat.shape[:-1]
    x = _linear(w, "observation_decoder.linear", feat.reshape(-1, feat.shape[-1]))
    x = x.reshape(x.shape[0], -1, 1, 1)
    for i in (0, 2, 4, 6):
        p = f"observation_decoder.decoder.{i}"
        x = F.conv_transpose2d(x, w[p + ".weight"], w[p + ".bias"], stride=2)
        if i != 6:
            x = F.relu(x)
    return x.reshape(*batch_shape, *shape)


def dreamer_loss(
    w,
    all_observation,
    all_action,
    all_reward,
    done,
    noise,
    horizon=15,
    dist="tanh_normal",
    discount=0.99,
    lambda_=0.95,
    free_nats=3.0,
    kl_scale=1.0,
    use_pcont=False,
    pcont_scale=10.0,
    obs_embed=None,
):
    """``Dreamer.loss`` (dreamer_algo.py:213-371).

    ``noise`` = dict(eps_prior_obs [T,B,S], eps_post_obs [T,B,S], noise_act [H,N,A], eps_prior_img [H,N,S]).
    With ``obs_embed`` given ([T,B,E]) the conv encoder/decoder and the image loss are skipped
    (RSSM-only configuration).
    Returns dict of losses and intermediate tensors.
    """
    action = all_action[1:]
    reward = all_reward[1:].unsqueeze(2)
    T, B = action.shape[:2]
    if obs_embed is None:
        observation = all_observation[:-1].to(action.dtype) / 255.0 - 0.5
        embed = observation_encoder(w, observation)
    else:
        observation, embed = None, obs_embed
    S = w[TR + "_stochastic_prior_model.2.bias"].shape[0] // 2
    D = w[TR + "_cell.weight_hh"].shape[1]
    zeros_s = action.new_zeros(B, S)
    zeros_h = action.new_zeros(B, D)
    prior, post = rollout_representation(
        w, T, embed, action, zeros_s, zeros_h, noise["eps_prior_obs"], noise["eps_post_obs"]
    )
    feat = get_feat(post)
    reward_pred = dense_mean(w, "reward_model", feat)
    reward_loss = -torch.mean(normal_unit_logprob(reward_pred, reward))
    if observation is not None:
        image_mean = observation_decoder(w, feat)
        image_logprob = (-0.5 * (observation - image_mean) ** 2 - 0.5 * LOG_2PI).sum((-3, -2, -1))
        image_loss = -torch.mean(image_logprob)
    else:
        image_loss = reward_loss.new_zeros(())
    pcont_loss = reward_loss.new_zeros(())
    if use_pcont:
        pcont_logits = dense_mean(w, "pcont", feat)
        pcont_target = discount * (1 - done.unsqueeze(2).to(action.dtype))
        pcont_loss = -torch.mean(bernoulli_logprob(pcont_logits, pcont_target))
    div = torch.mean(kl_normal(post.mean, post.std, prior.mean, prior.std))
    div = torch.max(div, div.new_full(div.size(), free_nats))
    model_loss = kl_scale * div + reward_loss + image_loss
    if use_pcont:
        model_loss = model_loss + pcont_scale * pcont_loss

    with torch.no_grad():
        if use_pcont:
            flat_stoch = post.stoch[:-1].reshape((T - 1) * B, -1)
            flat_deter = post.deter[:-1].reshape((T - 1) * B, -1)
        else:
            flat_stoch = post.stoch.reshape(T * B, -1)
            flat_deter = post.deter.reshape(T * B, -1)
    w_actor = {k: v for k, v in w.items() if k.startswith("action_decoder.")}
    w_value = {k: v for k, v in w.items() if k.startswith("value_model.")}
    actor_loss, value_loss, aux = imagination_losses(
        w,
        flat_stoch,
        flat_deter,
        noise["noise_act"],
        noise["eps_prior_img"],
        horizon,
        dist,
        discount,
        lambda_,
        use_pcont,
        w_actor,
        w_value,
    )
    with torch.no_grad():
        prior_ent = torch.mean(normal_entropy(prior.std))
        post_ent = torch.mean(normal_entropy(post.std))
    out = dict(
        model_loss=model_loss,
        actor_loss=actor_loss,
        value_loss=value_loss,
        prior_entropy=prior_ent,
        post_entropy=post_ent,
        divergence=div,
        reward_loss=reward_loss,
        image_loss=image_loss,
        pcont_loss=pcont_loss,
        prior=prior,
        post=post,
        embed=embed,
    )
    out.update(aux)
    return out


# --------------------------------------------------------------------------------------------
# deterministic synthetic weights / inputs (shared by the fixture generator, tests and bench)
# --------------------------------------------------------------------------------------------
def weight_shapes(A, S=30, D=200, Hd=200, E=1024, dist="tanh_normal", act_hidden=200, act_layers=3,
                  reward_hidden=300, reward_layers=3, value_hidden=200, value_layers=3,
                  use_pcont=False, pcont_hidden=200, pcont_layers=3, conv=False):
    """Shapes of the reference ``AgentModel.state_dict()`` entries the path touches
    (agent.py:48-87; rnns.py:72-88,134-139; action.py:37-49; dense.py:28-35).  Only the canonical
    (un-aliased) key of each shared tensor is listed."""
    F_ = S + D
    sh = {
        TR + "_cell.weight_ih": (3 * D, Hd),
        TR + "_cell.weight_hh": (3 * D, D),
        TR + "_cell.bias_ih": (3 * D,),
        TR + "_cell.bias_hh": (3 * D,),
        TR + "_rnn_input_model.0.weight": (Hd, A + S),
        TR + "_rnn_input_model.0.bias": (Hd,),
        TR + "_stochastic_prior_model.0.weight": (Hd, D),
        TR + "_stochastic_prior_model.0.bias": (Hd,),
        TR + "_stochastic_prior_model.2.weight": (2 * S, Hd),
        TR + "_stochastic_prior_model.2.bias": (2 * S,),
        POST + "0.weight": (Hd, D + E),
        POST + "0.bias": (Hd,),
        POST + "2.weight": (2 * S, Hd),
        POST + "2.bias": (2 * S,),
    }

    def add_mlp(prefix, inp, hidden, layers, out):
        d = inp
        for i in range(layers):
            sh[f"{prefix}{2 * i}.weight"] = (hidden, d)
            sh[f"{prefix}{2 * i}.bias"] = (hidden,)
            d = hidden
        sh[f"{prefix}{2 * layers}.weight"] = (out, d)
        sh[f"{prefix}{2 * layers}.bias"] = (out,)

    add_mlp(ACT, F_, act_hidden, act_layers, 2 * A if dist == "tanh_normal" else A)
    add_mlp("reward_model.model.", F_, reward_hidden, reward_layers, 1)
    add_mlp("value_model.model.", F_, value_hidden, value_layers, 1)
    if use_pcont:
        add_mlp("pcont.model.", F_, pcont_hidden, pcont_layers, 1)
    if conv:
        for i, (o, c) in zip((0, 2, 4, 6), ((32, 3), (64, 32), (128, 64), (256, 128))):
            sh[f"observation_encoder.convolutions.{i}.weight"] = (o, c, 4, 4)
            sh[f"observation_encoder.convolutions.{i}.bias"] = (o,)
        sh["observation_decoder.linear.weight"] = (1024, F_)
        sh["observation_decoder.linear.bias"] = (1024,)
        for i, (c, o, k) in zip((0, 2, 4, 6), ((1024, 128, 5), (128, 64, 5), (64, 32, 6), (32, 3, 6))):
            sh[f"observation_decoder.decoder.{i}.weight"] = (c, o, k, k)
            sh[f"observation_decoder.decoder.{i}.bias"] = (o,)
    return sh


def make_weights(shapes: Dict[str, tuple], seed: int, dtype=torch.float32, gain: float = 1.0):
    """U(-b, b) with b = gain / sqrt(fan_in), drawn key by key (sorted) from one seeded CPU generator:
    the same scale as torch's default Linear / GRUCell / Conv init, but reproducible from the
    seed alone, independent of module construction order."""
    g = torch.Generator().manual_seed(seed)
    out = {}
    for k in sorted(shapes):
        shp = shapes[k]
        wkey = k[: -len(".bias")] + ".weight" if k.endswith(".bias") else k
        if k.endswith("bias_ih") or k.endswith("bias_hh"):
            wkey = k.replace("bias", "weight")
        ws = shapes.get(wkey, shp)
        if "observation_decoder.decoder" in wkey:  # ConvTranspose2d: fan_in = out_ch * k * k
            fan_in = ws[1] * ws[2] * ws[3]
        elif len(ws) > 1:
            fan_in = 1
            for d in ws[1:]:
                fan_in *= d
        else:
            fan_in = ws[0]
        b = gain / math.sqrt(fan_in)
        out[k] = ((torch.rand(shp, generator=g, dtype=torch.float64) * 2 - 1) * b).to(dtype)
    return out


def expand_aliases(w: Dict[str, torch.Tensor]) -> Dict[str, torch.Tensor]:
    """Full 90-key (non-pcont) reference state_dict: the transition's 10 tensors also appear under
    three more prefixes and the posterior's 4 under one more (checkpoint row)."""
    full = dict(w)
    for k, v in w.items():
        if k.startswith(TR):
            tail = k[len(TR):]
            full["representation._transition_model." + tail] = v
            full["rollout.representation_model._transition_model." + tail] = v
            full["rollout.transition_model." + tail] = v
        if k.startswith(POST):
            tail = k[len(POST):]
            full["rollout.representation_model._stochastic_posterior_model." + tail] = v
    return full


def make_noise(T, B, N, H, S, A, seed, dist="tanh_normal", dtype=torch.float32):
    g = torch.Generator().manual_seed(seed)
    n = dict(
        eps_prior_obs=torch.randn(T, B, S, generator=g, dtype=torch.float64),
        eps_post_obs=torch.randn(T, B, S, generator=g, dtype=torch.float64),
        eps_prior_img=torch.randn(H, N, S, generator=g, dtype=torch.float64),
    )
    if dist == "tanh_normal":
        n["noise_act"] = torch.randn(H, N, A, generator=g, dtype=torch.float64)
    else:
        u = torch.rand(H, N, A, generator=g, dtype=torch.float64).clamp(1e-9, 1 - 1e-9)
        n["noise_act"] = -torch.log(-torch.log(u))
    return {k: v.to(dtype) for k, v in n.items()}


def normwise_err(x: torch.Tensor, ref: torch.Tensor) -> float:
    """SURVEY.md section 8(c) tolerance definition: max|X - X_ref| / max(max|X_ref|, tiny)."""
    x = x.detach().double().cpu()
    ref = ref.detach().double().cpu()
    return float((x - ref).abs().max() / max(float(ref.abs().max()), 1e-30))


# ------------------------------------------------------------------------------------------------
# optimizer tail: dreamer_algo.py:182-194 = clip_grad_norm_ + Adam.step, on CPU tensors.
# The arithmetic lives in torch (clip_grad.py, adam.py); the oracle simply runs it.
# ------------------------------------------------------------------------------------------------
def clip_adam_steps(params, grads_per_step, lr, grad_clip, betas=(0.9, 0.999), eps=1e-8):
    """params: list of CPU fp32 tensors (cloned); grads_per_step: list (one entry per step) of lists of gradients.
    Returns (final params, list of the total norms clip_grad_norm_ reported, final exp_avg, final exp_avg_sq)."""
    ps = [torch.nn.Parameter(p.detach().clone().float()) for p in params]
    opt = torch.optim.Adam(ps, lr=lr, betas=betas, eps=eps)
    norms = []
    for grads in grads_per_step:
        for p, g in zip(ps, grads):
            p.grad = g.detach().clone().float()
        norms.append(torch.nn.utils.clip_grad_norm_(ps, grad_clip).detach().clone())
        opt.step()
    return ([p.detach() for p in ps], norms, [opt.state[p]["exp_avg"] for p in ps], [opt.state[p]["exp_avg_sq"] for p in ps])
