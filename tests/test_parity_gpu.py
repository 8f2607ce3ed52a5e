"""GPU parity tests: every CUDA entry point (through the C ABI) against the CPU oracle on the same
seeded inputs, and the full loss section against golden outputs of the unmodified reference.

Tolerance (BASELINE.json north_star / SURVEY.md 8(c)): normwise max|X - X_ref| / max|X_ref| <= 1e-4
per output tensor, loss scalar and parameter gradient on the fp32 path.
"""
import math

import pytest
import torch

from _cases import CASES, FP32_TOL, case_inputs, load_golden, model_kwargs, oracle_loss, rel_err
from oracle import dreamer_oracle as O

pytestmark = pytest.mark.gpu


def _w(shapes, seed, dev):
    return {k: v.to(dev) for k, v in O.make_weights(shapes, seed).items()}


# ------------------------------------------------------------------------------------------------
# batched MLP
# ------------------------------------------------------------------------------------------------
@pytest.mark.parametrize("M,k0,k1,hidden,out,layers", [
    (777, 30, 200, 200, 1, 3),     # value head shape, ragged M
    (50, 30, 200, 300, 1, 3),      # reward head shape, M smaller than a wave
    (1, 8, 0, 32, 5, 1),           # single row, single segment, wide-ish output
    (2500, 30, 200, 200, 12, 3),   # actor-shaped (out = 2A)
    (333, 7, 9, 36, 40, 2),        # odd segment sizes (padding paths), out >= 32
])
def test_mlp_fwd_bwd(cuda, M, k0, k1, hidden, out, layers):
    from dreamer_pytorch_b200.functional import MlpFn

    g = torch.Generator().manual_seed(5)
    F_ = k0 + k1
    ws = []
    d = F_
    for l in range(layers):
        ws += [torch.randn(hidden, d, generator=g) / math.sqrt(d), torch.randn(hidden, generator=g) * 0.1]
        d = hidden
    ws += [torch.randn(out, d, generator=g) / math.sqrt(d), torch.randn(out, generator=g) * 0.1]
    x0 = torch.randn(M, k0, generator=g)
    x1 = torch.randn(M, k1, generator=g) if k1 else None
    gy = torch.randn(M, out, generator=g)

    def ref(x0, x1, ws):
        x = torch.cat([x0, x1], -1) if x1 is not None else x0
        for l in range(layers):
            x = torch.nn.functional.elu(torch.nn.functional.linear(x, ws[2 * l], ws[2 * l + 1]))
        return torch.nn.functional.linear(x, ws[-2], ws[-1])

    r_in = [t.double().requires_grad_(True) for t in ([x0] + ([x1] if k1 else []) + ws)]
    ry = ref(r_in[0], r_in[1] if k1 else None, r_in[(2 if k1 else 1):])
    rg = torch.autograd.grad(ry, r_in, gy.double())

    c_in = [t.to(cuda).requires_grad_(True) for t in ([x0] + ([x1] if k1 else []) + ws)]
    cy = MlpFn.apply(c_in[0], c_in[1] if k1 else None, *c_in[(2 if k1 else 1):])
    cg = torch.autograd.grad(cy, c_in, gy.to(cuda))
    assert rel_err(cy, ry) < FP32_TOL
    for a, b in zip(cg, rg):
        assert rel_err(a, b) < FP32_TOL


# ------------------------------------------------------------------------------------------------
# rollouts
# ------------------------------------------------------------------------------------------------
def _rssm_list(w, repr_):
    TR, PO = O.TR, O.POST
    keys = [TR + "_rnn_input_model.0.weight", TR + "_rnn_input_model.0.bias", TR + "_cell.weight_ih",
            TR + "_cell.weight_hh", TR + "_cell.bias_ih", TR + "_cell.bias_hh",
            TR + "_stochastic_prior_model.0.weight", TR + "_stochastic_prior_model.0.bias",
            TR + "_stochastic_prior_model.2.weight", TR + "_stochastic_prior_model.2.bias"]
    if repr_:
        keys += [PO + "0.weight", PO + "0.bias", PO + "2.weight", PO + "2.bias"]
    return keys


def _actor_list(layers=3):
    return [f"{O.ACT}{2 * i}.{p}" for i in range(layers + 1) for p in ("weight", "bias")]


SIZES = [
    dict(A=6, S=30, D=200, Hd=200, E=1024, n=53, T=7),   # reference widths, ragged n
    dict(A=1, S=30, D=200, Hd=200, E=1024, n=150, T=3),  # cartpole action size (A+S < 32)
    dict(A=10, S=12, D=40, Hd=36, E=100, n=4, T=10),     # the reference's own test sizes (tests/test_rnns.py)
]


@pytest.mark.parametrize("sz", SIZES)
@pytest.mark.parametrize("mode", ["transition", "representation"])
def test_rollout_data(cuda, sz, mode):
    from dreamer_pytorch_b200 import _lib
    from dreamer_pytorch_b200.functional import RolloutCfg, RolloutFn

    A, S, D, Hd, E, n, T = (sz[k] for k in ("A", "S", "D", "Hd", "E", "n", "T"))
    repr_ = mode == "representation"
    shapes = O.weight_shapes(A, S, D, Hd, E)
    w = O.make_weights(shapes, 21)
    g = torch.Generator().manual_seed(22)
    start_s, start_h = torch.randn(n, S, generator=g), torch.tanh(torch.randn(n, D, generator=g))
    action = torch.rand(T, n, A, generator=g) * 2 - 1
    embed = torch.randn(T, n, E, generator=g)
    eps1, eps2 = torch.randn(T, n, S, generator=g), torch.randn(T, n, S, generator=g)
    keys = _rssm_list(w, repr_)

    def run_oracle():
        wd = {k: v.double().requires_grad_(True) for k, v in w.items()}
        ins = [t.double().requires_grad_(True) for t in (start_s, start_h, action, embed)]
        if repr_:
            prior, post = O.rollout_representation(wd, T, ins[3], ins[2], ins[0], ins[1], eps1.double(), eps2.double())
            outs = list(prior) + [post.mean, post.std, post.stoch]
        else:
            prior = O.rollout_transition(wd, T, ins[2], ins[0], ins[1], eps1.double())
            outs = list(prior)
        return outs, ins, [wd[k] for k in keys]

    def run_cuda():
        wc = [w[k].to(cuda).requires_grad_(True) for k in keys] + ([None] * 4 if not repr_ else [])
        ins = [t.to(cuda).requires_grad_(True) for t in (start_s, start_h, action, embed)]
        cfg = RolloutCfg(_lib.MODE_REPRESENTATION if repr_ else _lib.MODE_TRANSITION, T, n, (A, S, D, Hd, E if repr_ else 0))
        outs = RolloutFn.apply(cfg, ins[0], ins[1], ins[2], ins[3] if repr_ else None, eps1.to(cuda),
                               eps2.to(cuda) if repr_ else None, None, *wc)
        return list(outs), ins, [x for x in wc if x is not None]

    ro, ri, rw = run_oracle()
    co, ci, cw = run_cuda()
    for a, b in zip(co, ro):
        assert rel_err(a, b) < FP32_TOL
    # backward: random cotangents on every output
    cots = [torch.randn(o.shape, generator=g) for o in ro]
    r_leaves = ri[:3] + ([ri[3]] if repr_ else []) + rw
    c_leaves = ci[:3] + ([ci[3]] if repr_ else []) + cw
    rg = torch.autograd.grad(ro, r_leaves, [c.double() for c in cots])
    cg = torch.autograd.grad(co, c_leaves, [c.to(cuda) for c in cots])
    for i, (a, b) in enumerate(zip(cg, rg)):
        assert rel_err(a, b) < FP32_TOL, i


@pytest.mark.parametrize("dist", ["tanh_normal", "one_hot", "relaxed_one_hot"])
@pytest.mark.parametrize("sz", SIZES[:2])
def test_rollout_policy(cuda, sz, dist):
    from dreamer_pytorch_b200 import _lib
    from dreamer_pytorch_b200.functional import RolloutCfg, RolloutFn

    A, S, D, Hd, n, H = sz["A"], sz["S"], sz["D"], sz["Hd"], sz["n"], sz["T"]
    if dist != "tanh_normal" and A == 1:
        pytest.skip("categorical policy needs A > 1")
    shapes = O.weight_shapes(A, S, D, Hd, 1024, dist)
    w = O.make_weights(shapes, 31)
    nz = O.make_noise(1, 1, n, H, S, A, 32, dist)
    g = torch.Generator().manual_seed(33)
    start_s, start_h = torch.randn(n, S, generator=g), torch.tanh(torch.randn(n, D, generator=g))
    rk, ak = _rssm_list(w, False), _actor_list()

    wd = {k: v.double().requires_grad_(True) for k, v in w.items()}
    imag, actions = O.rollout_policy(wd, H, start_s.double(), start_h.double(), nz["noise_act"].double(),
                                     nz["eps_prior_img"].double(), dist)
    ro = list(imag) + [actions]

    wr = [w[k].to(cuda).requires_grad_(True) for k in rk]
    wa = [w[k].to(cuda).requires_grad_(True) for k in ak]
    cfg = RolloutCfg(_lib.MODE_POLICY, H, n, (A, S, D, Hd, 0), act_dist=dist)
    co = list(RolloutFn.apply(cfg, start_s.to(cuda), start_h.to(cuda), None, None, nz["eps_prior_img"].to(cuda), None,
                              nz["noise_act"].to(cuda), *wr, None, None, None, None, *wa))
    for i, (a, b) in enumerate(zip(co, ro)):
        assert rel_err(a, b) < FP32_TOL, i
    cots = [torch.randn(o.shape, generator=g) for o in ro]
    rg = torch.autograd.grad(ro, [wd[k] for k in rk + ak], [c.double() for c in cots])
    cg = torch.autograd.grad(co, wr + wa, [c.to(cuda) for c in cots])
    for k, a, b in zip(rk + ak, cg, rg):
        assert rel_err(a, b) < FP32_TOL, k


def test_policy_rollout_frozen_world_model(cuda):
    """FreezeParameters semantics: frozen transition weights get no gradient, actor weights do."""
    from dreamer_pytorch_b200 import _lib
    from dreamer_pytorch_b200.functional import RolloutCfg, RolloutFn

    A, S, D, Hd, n, H = 6, 30, 200, 200, 40, 4
    w = O.make_weights(O.weight_shapes(A, S, D, Hd), 41)
    nz = O.make_noise(1, 1, n, H, S, A, 42)
    wr = [w[k].to(cuda) for k in _rssm_list(w, False)]
    wa = [w[k].to(cuda).requires_grad_(True) for k in _actor_list()]
    cfg = RolloutCfg(_lib.MODE_POLICY, H, n, (A, S, D, Hd, 0))
    outs = RolloutFn.apply(cfg, torch.zeros(n, S, device=cuda), torch.zeros(n, D, device=cuda), None, None,
                           nz["eps_prior_img"].to(cuda), None, nz["noise_act"].to(cuda), *wr, None, None, None, None, *wa)
    torch.cat([outs[2], outs[3]], -1).sum().backward()
    assert all(p.grad is not None and torch.isfinite(p.grad).all() for p in wa)
    assert all(p.grad is None for p in wr)


# ------------------------------------------------------------------------------------------------
# loss section
# ------------------------------------------------------------------------------------------------
@pytest.mark.parametrize("use_pcont", [False, True])
def test_returns_loss(cuda, use_pcont):
    from dreamer_pytorch_b200.functional import returns_loss_raw

    H, N = 15, 2500
    g = torch.Generator().manual_seed(51)
    reward, value, pc = (torch.randn(H, N, generator=g) for _ in range(3))
    r, v, p = (t.double().requires_grad_(True) for t in (reward, value, pc))
    disc_arr = torch.sigmoid(p) if use_pcont else 0.99 * torch.ones_like(r)
    returns = O.compute_return(r[:-1], v[:-1], disc_arr[:-1], v[-1], 0.95)
    disc = torch.cumprod(torch.cat([torch.ones_like(disc_arr[:1]), disc_arr[1:]])[:-1], 0)
    actor = -torch.mean(disc * returns)
    vp = value.double().requires_grad_(True)
    vloss = -torch.mean(disc.detach() * (-0.5 * (returns.detach() - vp[:-1]) ** 2 - 0.5 * O.LOG_2PI))
    gr, gv, gp = torch.autograd.grad(actor, [r, v, p], allow_unused=True)
    gvp, = torch.autograd.grad(vloss, [vp])
    out = returns_loss_raw(reward.to(cuda), value.to(cuda), pc.to(cuda) if use_pcont else None, 0.99, 0.95)
    assert rel_err(out["losses"][0], actor) < FP32_TOL
    assert rel_err(out["losses"][1], vloss) < FP32_TOL
    assert rel_err(out["returns"], returns) < FP32_TOL
    assert rel_err(out["disc"], disc) < FP32_TOL
    assert rel_err(out["d_reward"], gr) < FP32_TOL
    assert rel_err(out["d_value"], gv) < FP32_TOL
    assert rel_err(out["d_value_pred"], gvp) < FP32_TOL
    if use_pcont:
        assert rel_err(out["d_pcont"], gp) < FP32_TOL


@pytest.mark.parametrize("scale", [1.0, 0.05])
def test_kl_entropy(cuda, scale):
    from dreamer_pytorch_b200.functional import KLFn

    g = torch.Generator().manual_seed(61)
    M, S = 2500, 30
    qm, pm = torch.randn(M, S, generator=g) * scale, torch.randn(M, S, generator=g) * scale
    ps = torch.rand(M, S, generator=g) + 0.6
    qs = torch.rand(M, S, generator=g) + 0.1 if scale == 1.0 else ps + 0.01 * torch.rand(M, S, generator=g)
    leaves = [t.double().requires_grad_(True) for t in (qm, qs, pm, ps)]
    raw = torch.mean(O.kl_normal(*leaves))
    div = torch.max(raw, raw.new_full((), 3.0))
    rg = torch.autograd.grad(div, leaves)
    cl = [t.to(cuda).requires_grad_(True) for t in (qm, qs, pm, ps)]
    cdiv, craw, pent, qent = KLFn.apply(*cl, 3.0)
    cg = torch.autograd.grad(cdiv, cl)
    assert rel_err(cdiv, div) < FP32_TOL and rel_err(craw, raw) < FP32_TOL
    assert rel_err(pent, torch.mean(O.normal_entropy(leaves[3]))) < FP32_TOL
    assert rel_err(qent, torch.mean(O.normal_entropy(leaves[1]))) < FP32_TOL
    clamped = float(raw) < 3.0
    for a, b in zip(cg, rg):
        if clamped:
            assert float(a.abs().max()) == 0.0 and float(b.abs().max()) == 0.0
        else:
            assert rel_err(a, b) < FP32_TOL


# ------------------------------------------------------------------------------------------------
# the whole loss section against golden outputs of the unmodified reference
# ------------------------------------------------------------------------------------------------
def _build_model(c, w, dev):
    from dreamer_pytorch_b200.dropin.agent import AtariDreamerModel

    model = AtariDreamerModel(action_shape=(c["A"],), action_dist=c["dist"], use_pcont=c["use_pcont"], **model_kwargs(c))
    model.load_state_dict(O.expand_aliases(w))
    return model.to(dev).train()


def _canon_grads(model):
    out = {}
    for k, p in model.named_parameters():  # named_parameters de-duplicates shared tensors
        if p.grad is not None and not k.startswith("observation_"):
            out[k] = p.grad
    return out


@pytest.mark.parametrize("name", list(CASES))
def test_dreamer_loss_matches_reference_golden(cuda, name):
    from dreamer_pytorch_b200 import noise
    from dreamer_pytorch_b200.algo import dreamer_loss

    c = CASES[name]
    gold = load_golden(name)
    w, obs, act, rew, done, nz = case_inputs(c)
    model = _build_model(c, w, cuda)
    observation = obs[:-1].to(cuda).float() / 255.0 - 0.5
    with noise.injected(eps_prior=[nz["eps_prior_obs"], nz["eps_prior_img"]], eps_post=[nz["eps_post_obs"]],
                        noise_act=[nz["noise_act"]]):
        model_loss, actor_loss, value_loss, info = dreamer_loss(
            model, observation, act[1:].to(cuda), rew[1:].unsqueeze(2).to(cuda), done.unsqueeze(2).to(cuda),
            horizon=c["H"], use_pcont=c["use_pcont"])
    model_loss.backward()
    actor_loss.backward()
    value_loss.backward()
    for k, v in (("model_loss", model_loss), ("actor_loss", actor_loss), ("value_loss", value_loss),
                 ("reward_loss", info.reward_loss), ("pcont_loss", info.pcont_loss), ("div", info.divergence),
                 ("prior_ent", info.prior_entropy), ("post_ent", info.post_entropy)):
        assert rel_err(v, gold[k]) < FP32_TOL, (k, float(v), float(gold[k]))
    grads = _canon_grads(model)
    assert set(grads) == set(gold["grads"]), set(grads) ^ set(gold["grads"])
    for k, gten in grads.items():
        ref = gold["grads"][k]
        if gold["grads_full"]:
            assert rel_err(gten, ref) < FP32_TOL, k
        else:
            assert abs(float(gten.abs().max()) - float(ref[0])) <= FP32_TOL * max(float(ref[0]), 1e-12), k


@pytest.mark.parametrize("name", ["tanh_default", "onehot_pcont_small"])
def test_rollouts_match_reference_golden(cuda, name):
    from dreamer_pytorch_b200 import noise
    from dreamer_pytorch_b200.dropin.rnns import RSSMState

    c = CASES[name]
    gold = load_golden(name)
    w, obs, act, rew, done, nz = case_inputs(c)
    model = _build_model(c, w, cuda)
    T, B = c["T"], c["B"]
    with torch.no_grad():
        embed = model.observation_encoder(obs[:-1].to(cuda).float() / 255.0 - 0.5)
        prev = model.representation.initial_state(B, device=cuda)
        with noise.injected(eps_prior=[nz["eps_prior_obs"], nz["eps_prior_img"]], eps_post=[nz["eps_post_obs"]],
                            noise_act=[nz["noise_act"]]):
            prior, post = model.rollout.rollout_representation(T, embed, act[1:].to(cuda), prev)
            src = post[:-1] if c["use_pcont"] else post
            flat = RSSMState(*(f.reshape(-1, f.shape[-1]) for f in src))
            imag, actions = model.rollout.rollout_policy(c["H"], model.policy, flat)
    for nm, st in (("prior", prior), ("post", post), ("imag", imag)):
        for f in ("mean", "std", "stoch", "deter"):
            # the conv embedding itself differs (cuDNN vs CPU) at the 1e-6 level; budget unchanged
            assert rel_err(getattr(st, f), gold[f"{nm}_{f}"]) < FP32_TOL, (nm, f)
    assert rel_err(actions, gold["actions"]) < FP32_TOL


def test_full_size_loss_section_vs_oracle(cuda):
    """BASELINE.json configs[0]: cartpole-sized RSSM, B=50 T=50 H=15 (N=2500), RSSM only (synthetic
    obs_embed).  Oracle runs on the GPU box's CPU in fp32; losses and every gradient must agree."""
    from dreamer_pytorch_b200 import noise
    from dreamer_pytorch_b200.algo import dreamer_loss

    torch.set_num_threads(max(1, min(16, torch.get_num_threads())))
    c = dict(A=1, dist="tanh_normal", use_pcont=False, T=50, B=50, H=15, S=30, D=200, Hd=200, act_hidden=200,
             reward_hidden=300, value_hidden=200, pcont_hidden=200)
    shapes = O.weight_shapes(1, conv=True)
    w = O.make_weights(shapes, 0)
    g = torch.Generator().manual_seed(1)
    embed = torch.randn(50, 50, 1024, generator=g)
    action = torch.rand(50, 50, 1, generator=g) * 2 - 1
    reward = torch.randn(50, 50, 1, generator=g)
    nz = O.make_noise(50, 50, 2500, 15, 30, 1, 2)
    wl = {k: v.clone().requires_grad_(True) for k, v in w.items()}
    all_action = torch.cat([action[:1], action])
    all_reward = torch.cat([reward[:1, :, 0], reward[:, :, 0]])
    out = O.dreamer_loss(wl, None, all_action, all_reward, None, nz, horizon=15, obs_embed=embed)
    names = [k for k in wl if not k.startswith("observation_")]
    total = out["model_loss"] + out["actor_loss"] + out["value_loss"]
    rgrads = dict(zip(names, torch.autograd.grad(total, [wl[k] for k in names], allow_unused=True)))

    model = _build_model(c, w, cuda)
    with noise.injected(eps_prior=[nz["eps_prior_obs"], nz["eps_prior_img"]], eps_post=[nz["eps_post_obs"]],
                        noise_act=[nz["noise_act"]]):
        ml, al, vl, info = dreamer_loss(model, None, action.to(cuda), reward.to(cuda), None, horizon=15,
                                        obs_embed=embed.to(cuda))
    ml.backward(); al.backward(); vl.backward()
    assert rel_err(ml, out["model_loss"]) < FP32_TOL
    assert rel_err(al, out["actor_loss"]) < FP32_TOL
    assert rel_err(vl, out["value_loss"]) < FP32_TOL
    cg = _canon_grads(model)
    for k, gr in rgrads.items():
        if gr is None:
            continue
        assert k in cg, k
        assert rel_err(cg[k], gr) < FP32_TOL, (k, rel_err(cg[k], gr))


@pytest.mark.parametrize("prec,tol", [("fp32", FP32_TOL), ("bf16", 1e-2)])
def test_scaled_rssm_config_vs_oracle(cuda, prec, tol):
    """BASELINE.json configs[4]: stoch 60 / deter 600 / hidden 600 (GRU weights alone 2.16 M parameters: beyond what the
    tcgen05 rollout keeps per CTA, and 660 input features beyond the tcgen05 MLP's 512), heads at the reference defaults.
    Both precisions run the fp32 kernels there; the whole loss section against the fp64 oracle."""
    from dreamer_pytorch_b200 import noise
    from dreamer_pytorch_b200.algo import dreamer_loss
    from dreamer_pytorch_b200.dropin.agent import AtariDreamerModel
    from dreamer_pytorch_b200.precision import precision

    A, S, D, Hd, T, B, H = 6, 60, 600, 600, 4, 5, 6
    w = O.make_weights(O.weight_shapes(A, S=S, D=D, Hd=Hd, conv=True), 21)
    g = torch.Generator().manual_seed(22)
    embed = torch.randn(T, B, 1024, generator=g)
    action = torch.rand(T, B, A, generator=g) * 2 - 1
    reward = torch.randn(T, B, 1, generator=g)
    nz = O.make_noise(T, B, T * B, H, S, A, 23)
    wl = {k: v.double().requires_grad_(True) for k, v in w.items()}
    nz64 = {k: v.double() for k, v in nz.items()}
    ref = O.dreamer_loss(wl, None, torch.cat([action[:1], action]).double(),
                         torch.cat([reward[:1, :, 0], reward[:, :, 0]]).double(), None, nz64, horizon=H,
                         obs_embed=embed.double())
    names = [k for k in wl if not k.startswith("observation_")]
    rgrads = dict(zip(names, torch.autograd.grad(ref["model_loss"] + ref["actor_loss"] + ref["value_loss"],
                                                 [wl[k] for k in names], allow_unused=True)))
    model = AtariDreamerModel(action_shape=(A,), action_dist="tanh_normal", stochastic_size=S, deterministic_size=D,
                              hidden_size=Hd)
    model.load_state_dict(O.expand_aliases(w))
    model = model.to(cuda).train()
    with precision(prec), noise.injected(eps_prior=[nz["eps_prior_obs"], nz["eps_prior_img"]],
                                         eps_post=[nz["eps_post_obs"]], noise_act=[nz["noise_act"]]):
        ml, al, vl, _ = dreamer_loss(model, None, action.to(cuda), reward.to(cuda), None, horizon=H,
                                     obs_embed=embed.to(cuda))
        ml.backward(); al.backward(); vl.backward()
    for got, want in ((ml, ref["model_loss"]), (al, ref["actor_loss"]), (vl, ref["value_loss"])):
        assert rel_err(got, want) < tol
    cg = _canon_grads(model)
    for k, gr in rgrads.items():
        if gr is not None:
            assert rel_err(cg[k], gr) < tol, (k, rel_err(cg[k], gr))
