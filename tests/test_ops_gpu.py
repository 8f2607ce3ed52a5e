"""torch.ops.dreamer_b200.* (SURVEY 8(b) "Torch op layer") against the module path on identical noise: same values,
same gradients, FreezeParameters honoured."""
import pytest
import torch

from _cases import FP32_TOL, rel_err

pytestmark = pytest.mark.gpu


def _model(dev, dist="tanh_normal"):
    from dreamer_pytorch_b200.dropin.agent import AtariDreamerModel

    torch.manual_seed(0)
    return AtariDreamerModel(action_shape=(6,), action_dist=dist).to(dev).train()


def test_ops_are_registered():
    import dreamer_pytorch_b200.ops as ops

    for name in ("transition_step", "transition_rollout", "representation_rollout", "policy_rollout", "mlp", "kl_entropy",
                 "heads_returns_loss", "image_logprob"):
        assert name in ops.OP_NAMES and hasattr(torch.ops.dreamer_b200, name)


def test_rollout_ops_match_modules_and_differentiate(cuda):
    import dreamer_pytorch_b200.ops  # noqa: F401  (registers the ops)
    from dreamer_pytorch_b200 import noise
    from dreamer_pytorch_b200.dropin.module import FreezeParameters
    from dreamer_pytorch_b200.dropin.rnns import RSSMState
    from dreamer_pytorch_b200.precision import precision

    O = torch.ops.dreamer_b200
    m = _model(cuda)
    T, B, H, S, A = 6, 5, 4, 30, 6
    g = torch.Generator().manual_seed(1)
    embed = torch.randn(T, B, 1024, generator=g).to(cuda)
    act = (torch.rand(T, B, A, generator=g) * 2 - 1).to(cuda)
    ep, eq = torch.randn(T, B, S, generator=g).to(cuda), torch.randn(T, B, S, generator=g).to(cuda)
    na, ei = torch.randn(H, T * B, A, generator=g).to(cuda), torch.randn(H, T * B, S, generator=g).to(cuda)
    prev = m.representation.initial_state(B, device=cuda)
    rep_w, tr_w, actor_w = m.representation.weights(), m.transition.weights(), m.action_decoder.weights()
    dec = m.action_decoder
    with precision("fp32"):
        # module path
        with noise.injected(eps_prior=[ep, ei], eps_post=[eq], noise_act=[na]):
            prior, post = m.rollout.rollout_representation(T, embed, act, prev)
            flat = RSSMState(*(f.reshape(-1, f.shape[-1]).detach() for f in post))
            with FreezeParameters([m.transition]):
                imag, actions = m.rollout.rollout_policy(H, m.policy, flat)
        loss = post.stoch.square().mean() + prior.std.mean() + imag.deter.mean() + actions.square().mean()
        params = list(m.representation.parameters()) + list(dec.parameters())
        ref = torch.autograd.grad(loss, params, allow_unused=True)
        # op path
        pm, ps, pst, deter, qm, qs, qst = O.representation_rollout(prev.stoch, prev.deter, act, embed, ep, eq, rep_w)
        with FreezeParameters([m.transition]):
            frozen = [w for w in tr_w]
            im, isd, ist, idt, iact = O.policy_rollout(qst.reshape(-1, S), deter.reshape(-1, 200), na, ei, frozen, actor_w,
                                                       "tanh_normal", H, float(dec.mean_scale), float(dec.raw_init_std),
                                                       float(dec.min_std))
        loss2 = qst.square().mean() + ps.mean() + idt.mean() + iact.square().mean()
        got = torch.autograd.grad(loss2, params, allow_unused=True)
        assert torch.equal(qst, post.stoch) and torch.equal(pst, prior.stoch) and torch.equal(iact, actions)
        for a, b, p in zip(got, ref, params):
            assert (a is None) == (b is None)
            if a is not None:
                assert rel_err(a, b) < FP32_TOL, tuple(p.shape)
        # single step == first step of the rollout
        s1 = O.transition_step(act[0], prev.stoch, prev.deter, ep[0], tr_w)
        s_roll = O.transition_rollout(prev.stoch, prev.deter, act[:1], ep[:1], tr_w)
        assert all(torch.equal(a, b[0]) for a, b in zip(s1, s_roll))
        assert rel_err(s1[3], deter[0]) < 1e-6  # the prior half of the first representation step


def test_head_and_loss_ops(cuda):
    import dreamer_pytorch_b200.ops  # noqa: F401
    from dreamer_pytorch_b200.precision import precision

    O = torch.ops.dreamer_b200
    m = _model(cuda)
    g = torch.Generator().manual_seed(2)
    feat = torch.randn(40, 230, generator=g).to(cuda)
    with precision("fp32"):
        y = O.mlp(feat[:, :30].contiguous(), feat[:, 30:].contiguous(), [p for p in m.value_model.model.parameters()])
        want = m.value_model.model(feat)
        assert rel_err(y, want) < FP32_TOL
        (gw,) = torch.autograd.grad(y.sum(), [m.value_model.model[0].weight])
        (gr,) = torch.autograd.grad(want.sum(), [m.value_model.model[0].weight])
        assert rel_err(gw, gr) < FP32_TOL
    mq, sq = torch.randn(12, 30, generator=g).to(cuda), torch.rand(12, 30, generator=g).to(cuda) + 0.1
    mp, sp = torch.randn(12, 30, generator=g).to(cuda), torch.rand(12, 30, generator=g).to(cuda) + 0.1
    div, raw, ent_p, ent_q = O.kl_entropy(mq, sq, mp, sp, 3.0)
    import torch.distributions as td
    kl = td.kl_divergence(td.Independent(td.Normal(mq, sq), 1), td.Independent(td.Normal(mp, sp), 1)).mean()
    assert rel_err(raw, kl) < FP32_TOL and rel_err(div, torch.clamp(kl, min=3.0)) < FP32_TOL
    r, v = torch.randn(15, 64, generator=g).to(cuda), torch.randn(15, 64, generator=g).to(cuda)
    out = O.heads_returns_loss(r, v, None, 0.99, 0.95)
    assert out[0].shape == (2,) and out[1].shape == (14, 64) and out[6].numel() == 0 and torch.isfinite(out[0]).all()
    mean, obs = torch.randn(5, 3 * 8 * 8, generator=g).to(cuda), torch.rand(5, 3 * 8 * 8, generator=g).to(cuda)
    lp = O.image_logprob(mean, obs)
    want = td.Independent(td.Normal(mean, 1.0), 1).log_prob(obs)
    assert rel_err(lp, want) < FP32_TOL
