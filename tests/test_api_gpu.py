"""API-contract tests of the drop-in modules on CUDA tensors: the reference's own shape tests
(tests/test_rnns.py, test_action.py, test_agent.py, test_dense.py, test_distribution.py in the
reference checkout) restated for our modules, plus size-independent properties of the kernels at
the full BASELINE.json size."""
import pytest
import torch

pytestmark = pytest.mark.gpu


def _models(dev, action_size=10, obs_embed=100, stoch=30, deter=200, hidden=200):
    from dreamer_pytorch_b200.dropin.rnns import RSSMRepresentation, RSSMRollout, RSSMTransition

    tr = RSSMTransition(action_size, stoch, deter, hidden).to(dev)
    rep = RSSMRepresentation(tr, obs_embed, action_size, stoch, deter, hidden).to(dev)
    return tr, rep, RSSMRollout(rep, tr).to(dev)


def test_rssm_step_shapes(cuda):  # reference tests/test_rnns.py:12-35
    tr, rep, _ = _models(cuda)
    init = rep.initial_state(4, device=cuda)
    prior, post = rep(torch.randn(4, 100, device=cuda), torch.randn(4, 10, device=cuda), init)
    for st in (prior, post):
        assert st.stoch.shape == (4, 30) and st.deter.shape == (4, 200) and st.mean.shape == (4, 30)
    assert torch.equal(prior.deter, post.deter)
    assert (prior.std > 0.1).all() and (post.std > 0.1).all()


def test_rollout_shapes_and_free_callable_policy(cuda):  # reference tests/test_rnns.py:38-98
    from dreamer_pytorch_b200.dropin.distribution import SampleDist
    from dreamer_pytorch_b200.dropin.rnns import RSSMState

    tr, rep, roll = _models(cuda)
    init = rep.initial_state(4, device=cuda)
    obs, act = torch.randn(10, 4, 100, device=cuda), torch.randn(10, 4, 10, device=cuda)
    prior, post = roll(10, obs, act, init)
    assert isinstance(prior, RSSMState) and isinstance(post, RSSMState)
    for st in (prior, post):
        assert st.mean.shape == (10, 4, 30) and st.std.shape == (10, 4, 30)
        assert st.stoch.shape == (10, 4, 30) and st.deter.shape == (10, 4, 200)
    pr = roll.rollout_transition(10, act, init)
    assert pr.stoch.shape == (10, 4, 30) and pr.deter.shape == (10, 4, 200)

    def policy(state):  # an arbitrary Python closure, as in the reference test
        a = torch.randn(state.stoch.shape[0], 10, device=cuda)
        d = torch.distributions.Independent(torch.distributions.Normal(a, 1.0), 1)
        return a, SampleDist(d)

    nxt, actions = roll.rollout_policy(10, policy, post[-1])
    assert nxt.stoch.shape == (10, 4, 30) and nxt.deter.shape == (10, 4, 200) and actions.shape == (10, 4, 10)


@pytest.mark.parametrize("dist", ["tanh_normal", "one_hot", "relaxed_one_hot"])
def test_action_decoder(cuda, dist):  # reference tests/test_action.py:7-57
    from dreamer_pytorch_b200.dropin.action import ActionDecoder

    dec = ActionDecoder(10, 20, 24, 2, dist).to(cuda)
    d = dec(torch.randn(4, 20, device=cuda))
    if dist == "tanh_normal":
        m, mo, e = d.mean(), d.mode(), d.entropy()
        assert m.shape == (4, 10) and mo.shape == (4, 10) and e.shape == (4,)
        (m.sum() + mo.sum() + e.sum()).backward()
    else:
        s = d.sample() if dist == "one_hot" else d.rsample()
        assert s.shape == (4, 10)
        d.logits.sum().backward() if dist == "one_hot" else s.sum().backward()
    assert all(p.grad is not None for p in dec.parameters())


@pytest.mark.parametrize("dist", ["tanh_normal", "one_hot", "relaxed_one_hot"])
def test_agent_model(cuda, dist):  # reference tests/test_agent.py:7-37
    from dreamer_pytorch_b200.dropin.agent import AgentModel

    m = AgentModel(action_shape=(2,), action_dist=dist).to(cuda)
    obs = torch.randn(1, 3, 64, 64, device=cuda)
    st = m.get_state_representation(obs, None, None)
    assert st.deter.shape == (1, 200) and st.stoch.shape == (1, 30)
    a, _ = m.policy(st)
    assert a.shape == (1, 2)
    with pytest.raises(NotImplementedError):
        m(obs, None, None)
    st2 = m.get_state_transition(a, st)
    assert st2.deter.shape == (1, 200)
    m.eval()
    a, _ = m.policy(st)
    assert a.shape == (1, 2)


@pytest.mark.parametrize("dist", ["normal", "binary"])
def test_dense_model(cuda, dist):  # reference tests/test_dense.py:7-25
    from dreamer_pytorch_b200.dropin.dense import DenseModel

    d = DenseModel(20, (1,), 3, 24, dist).to(cuda)
    out = d(torch.randn(2, 20, device=cuda))
    assert out.sample().shape == (2, 1)
    out2 = d(torch.randn(5, 2, 20, device=cuda))
    assert out2.mean.shape == (5, 2, 1)


def test_atari_model_forward(cuda):  # agent.py:167-188 sampling-time step
    from dreamer_pytorch_b200.dropin.agent import AtariDreamerModel

    m = AtariDreamerModel(action_shape=(6,), action_dist="one_hot").to(cuda).eval()
    with torch.no_grad():
        out = m(torch.randint(0, 255, (3, 64, 64), dtype=torch.uint8, device=cuda), torch.zeros(6, device=cuda), None)
    assert out.action.shape == (6,) and out.state.deter.shape == (200,)


def test_cpu_tensors_raise(cuda):
    tr, rep, _ = _models("cpu")
    with pytest.raises(RuntimeError):
        tr(torch.randn(4, 10), tr.initial_state(4))


# ------------------------------------------------------------------------------------------------
# size-independent properties at BASELINE.json's full size (N = 2500 start states, H = 15)
# ------------------------------------------------------------------------------------------------
def _full(dev):
    from dreamer_pytorch_b200.dropin.agent import AtariDreamerModel

    torch.manual_seed(0)
    m = AtariDreamerModel(action_shape=(6,), action_dist="tanh_normal").to(dev).train()
    g = torch.Generator().manual_seed(7)
    N, H = 2500, 15
    stoch, deter = torch.randn(N, 30, generator=g).to(dev), torch.tanh(torch.randn(N, 200, generator=g)).to(dev)
    na, ep = torch.randn(H, N, 6, generator=g).to(dev), torch.randn(H, N, 30, generator=g).to(dev)
    return m, stoch, deter, na, ep


def test_rows_are_independent_and_sharding_is_exact(cuda):
    """Sharding the start states across ranks must not change any row (this is what makes the multi-GPU split
    collective-free): rollout(all)[rows] is bit-identical to rollout(rows) whenever both run with the same tile height
    (equal shards do), wherever the rows fall inside a tile.  With another tile height the fp32 kernels split the
    reduction index differently (csrc/drm_common.cuh, split K), so the same row agrees to fp32 rounding instead."""
    from dreamer_pytorch_b200 import noise
    from dreamer_pytorch_b200.dropin.rnns import RSSMState

    m, stoch, deter, na, ep = _full(cuda)
    z = torch.zeros_like(stoch)

    def shard(lo, hi):
        with noise.injected(noise_act=[na[:, lo:hi].contiguous()], eps_prior=[ep[:, lo:hi].contiguous()]):
            return m.rollout.rollout_policy(15, m.policy, RSSMState(z[lo:hi], z[lo:hi], stoch[lo:hi], deter[lo:hi]))

    with torch.no_grad():
        full, act = shard(0, 2500)
        lo, hi = 107, 2500  # 2 393 rows: 20-row tiles like the whole batch, shifted by 7 rows inside the tiles
        part, pact = shard(lo, hi)
        assert torch.equal(full.stoch[:, lo:hi], part.stoch) and torch.equal(full.deter[:, lo:hi], part.deter)
        assert torch.equal(act[:, lo:hi], pact)
        lo, hi = 1250, 2500  # half the batch: 12-row tiles
        part, pact = shard(lo, hi)
        for a, b in ((full.stoch[:, lo:hi], part.stoch), (full.deter[:, lo:hi], part.deter), (act[:, lo:hi], pact)):
            assert float((a - b).abs().max()) < 2e-5 * float(a.abs().max())
    assert torch.isfinite(full.deter).all() and (full.deter.abs() <= 1).all() and (act.abs() <= 1).all()


def test_multi_step_rollout_equals_chained_single_steps(cuda):
    """rollout_transition(T) == T chained RSSMTransition.forward calls, bit for bit."""
    from dreamer_pytorch_b200 import noise
    from dreamer_pytorch_b200.dropin.rnns import RSSMState

    m, stoch, deter, na, ep = _full(cuda)
    z = torch.zeros_like(stoch)
    T = 5
    acts = torch.tanh(na[:T])
    with torch.no_grad():
        with noise.injected(eps_prior=[ep[:T].contiguous()]):
            multi = m.rollout.rollout_transition(T, acts, RSSMState(z, z, stoch, deter))
        st = RSSMState(z, z, stoch, deter)
        for t in range(T):
            with noise.injected(eps_prior=[ep[t:t + 1].contiguous()]):
                st = m.transition(acts[t], st)
            assert torch.equal(st.stoch, multi.stoch[t]) and torch.equal(st.deter, multi.deter[t])
