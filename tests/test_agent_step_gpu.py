"""the environment-time agent step.  The fused SampleDist.mode() kernel against torch.distributions on the
same draws, and the whole step (encoder -> posterior step -> actor -> mode) replayed from one CUDA graph against the eager
drop-in forward."""
import pytest
import torch

from _cases import FP32_TOL, rel_err

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("n,A", [(1, 6), (7, 1), (64, 6), (33, 17)])
def test_fused_mode_matches_torch_distributions(cuda, n, A):
    import torch.distributions as td

    from dreamer_pytorch_b200.functional import RAW_INIT_STD, policy_mode
    from dreamer_pytorch_b200 import noise

    g = torch.Generator().manual_seed(9)
    raw = (torch.randn(n, 2 * A, generator=g) * 2).to(cuda)
    eps = torch.randn(100, n, A, generator=g).to(cuda)
    with noise.injected(mode_noise=[eps]):
        action, logp = policy_mode(raw, 100, 5.0, RAW_INIT_STD, 1e-4, return_logp=True)
    # the reference's formulas (action.py:54-61, distribution.py:24-35) in torch fp32 on the same draws: tanh saturates
    # in fp32 for |z| > 9 and TanhTransform._inverse clamps, so the comparison has to run in the reference's own precision
    rm, rs = raw.chunk(2, -1)
    mean = 5.0 * torch.tanh(rm / 5.0)
    std = torch.nn.functional.softplus(rs + RAW_INIT_STD) + 1e-4
    dist = td.Independent(td.TransformedDistribution(td.Normal(mean, std), td.TanhTransform()), 1)
    wide = dist.expand((100, n))
    sample = torch.tanh(mean + std * eps)
    ref_lp = wide.log_prob(sample).double()
    ok = torch.isfinite(ref_lp)
    assert ok.any()
    assert float((logp.double() - ref_lp)[ok].abs().max()) < 2e-3 * max(1.0, float(ref_lp[ok].abs().max()))
    assert torch.equal(torch.isnan(logp), torch.isnan(ref_lp))  # saturated draws: NaN in the reference, NaN here
    # the winner, torch.argmax semantics (NaN = maximum, first occurrence): identical wherever it is unambiguous --
    # a row with a NaN draw, or one whose top two finite log-probs are further apart than the fp32 noise
    ref_idx = torch.argmax(ref_lp, dim=0)
    has_nan = torch.isnan(ref_lp).any(0)
    top2 = torch.nan_to_num(ref_lp, nan=-1e30, neginf=-1e30).topk(2, dim=0)
    clear = has_nan | ((top2.values[0] - top2.values[1]) > 1e-2)
    want = sample.double().gather(0, ref_idx.reshape(1, n, 1).expand(1, n, A)).squeeze(0)
    assert clear.any()
    assert rel_err(action[clear], want[clear]) < FP32_TOL


def test_sample_dist_mode_uses_the_fused_kernel_and_falls_through_with_grad(cuda):
    from dreamer_pytorch_b200.dropin.action import ActionDecoder

    dec = ActionDecoder(6, 230, 200, 3, "tanh_normal").to(cuda)
    feat = torch.randn(5, 230, device=cuda)
    with torch.no_grad():
        a = dec(feat).mode()
    assert a.shape == (5, 6) and bool((a.abs() <= 1).all())
    b = dec(feat).mode()  # grad enabled and parameters require grad: the reference's torch path
    assert b.shape == (5, 6)


@pytest.mark.parametrize("training", [False, True])
def test_env_step_graph_matches_eager_forward(cuda, training):
    from dreamer_pytorch_b200 import noise
    from dreamer_pytorch_b200.agent_step import EnvStepGraph
    from dreamer_pytorch_b200.dropin.agent import AtariDreamerModel

    torch.manual_seed(2)
    model = AtariDreamerModel(action_shape=(6,), action_dist="tanh_normal").to(cuda)
    model.train(training)
    g = torch.Generator().manual_seed(3)
    obs = torch.randint(0, 256, (2, 1, 3, 64, 64), generator=g, dtype=torch.uint8).to(cuda)
    eps_p, eps_q = torch.randn(1, 1, 30, generator=g).to(cuda), torch.randn(1, 1, 30, generator=g).to(cuda)
    nz_mode = torch.randn(100, 1, 6, generator=g).to(cuda)
    nz_act = torch.randn(1, 6, generator=g).to(cuda)

    def inj(k):
        d = dict(eps_prior=[eps_p] * k, eps_post=[eps_q] * k)
        if not training:
            d["mode_noise"] = [nz_mode] * k
        return d

    import torch.distributions.normal as tdn

    orig = tdn._standard_normal
    if training:  # train mode draws the action with dist.rsample(): pin that draw too
        tdn._standard_normal = lambda shape, dtype, device: nz_act.to(device=device, dtype=dtype).reshape(shape)
    try:
        with noise.injected(**inj(4)):  # 3 warm-up passes + the capture: the injected tensors become the graph's noise
            graph = EnvStepGraph(model, batch=1)
        outs = []
        for t in range(2):
            a, st = graph.step(obs[t])
            outs.append((a.clone(), [f.clone() for f in st]))
        # eager: the same two steps with the same noise, state and action carried by hand
        with torch.no_grad(), noise.injected(**inj(2)):
            prev_a = torch.zeros(1, 6, device=cuda)
            prev_s = None
            for t in range(2):
                out = model(obs[t], prev_a, prev_s)
                assert rel_err(outs[t][0], out.action) < 1e-6
                for fa, fb in zip(outs[t][1], out.state):
                    assert rel_err(fa, fb) < 1e-6
                prev_a, prev_s = out.action, out.state
    finally:
        tdn._standard_normal = orig
