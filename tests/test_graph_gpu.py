"""ImaginationGraph (one CUDA graph per imagination step) against the eager fused path on identical noise."""
import pytest
import torch

pytestmark = pytest.mark.gpu


def rel_err(a, b):
    a, b = a.detach().double().cpu(), b.detach().double().cpu()
    return float((a - b).abs().max() / b.abs().max().clamp_min(1e-30))


@pytest.mark.parametrize("prec", ["bf16", "fp32"])
def test_graph_step_matches_eager(prec):
    from dreamer_pytorch_b200 import noise
    from dreamer_pytorch_b200.algo import imagination_losses
    from dreamer_pytorch_b200.dropin.agent import AtariDreamerModel
    from dreamer_pytorch_b200.dropin.rnns import RSSMState
    from dreamer_pytorch_b200.graph import ImaginationGraph
    from dreamer_pytorch_b200.precision import precision

    dev = torch.device("cuda:0")
    torch.manual_seed(0)
    A, N, H = 6, 300, 5
    model = AtariDreamerModel(action_shape=(A,), action_dist="tanh_normal").to(dev).train()
    g = torch.Generator().manual_seed(1)
    stoch = torch.randn(N, 30, generator=g).to(dev)
    deter = torch.tanh(torch.randn(N, 200, generator=g)).to(dev)
    eps = torch.randn(H, N, 30, generator=g).to(dev)
    nza = torch.randn(H, N, A, generator=g).to(dev)
    params = list(model.action_decoder.parameters()) + list(model.value_model.parameters())
    with precision(prec):
        with noise.injected(eps_prior=[eps], noise_act=[nza]):
            al, vl, _ = imagination_losses(model, RSSMState(stoch, stoch, stoch, deter), H)
            al.backward(); vl.backward()
        ref = [p.grad.clone() for p in params]
        ref_losses = (al.detach().clone(), vl.detach().clone())
        del al, vl  # frees the eager autograd graph (its AccumulateGrad nodes are bound to the default stream)
        warm = 2
        with noise.injected(eps_prior=[eps] * (warm + 1), noise_act=[nza] * (warm + 1)):
            graph = ImaginationGraph(model, N, H, warmup=warm)
        assert graph.launches_per_replay > 5
        for p in params:
            p.grad.zero_()
        host_s, host_d = stoch.cpu().pin_memory(), deter.cpu().pin_memory()
        for _ in range(2):  # replays are idempotent: gradients are overwritten, not accumulated
            gal, gvl = graph.step(host_s, host_d)
    torch.cuda.synchronize()
    assert rel_err(gal, ref_losses[0]) < 1e-5 and rel_err(gvl, ref_losses[1]) < 1e-5
    for p, r in zip(params, ref):
        assert rel_err(p.grad, r) < 1e-4, (p.shape, rel_err(p.grad, r))  # fp32 atomics reorder between runs


@pytest.mark.parametrize("use_pcont", [False, True])
def test_kernels_beside_each_other_match_stream_order(use_pcont):
    """Forward: head MLPs gated by the rollout's per-step completion counters.  Backward: weight packing issued beside the
    forward pass, head data-gradient kernels beside the reverse-time rollout (tile hand-over between heads, per-step row
    counters towards the rollout).  All on side streams with capped grids; they must give the results of the plain
    stream-ordered launch sequence: same kernels, same data, only the schedule differs."""
    from dreamer_pytorch_b200 import algo, noise
    from dreamer_pytorch_b200.dropin.agent import AtariDreamerModel
    from dreamer_pytorch_b200.dropin.rnns import RSSMState
    from dreamer_pytorch_b200.precision import precision

    dev = torch.device("cuda:0")
    torch.manual_seed(3)
    A, N, H = 6, 1000, 15
    model = AtariDreamerModel(action_shape=(A,), action_dist="tanh_normal", use_pcont=use_pcont).to(dev).train()
    g = torch.Generator().manual_seed(4)
    stoch = torch.randn(N, 30, generator=g).to(dev)
    deter = torch.tanh(torch.randn(N, 200, generator=g)).to(dev)
    eps = torch.randn(H, N, 30, generator=g).to(dev)
    nza = torch.randn(H, N, A, generator=g).to(dev)
    params = list(model.action_decoder.parameters()) + list(model.value_model.parameters())
    with precision("bf16"):  # the first pass on a device always runs in plain stream order (algo._WARM): get it out of the way
        al, vl, _ = algo.imagination_losses(model, RSSMState(stoch, stoch, stoch, deter), H, use_pcont=use_pcont)
        al.backward(); vl.backward()
    assert any(k[:3] == ("cuda", 0, "bf16") and algo._overlap_ok(k) for k in algo._WARM)
    res = {}
    flags = ("_PIPELINE_HEADS", "_PREPARE_BACKWARD", "_PIPELINE_BACKWARD")
    old = {k: getattr(algo, k) for k in flags}
    try:
        # everything beside everything / plain stream order / each mechanism alone
        for mode in ((True, True, True), (False, False, False), (True, False, False), (False, True, False),
                     (False, False, True)):
            for k, v in zip(flags, mode):
                setattr(algo, k, v)
            for p in params:
                p.grad = None
            with precision("bf16"), noise.injected(eps_prior=[eps], noise_act=[nza]):
                al, vl, info = algo.imagination_losses(model, RSSMState(stoch, stoch, stoch, deter), H,
                                                       use_pcont=use_pcont)
                al.backward(); vl.backward()
            torch.cuda.synchronize()
            res[mode] = (al.detach().clone(), vl.detach().clone(), info.reward.clone(), info.value.clone(),
                         [p.grad.clone() for p in params])
    finally:
        for k, v in old.items():
            setattr(algo, k, v)
    b = res[(False, False, False)]
    for mode, a in res.items():
        assert torch.equal(a[2], b[2]) and torch.equal(a[3], b[3]), mode      # head outputs: bit-identical
        assert rel_err(a[0], b[0]) < 1e-6 and rel_err(a[1], b[1]) < 1e-6, mode
        for ga, gb in zip(a[4], b[4]):
            assert rel_err(ga, gb) < 1e-4, mode                                 # fp32 atomics reorder between runs
