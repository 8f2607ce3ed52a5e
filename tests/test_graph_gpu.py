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


def test_two_passes_on_two_streams_of_one_device():
    """Single-tenant rule: a pass issued on another stream while the previous pass on the device is
    still in flight must not also count on idle SMs -- it is demoted to plain stream order (algo._exclusive) and both
    passes finish with the results they give alone; no bounded-spin trap, no hang."""
    from dreamer_pytorch_b200 import algo, noise
    from dreamer_pytorch_b200.dropin.agent import AtariDreamerModel
    from dreamer_pytorch_b200.dropin.rnns import RSSMState
    from dreamer_pytorch_b200.precision import precision

    dev = torch.device("cuda:0")
    torch.manual_seed(5)
    A, N, H = 6, 2500, 15
    model = AtariDreamerModel(action_shape=(A,), action_dist="tanh_normal").to(dev).train()
    actor_p, value_p = list(model.action_decoder.parameters()), list(model.value_model.parameters())
    g = torch.Generator().manual_seed(6)
    inputs = []
    for _ in range(2):
        inputs.append((torch.randn(N, 30, generator=g).to(dev), torch.tanh(torch.randn(N, 200, generator=g)).to(dev),
                       torch.randn(H, N, 30, generator=g).to(dev), torch.randn(H, N, A, generator=g).to(dev)))

    def forward(i):
        stoch, deter, eps, nza = inputs[i]
        with noise.injected(eps_prior=[eps], noise_act=[nza]):
            return algo.imagination_losses(model, RSSMState(stoch, stoch, stoch, deter), H)

    def backward(al, vl):
        return torch.autograd.grad(al, actor_p) + torch.autograd.grad(vl, value_p)

    with precision("bf16"):
        for _ in range(2):  # first pass of a configuration is plain anyway (algo._WARM); second one overlapped
            al, vl, _ = forward(0)
            backward(al, vl)
        alone = []
        for i in range(2):
            al, vl, info = forward(i)
            alone.append((al.detach().clone(), vl.detach().clone(), info.reward.clone(), backward(al, vl)))
        torch.cuda.synchronize()
        before = algo.demoted_passes
        s = [torch.cuda.Stream(), torch.cuda.Stream()]
        torch.cuda._sleep(200_000_000)  # ~0.1 s on the default stream: both streams start together, behind it
        for st in s:
            st.wait_stream(torch.cuda.current_stream())
        out = [None, None]
        for i in range(2):  # both forwards, then both backwards: everything is enqueued before anything runs
            with torch.cuda.stream(s[i]):
                out[i] = forward(i)
        grads = [None, None]
        for i in range(2):
            with torch.cuda.stream(s[i]):
                grads[i] = backward(out[i][0], out[i][1])
        for st in s:
            torch.cuda.current_stream().wait_stream(st)
        torch.cuda.synchronize()  # a trap in a spinning kernel would surface here as a CUDA error
    assert algo.demoted_passes - before >= 2  # the second stream's forward and at least one backward went plain
    for i in range(2):
        al, vl, info = out[i]
        assert torch.equal(info.reward, alone[i][2])  # head outputs: bit-identical whatever the schedule
        assert rel_err(al, alone[i][0]) < 1e-5 and rel_err(vl, alone[i][1]) < 1e-5  # one float atomic per CTA in the loss kernel
        for ga, gb in zip(grads[i], alone[i][3]):
            assert rel_err(ga, gb) < 1e-4  # fp32 atomics reorder between runs


@pytest.mark.parametrize("use_pcont", [False, True])
def test_whole_update_graph_matches_eager_update(use_pcont):
    """UpdateGraph (Dreamer.loss + 3 backwards + clip + Adam of the three groups, conv stack included, one replay) against the
    same update done eagerly -- dreamer_loss, .backward() x 3, ClipAdam.step() x 3 -- on identical noise, over two updates."""
    import copy

    from dreamer_pytorch_b200 import noise
    from dreamer_pytorch_b200.algo import dreamer_loss
    from dreamer_pytorch_b200.dropin.agent import AtariDreamerModel
    from dreamer_pytorch_b200.dropin.module import get_parameters
    from dreamer_pytorch_b200.graph import UpdateGraph
    from dreamer_pytorch_b200.optim import ClipAdam
    from dreamer_pytorch_b200.precision import precision
    from dreamer_pytorch_b200.replay import SamplesFromReplay

    dev = torch.device("cuda:0")
    torch.manual_seed(11)
    A, T, B, H, S = 6, 6, 8, 5, 30
    dist = "one_hot" if use_pcont else "tanh_normal"
    m_graph = AtariDreamerModel(action_shape=(A,), action_dist=dist, use_pcont=use_pcont).to(dev).train()
    m_eager = copy.deepcopy(m_graph)
    g = torch.Generator().manual_seed(12)
    obs = torch.randint(0, 256, (T + 1, B, 3, 64, 64), generator=g, dtype=torch.uint8).to(dev)
    if use_pcont:
        act = torch.nn.functional.one_hot(torch.randint(0, A, (T + 1, B), generator=g), A).float().to(dev)
    else:
        act = (torch.rand(T + 1, B, A, generator=g) * 2 - 1).to(dev)
    rew = torch.randn(T + 1, B, generator=g).to(dev)
    done = (torch.rand(T, B, generator=g) < 0.1).to(dev)
    N = (T - 1) * B if use_pcont else T * B
    eo, eq = torch.randn(T, B, S, generator=g).to(dev), torch.randn(T, B, S, generator=g).to(dev)
    ei = torch.randn(H, N, S, generator=g).to(dev)
    na = torch.randn(H, N, A, generator=g).to(dev)
    if use_pcont:
        na = -torch.log(-torch.log(torch.rand(H, N, A, generator=g).clamp(1e-6, 1 - 1e-6))).to(dev)

    def inj(k):
        return dict(eps_prior=[eo, ei] * k, eps_post=[eq] * k, noise_act=[na] * k)

    samples = SamplesFromReplay(obs, act, rew, None, done, None, None)
    with precision("fp32"):
        with noise.injected(**inj(4)):  # 3 warm-up passes + capture: the injected tensors become the graph's noise
            graph = UpdateGraph(m_graph, B, T, horizon=H, use_pcont=use_pcont)
        losses_g = [graph.step(samples).clone() for _ in range(2)]

        def modules(m):
            return [m.observation_encoder, m.observation_decoder, m.reward_model, m.representation, m.transition] + (
                [m.pcont] if use_pcont else [])

        opts = (ClipAdam(get_parameters(modules(m_eager)), lr=6e-4, grad_clip=100.0),
                ClipAdam(list(m_eager.action_decoder.parameters()), lr=8e-5, grad_clip=100.0),
                ClipAdam(list(m_eager.value_model.parameters()), lr=8e-5, grad_clip=100.0))
        losses_e = []
        for _ in range(2):
            m_eager.zero_grad(set_to_none=True)
            with noise.injected(**inj(1)):
                ml, al, vl, _ = dreamer_loss(m_eager, obs[:-1].float() / 255.0 - 0.5, act[1:], rew[1:].unsqueeze(2),
                                             done.unsqueeze(2), horizon=H, use_pcont=use_pcont)
            ml.backward(); al.backward(); vl.backward()
            for o in opts:
                o.step()
            losses_e.append(torch.stack([ml.detach(), al.detach(), vl.detach()]))
    torch.cuda.synchronize()
    for a, b in zip(losses_g, losses_e):
        assert rel_err(a, b) < 1e-5, (a, b)
    assert rel_err(losses_g[0], losses_g[1]) > 1e-7  # the second update saw the first one's parameters
    for (k, pg), (_, pe) in zip(m_graph.named_parameters(), m_eager.named_parameters()):
        # the first Adam steps move every weight by ~lr * sign(g): fp32-atomic reordering of a near-zero gradient shows up
        # as a few 1e-5 of the largest weight
        if k.startswith(("action_decoder", "value_model")):
            assert rel_err(pg, pe) < 1e-4, k
            continue
        # World-model group: every gradient in it flows through the conv decoder / encoder (cuDNN, outside the hot path),
        # whose kernels reduce with atomics and may pick another algorithm under capture, so those gradients carry ~1e-3
        # relative noise (5.9e-4 normwise was seen on transition._cell.weight_ih).  Comparing PARAMETERS after Adam is then
        # ill-conditioned exactly where |g| is at that noise level: Adam's first updates are lr * sign(g), a flipped sign
        # moves the weight by 2 lr per update (a normwise 1.2e-3 was observed on the decoder's Linear with ~0.1 % of its
        # elements flipped).  What must hold: no element further apart than two Adam trajectories can get in two updates
        # (2 x 2 lr), and all but a small fraction of the elements much closer than one update.
        lr = 6e-4
        d = (pg.detach() - pe.detach()).abs()
        assert float(d.max()) <= 4.04 * lr, (k, float(d.max()))
        assert int((d > 0.1 * lr).sum()) <= max(2, d.numel() // 100), (k, int((d > 0.1 * lr).sum()), d.numel())  # biases have 32 elements
