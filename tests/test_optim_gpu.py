"""ClipAdam (clip_grad_norm_ + Adam.step in one C-ABI call) against torch's own implementation."""
import pytest
import torch

pytestmark = pytest.mark.gpu


def rel_err(a, b):
    a, b = a.detach().double().cpu(), b.detach().double().cpu()
    return float((a - b).abs().max() / b.abs().max().clamp_min(1e-30))


@pytest.mark.parametrize("grad_clip,scale", [(100.0, 1.0), (100.0, 400.0), (0.5, 1.0)])
def test_clip_adam_matches_torch(grad_clip, scale):
    from oracle import dreamer_oracle as O
    from dreamer_pytorch_b200.optim import ClipAdam

    dev = torch.device("cuda:0")
    g = torch.Generator().manual_seed(7)
    shapes = [(200, 230), (200,), (200, 200), (200,), (200, 200), (200,), (12, 200), (12,)]   # the actor group
    params0 = [torch.randn(s, generator=g) * 0.1 for s in shapes]
    steps = 4
    grads = [[torch.randn(s, generator=g) * scale for s in shapes] for _ in range(steps)]   # scale 400: the clip engages
    ref_p, ref_norms, ref_m, ref_v = O.clip_adam_steps(params0, grads, lr=8e-5, grad_clip=grad_clip)

    ps = [torch.nn.Parameter(p.clone().to(dev)) for p in params0]
    opt = ClipAdam(ps, lr=8e-5, grad_clip=grad_clip)
    norms = []
    for k in range(steps):
        for p, gk in zip(ps, grads[k]):
            p.grad = gk.clone().to(dev)
        norms.append(opt.step().clone())
    torch.cuda.synchronize()
    for a, b in zip(norms, ref_norms):
        assert rel_err(a, b) < 1e-5
    for p, r in zip(ps, ref_p):
        assert rel_err(p, r) < 1e-5
    for p, m, v in zip(ps, ref_m, ref_v):
        assert rel_err(opt.state[p]["exp_avg"], m) < 1e-5 and rel_err(opt.state[p]["exp_avg_sq"], v) < 1e-5
    # gradients are scaled in place, as clip_grad_norm_ does
    coef = min(1.0, grad_clip / (float(ref_norms[-1]) + 1e-6))
    assert rel_err(ps[0].grad, grads[-1][0] * coef) < 1e-5


def test_clip_adam_rejects_cpu():
    from dreamer_pytorch_b200.optim import ClipAdam

    with pytest.raises(RuntimeError):
        ClipAdam([torch.nn.Parameter(torch.zeros(3))])


@pytest.mark.parametrize("grad_clip", [100.0, 0.05])
def test_world_model_group_with_duplicated_parameters_matches_torch(grad_clip):
    """The reference's world-model parameter list (dreamer_algo.py:112-118 via get_parameters) holds the transition
    parameters twice.  Pin ClipAdam against torch itself -- clip_grad_norm_ + torch.optim.Adam over the SAME list with
    duplicates, on the CPU, where torch's semantics are sequential and deterministic -- over three updates."""
    import warnings

    from dreamer_pytorch_b200.dropin.agent import AtariDreamerModel
    from dreamer_pytorch_b200.dropin.module import get_parameters
    from dreamer_pytorch_b200.optim import ClipAdam

    dev = torch.device("cuda:0")
    torch.manual_seed(5)
    cpu = AtariDreamerModel(action_shape=(6,), action_dist="tanh_normal")
    gpu = AtariDreamerModel(action_shape=(6,), action_dist="tanh_normal")
    gpu.load_state_dict(cpu.state_dict())
    gpu = gpu.to(dev)

    def modules(m):  # dreamer_algo.py:112-118
        return [m.observation_encoder, m.observation_decoder, m.reward_model, m.representation, m.transition]

    listed_cpu, listed_gpu = get_parameters(modules(cpu)), get_parameters(modules(gpu))
    assert len(listed_cpu) == len(listed_gpu) > len({id(p) for p in listed_cpu})  # duplicates are really there
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        ref = torch.optim.Adam(listed_cpu, lr=6e-4, foreach=False)
    opt = ClipAdam(listed_gpu, lr=6e-4, grad_clip=grad_clip)
    uniq_cpu = list({id(p): p for p in listed_cpu}.values())
    uniq_gpu = list({id(p): p for p in listed_gpu}.values())
    g = torch.Generator().manual_seed(6)
    for _ in range(3):
        for pc, pg in zip(uniq_cpu, uniq_gpu):
            gr = torch.randn(pc.shape, generator=g) * 0.01
            pc.grad, pg.grad = gr.clone(), gr.clone().to(dev)
        ref_norm = torch.nn.utils.clip_grad_norm_(listed_cpu, grad_clip, foreach=False)
        ref.step()
        norm = opt.step()
        # torch accumulates the norm of 5.3 M elements in fp32 on the CPU (4e-5 off the exact value here); the kernel sums
        # in double.  With the clip engaged that difference enters the gradients twice for the duplicated tensors.
        assert rel_err(norm, ref_norm) < 1e-4
    for pc, pg in zip(uniq_cpu, uniq_gpu):
        assert rel_err(pg, pc) < 1e-4
        assert rel_err(pg.grad, pc.grad) < 2e-4  # scaled in place, twice for the duplicated tensors
        assert rel_err(opt.state[pg]["exp_avg"], ref.state[pc]["exp_avg"]) < 2e-4
        assert rel_err(opt.state[pg]["exp_avg_sq"], ref.state[pc]["exp_avg_sq"]) < 4e-4
    # state_dict in torch.optim.Adam's layout: loads into torch and back
    sd, ref_sd = opt.state_dict(), ref.state_dict()
    assert sd["param_groups"][0]["params"] == ref_sd["param_groups"][0]["params"]
    assert set(sd["state"]) == set(ref_sd["state"])
    for k in ref_sd["state"]:
        assert float(sd["state"][k]["step"]) == float(ref_sd["state"][k]["step"])
    opt2 = ClipAdam(listed_gpu, lr=1.0)
    opt2.load_state_dict(ref_sd)
    assert opt2.step_count == 3 and abs(opt2.lr - 6e-4) < 1e-12
    assert rel_err(opt2.state[uniq_gpu[0]]["exp_avg"], ref.state[uniq_cpu[0]]["exp_avg"]) < 1e-6


def test_whole_update_in_one_graph_matches_eager_updates():
    """ImaginationGraph with recorded optimizers (one replay = losses, backward, clip + Adam for actor and value) against
    the same fused losses stepped eagerly with torch's clip_grad_norm_ + Adam, on identical noise, over three updates."""
    import copy

    from dreamer_pytorch_b200 import noise
    from dreamer_pytorch_b200.algo import imagination_losses
    from dreamer_pytorch_b200.dropin.agent import AtariDreamerModel
    from dreamer_pytorch_b200.dropin.rnns import RSSMState
    from dreamer_pytorch_b200.graph import ImaginationGraph
    from dreamer_pytorch_b200.optim import ClipAdam
    from dreamer_pytorch_b200.precision import precision

    dev = torch.device("cuda:0")
    torch.manual_seed(0)
    A, N, H, updates, warm = 6, 300, 5, 3, 2
    model = AtariDreamerModel(action_shape=(A,), action_dist="tanh_normal").to(dev).train()
    ref_model = copy.deepcopy(model)
    g = torch.Generator().manual_seed(1)
    stoch = torch.randn(N, 30, generator=g).to(dev)
    deter = torch.tanh(torch.randn(N, 200, generator=g)).to(dev)
    eps = [torch.randn(H, N, 30, generator=g).to(dev) for _ in range(updates)]
    nza = [torch.randn(H, N, A, generator=g).to(dev) for _ in range(updates)]
    lr, clip = 8e-5, 100.0
    with precision("bf16"):
        # eager reference: fused losses + torch optimizers
        groups = [list(ref_model.action_decoder.parameters()), list(ref_model.value_model.parameters())]
        opts = [torch.optim.Adam(ps, lr=lr) for ps in groups]
        ref_norms = []
        with noise.injected(eps_prior=eps, noise_act=nza):
            for _ in range(updates):
                for o in opts:
                    o.zero_grad()
                al, vl, _ = imagination_losses(ref_model, RSSMState(stoch, stoch, stoch, deter), H)
                al.backward(); vl.backward()
                ref_norms.append([torch.nn.utils.clip_grad_norm_(ps, clip).clone() for ps in groups])
                for o in opts:
                    o.step()
        del al, vl
        # one graph per update
        a_opt = ClipAdam(model.action_decoder.parameters(), lr=lr, grad_clip=clip, capturable=True)
        v_opt = ClipAdam(model.value_model.parameters(), lr=lr, grad_clip=clip, capturable=True)
        eps_buf, nza_buf = eps[0].clone(), nza[0].clone()  # the graph records these buffers: refilled before every replay
        with noise.injected(eps_prior=[eps_buf] * (warm + 1), noise_act=[nza_buf] * (warm + 1)):
            graph = ImaginationGraph(model, N, H, warmup=warm, actor_opt=a_opt, value_opt=v_opt)
        assert a_opt.step_count == 0 and v_opt.step_count == 0      # warm-up passes left no trace
        norms = []
        for k in range(updates):
            eps_buf.copy_(eps[k]); nza_buf.copy_(nza[k])
            graph.step(stoch, deter)
            norms.append([n.clone() for n in graph.grad_norms])
    torch.cuda.synchronize()
    assert a_opt.step_count == updates and v_opt.step_count == updates
    for p, r in zip(model.action_decoder.parameters(), ref_model.action_decoder.parameters()):
        assert rel_err(p, r) < 1e-4
    for p, r in zip(model.value_model.parameters(), ref_model.value_model.parameters()):
        assert rel_err(p, r) < 1e-4
    for a, b in zip(norms, ref_norms):
        assert rel_err(a[0], b[0]) < 1e-3 and rel_err(a[1], b[1]) < 1e-3
