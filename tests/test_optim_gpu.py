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


def test_clip_adam_rejects_cpu_and_duplicates():
    from dreamer_pytorch_b200.optim import ClipAdam

    with pytest.raises(RuntimeError):
        ClipAdam([torch.nn.Parameter(torch.zeros(3))])
    p = torch.nn.Parameter(torch.zeros(3, device="cuda:0"))
    p.grad = torch.ones_like(p)
    with pytest.raises(NotImplementedError):
        ClipAdam([p, p], grad_clip=1.0).step()
