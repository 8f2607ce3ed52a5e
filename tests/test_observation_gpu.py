"""the fused image log-prob kernel against torch.distributions, and the channels-last bf16 conv
encoder / decoder against their fp32 selves (same parameters)."""
import pytest
import torch
import torch.distributions as td

from _cases import FP32_TOL, rel_err

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("rows,shape", [(7, (3, 64, 64)), (1, (1, 4, 4)), (2500, (3, 64, 64)), (33, (3, 20, 12))])
@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_image_logprob_matches_torch(cuda, rows, shape, dtype):
    from dreamer_pytorch_b200.dropin.observation import ImageNormal

    g = torch.Generator().manual_seed(3)
    mean = (torch.randn(rows, *shape, generator=g) * 0.3).to(cuda).to(dtype).requires_grad_(True)
    obs = (torch.rand(rows, *shape, generator=g) - 0.5).to(cuda)
    gy = torch.randn(rows, generator=g).to(cuda)
    lp = ImageNormal(mean, 3).log_prob(obs)
    (gm,) = torch.autograd.grad(lp, mean, gy)
    ref_mean = mean.detach().double().requires_grad_(True)
    ref = td.Independent(td.Normal(ref_mean, 1), 3).log_prob(obs.double())
    (rg,) = torch.autograd.grad(ref, ref_mean, gy.double())
    assert lp.shape == ref.shape and gm.dtype == dtype
    assert rel_err(lp, ref) < FP32_TOL
    assert rel_err(gm.float(), rg) < (FP32_TOL if dtype == torch.float32 else 5e-3)  # bf16 output rounding 2^-9


def test_image_logprob_keeps_leading_dims_and_mean(cuda):
    from dreamer_pytorch_b200.dropin.observation import ImageNormal

    mean = torch.randn(5, 4, 3, 8, 8, device=cuda)
    d = ImageNormal(mean.bfloat16(), 3)
    assert d.mean.dtype == torch.float32 and d.batch_shape == (5, 4) and d.event_shape == (3, 8, 8)
    assert d.log_prob(torch.zeros_like(mean)).shape == (5, 4)


def test_bf16_channels_last_conv_stack_tracks_fp32(cuda):
    from dreamer_pytorch_b200.dropin.observation import ObservationDecoder, ObservationEncoder
    from dreamer_pytorch_b200.precision import conv_mode, precision

    torch.manual_seed(0)
    enc, dec = ObservationEncoder().to(cuda), ObservationDecoder(embed_size=230).to(cuda)
    keys = list(enc.state_dict()) + list(dec.state_dict())
    obs = torch.rand(6, 5, 3, 64, 64, device=cuda) - 0.5
    feat = torch.randn(6, 5, 230, device=cuda)
    with precision("fp32"):
        e32 = enc(obs)
        lp32 = dec(feat).log_prob(obs)
    with precision("bf16"), conv_mode("channels_last"):
        ecl = enc(obs)
        lpcl = dec(feat).log_prob(obs)
    assert rel_err(ecl, e32) < 1e-4 and rel_err(lpcl, lp32) < 1e-4  # fp32 channels-last (TF32 off in the test fixture)
    with precision("bf16"), conv_mode("bf16"):
        e16 = enc(obs)
        d16 = dec(feat)
        lp16 = d16.log_prob(obs)
        loss = -lp16.mean() + e16.square().mean()
        loss.backward()
    assert e16.dtype == torch.float32 and e16.shape == (6, 5, 1024) and lp16.shape == (6, 5)
    assert rel_err(e16, e32) < 1e-2 and rel_err(lp16, lp32) < 1e-2
    assert keys == list(enc.state_dict()) + list(dec.state_dict())  # channels-last conversion keeps the state_dict
    for p in list(enc.parameters()) + list(dec.parameters()):
        assert p.grad is not None and p.grad.dtype == torch.float32 and torch.isfinite(p.grad).all()
        assert p.shape == p.grad.shape


def test_fused_bias_relu_conv_stack_matches_torch_modules(cuda):
    """conv_mode("channels_last"): convolutions without bias + this library's in-place bias/ReLU pass (forward) and
    ReLU-mask + bias-gradient pass (backward) against the plain torch modules on the same parameters (fp32 NCHW)."""
    from dreamer_pytorch_b200 import _lib
    from dreamer_pytorch_b200.dropin.observation import ObservationDecoder, ObservationEncoder
    from dreamer_pytorch_b200.precision import conv_mode, precision

    torch.manual_seed(1)
    enc, dec = ObservationEncoder().to(cuda), ObservationDecoder(embed_size=230).to(cuda)
    params = list(enc.parameters()) + list(dec.parameters())
    obs = torch.rand(3, 7, 3, 64, 64, device=cuda) - 0.5
    feat = torch.randn(3, 7, 230, device=cuda)

    def run():
        e = enc(obs)
        lp = dec(feat).log_prob(obs)
        loss = -lp.mean() + e.square().mean()
        return e.detach(), lp.detach(), torch.autograd.grad(loss, params)

    with precision("fp32"), conv_mode("nchw"):
        e0, lp0, g0 = run()
    n0 = _lib.load().drm_launch_count()
    with precision("fp32"), conv_mode("channels_last"):
        e1, lp1, g1 = run()
    assert _lib.load().drm_launch_count() - n0 >= 2 * 7 + 2  # 7 fused layers fwd + bwd, image log-prob fwd + bwd
    assert rel_err(e1, e0) < 1e-4 and rel_err(lp1, lp0) < 1e-4
    for p, a, b in zip(params, g1, g0):
        assert a.shape == p.shape and rel_err(a, b) < 2e-3, tuple(p.shape)  # cuDNN picks other algorithms per layout
    with torch.no_grad(), precision("fp32"), conv_mode("channels_last"):
        assert rel_err(enc(obs), e0) < 1e-4  # inference path (no autograd graph): same kernels
