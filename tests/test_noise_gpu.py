"""Production noise (last sentence): Philox4x32-10 streams keyed on (seed, offset, stream, element).

* the generator against the published known-answer vectors and the CPU restatement (oracle/philox_ref.py),
* the N(0,1) / Gumbel(0,1) streams against their moments,
* in-kernel generation (fp32 rollout kernels, forward AND backward) against the same rollout fed the materialised
  tensors of the same (seed, offset): bit for bit,
* offsets advance on the device (also inside a CUDA graph), seeds reproduce.
"""
import ctypes as C

import numpy as np
import pytest
import torch

from oracle import philox_ref as P

pytestmark = pytest.mark.gpu


def _pair(seed, offset, dev):
    return torch.tensor([seed, offset], dtype=torch.int64, device=dev)


def _words(pair, stream, g0, n):
    from dreamer_pytorch_b200 import _lib

    out = torch.zeros(4 * n, dtype=torch.int32, device=pair.device)
    _lib.check(_lib.load().drm_noise_words(pair.data_ptr(), stream, g0, n, out.data_ptr(),
                                           C.c_void_p(torch.cuda.current_stream().cuda_stream)), "drm_noise_words")
    return out.cpu().numpy().view(np.uint32).reshape(n, 4)


def test_philox_known_answers_and_cpu_restatement(cuda):
    for c, k, want in P.KAT:  # the restatement itself (also checked on CPU in test_oracle_golden.py)
        assert P.philox4x32_10(np.array(c, dtype=np.uint32), k).tolist() == list(want)
    # the kernel's counter layout: {group lo, group hi | stream << 24, offset lo, offset hi}, key = seed.  The all-zero
    # known answer is group 0 of stream 0 at offset 0 with seed 0:
    assert _words(_pair(0, 0, cuda), 0, 0, 1)[0].tolist() == list(P.KAT[0][2])
    seed, off = 0x1234567812345678 & 0x7FFFFFFFFFFFFFFF, (7 << 32) | 9
    for stream, g0 in ((0, 0), (2, 12345), (1, (1 << 33) + 5)):
        got = _words(_pair(seed, off, cuda), stream, g0, 257)
        assert np.array_equal(got, P.words(seed, off, stream, np.arange(g0, g0 + 257, dtype=np.uint64)))


@pytest.mark.parametrize("kind", ["normal", "gumbel"])
def test_streams_match_restatement_and_moments(cuda, kind):
    from dreamer_pytorch_b200 import noise

    pair = _pair(99, 3, cuda)
    n = 1_000_003  # not a multiple of 4: ragged tail
    x = noise.fill(cuda, [("eps_prior" if kind == "normal" else "noise_act", (n,), kind)], pair=pair)[0]
    ref = P.stream_values(99, 3, 0 if kind == "normal" else 2, kind, n)
    err = np.abs(x.double().cpu().numpy() - ref)
    assert err.max() < 2e-5 * max(1.0, np.abs(ref).max())  # float32 log / sincos vs float64
    x = x.double()
    m, v = float(x.mean()), float(x.var())
    if kind == "normal":
        assert abs(m) < 5e-3 and abs(v - 1) < 5e-3 and abs(float((x ** 4).mean()) - 3) < 5e-2
        assert abs(float((x[:-1] * x[1:]).mean())) < 5e-3 and abs(float((x[:-2] * x[2:]).mean())) < 5e-3  # Box-Muller partners
        assert float(x.abs().max()) > 4.0
    else:
        assert abs(m - 0.5772156649) < 5e-3 and abs(v - np.pi ** 2 / 6) < 2e-2
    assert torch.isfinite(x).all()
    again = noise.fill(cuda, [("eps_prior" if kind == "normal" else "noise_act", (n,), kind)], pair=pair)[0]
    other = noise.fill(cuda, [("eps_prior" if kind == "normal" else "noise_act", (n,), kind)], pair=_pair(99, 4, cuda))[0]
    assert torch.equal(again.double(), x) and not torch.equal(other.double(), x)


def test_state_advances_on_the_device_and_in_a_graph(cuda):
    from dreamer_pytorch_b200 import noise

    noise.manual_seed(2024)
    st = noise.device_state(cuda)
    assert st.tolist()[:2] == [2024, 0]
    a = noise.normal("eps_prior", (5, 7, 3), cuda)
    b = noise.normal("eps_prior", (5, 7, 3), cuda)
    p = noise.next_pair(cuda)
    assert st.tolist() == [2024, 3, 0, 0] and p.tolist() == [2024, 2]
    assert not torch.equal(a, b)
    noise.manual_seed(2024)
    assert torch.equal(noise.normal("eps_prior", (5, 7, 3), cuda), a)  # reproducible
    # recorded into a graph: every replay draws the next offset without the host
    s = torch.cuda.Stream()
    s.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(s):
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g, stream=s):
            x = noise.normal("noise_act", (64, 6), cuda)
    torch.cuda.current_stream().wait_stream(s)
    before = st.tolist()[1]
    g.replay(); first = x.clone()
    g.replay(); second = x.clone()
    torch.cuda.synchronize()
    assert st.tolist()[1] == before + 2 and not torch.equal(first, second)
    assert torch.equal(first, noise.fill(cuda, [("noise_act", (64, 6), "normal")], pair=_pair(2024, before, cuda))[0])


def _model(dist, dev, pcont=False):
    from dreamer_pytorch_b200.dropin.agent import AtariDreamerModel

    torch.manual_seed(0)
    return AtariDreamerModel(action_shape=(6,), action_dist=dist, use_pcont=pcont).to(dev).train()


@pytest.mark.parametrize("dist", ["tanh_normal", "one_hot", "relaxed_one_hot"])
def test_in_kernel_noise_equals_materialised_noise_policy_rollout(cuda, dist):
    """fp32 imagination: losses and actor / value gradients with the noise generated inside the rollout kernels (forward
    and backward) == the same pass fed the tensors drm_noise_fill writes for that (seed, offset)."""
    from dreamer_pytorch_b200 import noise
    from dreamer_pytorch_b200.algo import imagination_losses
    from dreamer_pytorch_b200.dropin.rnns import RSSMState
    from dreamer_pytorch_b200.precision import precision

    m = _model(dist, cuda)
    N, H, A, S = 333, 7, 6, 30
    g = torch.Generator().manual_seed(1)
    stoch, deter = torch.randn(N, S, generator=g).to(cuda), torch.tanh(torch.randn(N, 200, generator=g)).to(cuda)
    ps = list(m.action_decoder.parameters()) + list(m.value_model.parameters())
    kind = "normal" if dist == "tanh_normal" else "gumbel"

    def run(injected):
        with precision("fp32"):
            if injected is None:
                al, vl, info = imagination_losses(m, RSSMState(stoch, stoch, stoch, deter), H)
            else:
                with noise.injected(noise_act=[injected[0]], eps_prior=[injected[1]]):
                    al, vl, info = imagination_losses(m, RSSMState(stoch, stoch, stoch, deter), H)
        grads = torch.autograd.grad(al + vl, ps)
        return al.detach(), vl.detach(), info.actions, info.imag.stoch, grads

    noise.manual_seed(77)
    offset = noise.device_state(cuda).tolist()[1]
    inside = run(None)
    pair = _pair(77, offset, cuda)
    tensors = noise.fill(cuda, [("noise_act", (H, N, A), kind), ("eps_prior", (H, N, S), "normal")], pair=pair)
    outside = run(tensors)
    assert torch.equal(inside[2], outside[2]) and torch.equal(inside[3], outside[3])
    for a, b in zip(inside[:2], outside[:2]):  # one atomic per CTA in the loss kernel: last-bit differences
        assert abs(float(a) - float(b)) <= 1e-6 * abs(float(b))
    for a, b in zip(inside[4], outside[4]):
        assert float((a - b).abs().max()) <= 1e-5 * float(b.abs().max())  # fp32 atomics in the weight-gradient kernels
    assert float(inside[2].std()) > 0.05  # the noise did something


def test_in_kernel_noise_equals_materialised_noise_observation_rollout(cuda):
    from dreamer_pytorch_b200 import noise
    from dreamer_pytorch_b200.precision import precision

    m = _model("tanh_normal", cuda)
    T, B, S = 9, 5, 30
    g = torch.Generator().manual_seed(2)
    embed = torch.randn(T, B, 1024, generator=g).to(cuda)
    act = (torch.rand(T, B, 6, generator=g) * 2 - 1).to(cuda)
    ps = list(m.representation.parameters())

    def run(injected):
        prev = m.representation.initial_state(B, device=cuda)
        with precision("fp32"):
            if injected is None:
                prior, post = m.rollout.rollout_representation(T, embed, act, prev)
            else:
                with noise.injected(eps_prior=[injected[0]], eps_post=[injected[1]]):
                    prior, post = m.rollout.rollout_representation(T, embed, act, prev)
        loss = post.stoch.square().mean() + prior.stoch.sum() * 1e-3 + post.std.mean() + prior.std.mean()
        return prior.stoch.detach(), post.stoch.detach(), torch.autograd.grad(loss, ps, allow_unused=True)

    noise.manual_seed(5)
    offset = noise.device_state(cuda).tolist()[1]
    inside = run(None)
    tensors = noise.fill(cuda, [("eps_prior", (T, B, S), "normal"), ("eps_post", (T, B, S), "normal")],
                         pair=_pair(5, offset, cuda))
    outside = run(tensors)
    assert torch.equal(inside[0], outside[0]) and torch.equal(inside[1], outside[1])
    for a, b in zip(inside[2], outside[2]):
        if a is not None:
            assert float((a - b).abs().max()) <= 1e-5 * float(b.abs().max())
    assert not torch.equal(tensors[0], tensors[1])  # two streams of one pair


def test_bf16_rollout_materialises_pairs_and_stays_reproducible(cuda):
    """Tensor-core path: pairs handed to the bf16 policy rollout are materialised with the same values; the fused loss
    section draws both blocks with one launch; the same seed gives the same update."""
    from dreamer_pytorch_b200 import noise
    from dreamer_pytorch_b200.algo import imagination_losses
    from dreamer_pytorch_b200.dropin.rnns import RSSMState
    from dreamer_pytorch_b200.precision import precision

    m = _model("tanh_normal", cuda)
    N, H = 300, 5
    g = torch.Generator().manual_seed(3)
    stoch, deter = torch.randn(N, 30, generator=g).to(cuda), torch.tanh(torch.randn(N, 200, generator=g)).to(cuda)
    st = RSSMState(stoch, stoch, stoch, deter)
    with precision("bf16"), torch.no_grad():
        noise.manual_seed(11)
        off = noise.device_state(cuda).tolist()[1]
        imag, act = m.rollout.rollout_policy(H, m.policy, st)  # module path: one pair -> materialised for the TC kernel
        na, ep = noise.fill(cuda, [("noise_act", (H, N, 6), "normal"), ("eps_prior", (H, N, 30), "normal")],
                            pair=_pair(11, off, cuda))
        with noise.injected(noise_act=[na], eps_prior=[ep]):
            imag2, act2 = m.rollout.rollout_policy(H, m.policy, st)
        assert torch.equal(act, act2) and torch.equal(imag.deter, imag2.deter)
    with precision("bf16"):
        res = []
        for _ in range(2):
            noise.manual_seed(12)
            al, vl, info = imagination_losses(m, st, H)
            res.append((al.detach().clone(), info.actions.clone()))
        assert torch.equal(res[0][0], res[1][0]) and torch.equal(res[0][1], res[1][1])
        al3, _, info3 = imagination_losses(m, st, H)  # next offset: different noise
        assert not torch.equal(info3.actions, res[0][1])
