"""Host-side logic of the drop-in modules that needs no GPU: constructor contracts, state_dict
compatibility, FreezeParameters, the rlpyt shims, noise injection and the no-CPU-fallback rule."""
import os
import sys

import pytest
import torch


def test_state_dict_keys_and_aliases():
    from dreamer_pytorch_b200.dropin.agent import AtariDreamerModel
    from oracle import dreamer_oracle as O

    m = AtariDreamerModel(action_shape=(6,), action_dist="tanh_normal")
    sd = m.state_dict()
    assert len(sd) == 90  # SURVEY.md section 5, checkpoint row
    full = O.expand_aliases(O.make_weights(O.weight_shapes(6, conv=True), 0))
    assert set(full) == set(sd)
    assert all(tuple(full[k].shape) == tuple(sd[k].shape) for k in sd)
    m.load_state_dict(full)
    # the transition is one shared object
    assert m.representation._transition_model is m.transition is m.rollout.transition_model
    m2 = AtariDreamerModel(action_shape=6, action_dist="one_hot", use_pcont=True, output_size=6)  # Q8: int + stray kwarg
    assert any(k.startswith("pcont.model.") for k in m2.state_dict())


def test_constructor_errors():
    from dreamer_pytorch_b200.dropin.action import ActionDecoder
    from dreamer_pytorch_b200.dropin.dense import DenseModel

    with pytest.raises(NotImplementedError):
        ActionDecoder(10, 20, 20, 2, "not_implemented_dist")  # action.py:47-48
    d = DenseModel(20, (1,), 2, 16, dist="weird")
    assert d._dist == "weird"  # dense.py:50 raises at forward time, not in the constructor


def test_no_cpu_fallback():
    from dreamer_pytorch_b200.dropin.dense import DenseModel
    from dreamer_pytorch_b200.dropin.rnns import RSSMTransition

    tr = RSSMTransition(3, 8, 16, 16)
    with pytest.raises(RuntimeError):
        tr(torch.randn(2, 3), tr.initial_state(2))
    with pytest.raises(RuntimeError):
        DenseModel(8, (1,), 2, 16)(torch.randn(2, 8))


def test_product_never_imports_the_oracle():
    root = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "dreamer_pytorch_b200")
    for dirpath, _, files in os.walk(root):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                src = open(os.path.join(dirpath, f)).read()
                assert "oracle" not in src.replace("test infrastructure", ""), f"{f} mentions the oracle"


def test_freeze_parameters():  # reference tests/test_utils.py:7-48
    from dreamer_pytorch_b200.dropin.module import FreezeParameters, get_parameters

    a, b = torch.nn.Linear(4, 4), torch.nn.Linear(4, 4)
    b.bias.requires_grad = False
    with FreezeParameters([a, b]):
        assert not any(p.requires_grad for p in get_parameters([a, b]))
        assert b(a(torch.randn(2, 4))).grad_fn is None
    assert a.weight.requires_grad and a.bias.requires_grad and b.weight.requires_grad
    assert not b.bias.requires_grad  # the original per-parameter state is restored


def test_rssm_state_behaves_like_namedarraytuple():
    from dreamer_pytorch_b200.dropin._compat import buffer_method
    from dreamer_pytorch_b200.dropin.rnns import RSSMState, get_dist, get_feat, stack_states

    s = RSSMState(torch.zeros(5, 3, 2), torch.ones(5, 3, 2), torch.zeros(5, 3, 2), torch.zeros(5, 3, 4))
    assert s[-1].deter.shape == (3, 4) and s[:-1, :].stoch.shape == (4, 3, 2)
    flat = buffer_method(s, "reshape", 15, -1)
    assert flat.deter.shape == (15, 4) and isinstance(flat, RSSMState)
    assert get_feat(s).shape == (5, 3, 6) and get_dist(s).event_shape == (2,)
    st = stack_states([s[0], s[1]], dim=0)
    assert st.mean.shape == (2, 3, 2)
    assert buffer_method(s, "detach").std.requires_grad is False


def test_noise_injection_order_and_leftovers():
    from dreamer_pytorch_b200 import noise

    a, b = torch.zeros(2, 3), torch.ones(2, 3)
    with noise.injected(eps_prior=[a, b]):
        assert torch.equal(noise.normal("eps_prior", (2, 3), "cpu"), a)
        assert torch.equal(noise.normal("eps_prior", (2, 3), "cpu"), b)
        with pytest.raises(RuntimeError, match="no CPU fallback"):  # not injected: a fresh draw is a CUDA kernel
            noise.normal("eps_post", (4,), "cpu")
    with pytest.raises(AssertionError):
        with noise.injected(eps_prior=[a]):
            pass
    with pytest.raises(RuntimeError):
        with noise.injected(eps_prior=[a]):
            noise.normal("eps_prior", (5,), "cpu")


def test_sample_dist_forwards_attributes():  # reference tests/test_distribution.py:7-21
    from dreamer_pytorch_b200.dropin.distribution import SampleDist

    d = torch.distributions.Independent(torch.distributions.Normal(torch.zeros(4, 3), torch.ones(4, 3)), 1)
    sd = SampleDist(d)
    assert sd.event_shape == (3,) and sd.batch_shape == (4,) and sd.name == "SampleDist"
    assert sd.mean().shape == (4, 3) and sd.mode().shape == (4, 3) and sd.entropy().shape == (4,)


def test_install_aliases_reference_paths():
    import dreamer_pytorch_b200.dropin as dropin

    saved = {k: v for k, v in sys.modules.items() if k == "dreamer" or k.startswith("dreamer.")}
    try:
        dropin.install()
        from dreamer.models.rnns import RSSMRollout  # noqa
        from dreamer.models.agent import AgentModel  # noqa
        from dreamer.utils.module import FreezeParameters  # noqa

        assert RSSMRollout is dropin.rnns.RSSMRollout and AgentModel is dropin.agent.AgentModel
    finally:
        for k in [k for k in sys.modules if k == "dreamer" or k.startswith("dreamer.")]:
            del sys.modules[k]
        sys.modules.update(saved)


def test_torch_ops_are_registered_without_a_gpu():
    """torch.ops.dreamer_b200.* (SURVEY 8(b) torch op layer): registration and schemas need no device."""
    import dreamer_pytorch_b200.ops as ops

    assert set(ops.OP_NAMES) == {"transition_step", "transition_rollout", "representation_rollout", "policy_rollout", "mlp",
                                 "kl_entropy", "heads_returns_loss", "image_logprob"}
    schema = str(torch.ops.dreamer_b200.policy_rollout.default._schema)
    assert "Tensor[] rssm" in schema and "str act_dist" in schema and schema.endswith("-> Tensor[]")
    with pytest.raises(RuntimeError):  # CPU tensors: no fallback behind the ops either
        torch.ops.dreamer_b200.mlp(torch.randn(2, 8), None, [torch.randn(1, 8), torch.randn(1)])


def test_gradient_buckets_come_from_the_provider():
    """functional.flat_like hands out views of ONE flat buffer per group; with a bucket provider (parallel.PeerBucket on a
    multi-GPU box) the buffer is the provider's, e.g. peer-mapped memory."""
    from dreamer_pytorch_b200 import functional as F

    ws = [torch.zeros(5, 3), torch.zeros(5), torch.zeros(2, 5), torch.zeros(2)]
    own = F.flat_like(ws)
    assert all(g._base is own[0]._base for g in own) and [tuple(g.shape) for g in own] == [tuple(w.shape) for w in ws]
    assert all(g.data_ptr() % 16 == 0 for g in own)  # 16-byte aligned pieces (vector reductions, one-shot all-reduce)
    pool = torch.zeros(64)
    seen = []

    def provide(tensors):
        seen.append([tuple(t.shape) for t in tensors])
        return pool[:sum((t.numel() + 3) // 4 * 4 for t in tensors)]

    with F.bucket_provider(provide):
        got = F.flat_like(ws)
    got[0].fill_(1.0)
    assert float(pool[:15].sum()) == 15.0 and seen == [[(5, 3), (5,), (2, 5), (2,)]]
    assert F.flat_like(ws)[0]._base is not pool  # the provider is gone with the context
