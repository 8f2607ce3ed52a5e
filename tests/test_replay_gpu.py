"""the device-resident replay against rlpyt's UniformSequenceReplayBuffer (the reference's buffer,
imported unmodified from oracle/_ref or /root/reference) on the same appended data and the same np.random stream:
bit-exact batches, including ring wrap-around and the invalid zone around the cursor."""
import numpy as np
import pytest
import torch

from oracle import ref_harness as RH

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("size,chunks", [(600, 3), (600, 9)])  # 9 chunks of 100 steps x 2 envs overfill a 300-step ring
def test_device_replay_matches_rlpyt_buffer(cuda, size, chunks):
    if not RH.reference_available():
        pytest.skip("reference checkout not available")
    RH.import_reference()
    from dreamer.algos.replay import SamplesToBuffer
    from rlpyt.replays.sequence.uniform import UniformSequenceReplayBuffer

    from dreamer_pytorch_b200.replay import DeviceSequenceReplay

    B, A, img = 2, 3, (3, 8, 8)
    example = SamplesToBuffer(observation=np.zeros(img, np.uint8), action=np.zeros(A, np.float32),
                              reward=np.float32(0), done=np.bool_(False))
    ref = UniformSequenceReplayBuffer(example=example, size=size, B=B, rnn_state_interval=0, discount=0.99,
                                      n_step_return=1)
    ours = DeviceSequenceReplay(example, size, B, device=cuda, discount=0.99, n_step_return=1)
    assert ours.T == ref.T
    rng = np.random.RandomState(0)
    for _ in range(chunks):
        Tn = 100
        s = SamplesToBuffer(observation=rng.randint(0, 256, (Tn, B) + img).astype(np.uint8),
                            action=rng.randn(Tn, B, A).astype(np.float32), reward=rng.randn(Tn, B).astype(np.float32),
                            done=rng.rand(Tn, B) < 0.05)
        ref.append_samples(s)
        ours.append_samples(SamplesToBuffer(*(torch.from_numpy(x) for x in s)))
    assert ours.t == ref.t and ours._buffer_full == ref._buffer_full
    for seed in range(5):
        np.random.seed(seed)
        a = ref.sample_batch(7, 20)
        np.random.seed(seed)
        b = ours.sample_batch(7, 20)
        for name in ("all_observation", "all_action", "all_reward", "return_", "done", "done_n"):
            x, y = getattr(a, name), getattr(b, name)
            assert tuple(x.shape) == tuple(y.shape), name
            assert y.is_cuda and x.dtype == y.dtype, name
            assert torch.equal(x, y.cpu()), name
        assert b.init_rnn_state is None


def test_device_replay_explicit_wrap_indices(cuda):
    """Start times at both ends of the ring: t0 = 0 makes the prev-action sequence start at index -1 (wrap at the beginning),
    t0 = T - 3 wraps at the end (rlpyt/utils/misc.py:47-55)."""
    from collections import namedtuple

    from dreamer_pytorch_b200.replay import DeviceSequenceReplay

    Ex = namedtuple("Ex", ["observation", "action", "reward", "done"])
    ex = Ex(np.zeros((1, 5), np.uint8), np.zeros(2, np.float32), np.float32(0), np.bool_(False))
    rep = DeviceSequenceReplay(ex, 40, 2, device=cuda)
    T = rep.T
    obs = torch.arange(T * 2 * 5, dtype=torch.int64).reshape(T, 2, 1, 5).remainder(251).to(torch.uint8)
    act = torch.arange(T * 2 * 2, dtype=torch.float32).reshape(T, 2, 2)
    rew = torch.arange(T * 2, dtype=torch.float32).reshape(T, 2)
    rep.append_samples(Ex(obs, act, rew, torch.zeros(T, 2, dtype=torch.bool)))
    out = rep.extract_batch(np.array([0, T - 3]), np.array([1, 0]), 5)
    for i, (t0, b) in enumerate([(0, 1), (T - 3, 0)]):
        for j in range(6):
            assert torch.equal(out.all_observation[j, i].cpu(), obs[(t0 + j) % T, b])
            assert torch.equal(out.all_action[j, i].cpu(), act[(t0 - 1 + j) % T, b])
            assert float(out.all_reward[j, i]) == float(rew[(t0 - 1 + j) % T, b])
