"""Device-resident sequence replay.

The reference samples every update on the host: ``UniformSequenceReplayBuffer.sample_batch`` gathers ``[T+1, B]``
sequences with numpy (rlpyt/replays/sequence/uniform.py:15-39, n_step.py:68-106) and ``buffer_to`` then copies ~31 MB
synchronously to the GPU (dreamer_algo.py:166-169).  Here the ring buffer itself lives in HBM (1 M frames of 3x64x64
uint8 are 12 GB of the 180), new environment steps are appended with small asynchronous copies, and a batch is ONE
gather launch per field (``drm_replay_gather``) writing the ``[T+1, B, ...]`` tensors straight into device memory.

Same interface and same index arithmetic as the reference buffer (``append_samples``, ``sample_batch(batch_B,
batch_T)``, the cursor / wrap / invalid-zone rules of ``sample_idxs``; indices come from ``np.random`` like there, so a
seeded run samples the same sequences): it drops into ``Dreamer.initialize_replay_buffer`` (dreamer/algos/replay.py).
``n_step_return`` must be 1 (Dreamer's value: return_ = reward, done_n = done); RNN states are not stored
(``rnn_state_interval=0``, as replay.py:24 sets it).
"""
from __future__ import annotations

import ctypes as C
import math
from collections import namedtuple

import numpy as np
import torch

from . import _lib

SamplesFromReplay = namedtuple("SamplesFromReplay", ["all_observation", "all_action", "all_reward", "return_", "done",
                                                     "done_n", "init_rnn_state"])


class DeviceSequenceReplay:
    def __init__(self, example, size: int, B: int, device, rnn_state_interval: int = 0, discount: float = 1.0,
                 n_step_return: int = 1, batch_T=None):
        """example: namedtuple-like with fields observation / action / reward / done holding ONE entry each (no leading
        dims), as ``initialize_replay_buffer`` builds it."""
        if n_step_return != 1:
            raise NotImplementedError("DeviceSequenceReplay implements n_step_return=1 (Dreamer's setting)")
        if rnn_state_interval != 0:
            raise NotImplementedError("RNN states are not stored (rnn_state_interval=0, dreamer/algos/replay.py:24)")
        self.device = torch.device(device)
        if self.device.type != "cuda":
            raise RuntimeError("DeviceSequenceReplay keeps the ring in HBM: it needs a CUDA device (no CPU fallback)")
        self.T = math.ceil(size / B)  # rlpyt/replays/n_step.py: T = ceil(size / B)
        self.B = int(B)
        self.size = self.T * self.B
        self.batch_T = batch_T
        self.discount, self.n_step_return = discount, n_step_return
        self.t = 0
        self._buffer_full = False
        self.off_backward = n_step_return  # rlpyt/replays/n_step.py:59-60
        self.off_forward = 1
        self.rnn_state_interval = 0
        self.fields = {}
        for name in ("observation", "action", "reward", "done"):
            ex = torch.as_tensor(np.asarray(getattr(example, name)))
            self.fields[name] = torch.zeros(self.T, self.B, *ex.shape, dtype=ex.dtype, device=self.device)

    # ------------------------------------------------------------------ append (rlpyt/replays/n_step.py:62-79)
    def append_samples(self, samples):
        """samples: fields observation / action / reward / done with leading dims [T_new, B] (host or device tensors
        or numpy arrays).  Returns (T_new, idxs) like the reference."""
        first = torch.as_tensor(np.asarray(samples.reward) if not torch.is_tensor(samples.reward) else samples.reward)
        Tn, B = first.shape[:2]
        assert B == self.B
        t = self.t
        idxs = np.arange(t, t + Tn) % self.T if t + Tn > self.T else slice(t, t + Tn)
        for name, buf in self.fields.items():
            src = getattr(samples, name)
            src = src if torch.is_tensor(src) else torch.as_tensor(np.asarray(src))
            src = src.to(dtype=buf.dtype)
            if t + Tn > self.T:
                m = self.T - t
                buf[t:].copy_(src[:m], non_blocking=True)
                buf[:Tn - m].copy_(src[m:], non_blocking=True)
            else:
                buf[t:t + Tn].copy_(src, non_blocking=True)
        if not self._buffer_full and t + Tn >= self.T:
            self._buffer_full = True
        self.t = (t + Tn) % self.T
        return Tn, idxs

    # ------------------------------------------------------------------ sampling (rlpyt/replays/sequence/uniform.py:15-39)
    def set_batch_T(self, batch_T):
        self.batch_T = batch_T

    def sample_idxs(self, batch_B, batch_T):
        t, b, f = self.t, self.off_backward + batch_T, self.off_forward
        high = self.T - b - f if self._buffer_full else t - b - f
        T_idxs = np.random.randint(low=0, high=high, size=(batch_B,))
        T_idxs[T_idxs >= t - b] += min(t, b) + f
        B_idxs = np.random.randint(low=0, high=self.B, size=(batch_B,))
        return T_idxs, B_idxs

    def sample_batch(self, batch_B, batch_T=None):
        batch_T = self.batch_T if batch_T is None else batch_T
        T_idxs, B_idxs = self.sample_idxs(batch_B, batch_T)
        return self.extract_batch(T_idxs, B_idxs, batch_T)

    # ------------------------------------------------------------------ gather (rlpyt/replays/sequence/n_step.py:68-106)
    def _gather(self, buf, t0, b, T_seq, t_offset):
        lib = _lib.load()
        _lib.ensure_init()
        batch = t0.numel()
        out = torch.empty(T_seq, batch, *buf.shape[2:], dtype=buf.dtype, device=self.device)
        c = _lib.drm_replay_gather_call()
        c.T_cap, c.B_env, c.T_seq, c.batch, c.t_offset = self.T, self.B, int(T_seq), int(batch), int(t_offset)
        c.row_bytes = int(buf[0, 0].numel() * buf.element_size())
        c.src, c.dst, c.t0, c.b = buf.data_ptr(), out.data_ptr(), t0.data_ptr(), b.data_ptr()
        stream = C.c_void_p(torch.cuda.current_stream(self.device).cuda_stream)
        _lib.check(lib.drm_replay_gather(C.byref(c), stream), "drm_replay_gather")
        return out

    def extract_batch(self, T_idxs, B_idxs, T):
        t0 = torch.as_tensor(np.asarray(T_idxs, dtype=np.int64)).to(self.device, non_blocking=True)
        b = torch.as_tensor(np.asarray(B_idxs, dtype=np.int64)).to(self.device, non_blocking=True)
        n = self.n_step_return
        f = self.fields
        reward_seq = self._gather(f["reward"], t0, b, T, 0)
        done_seq = self._gather(f["done"], t0, b, T, 0)
        return SamplesFromReplay(
            all_observation=self._gather(f["observation"], t0, b, T + n, 0),
            all_action=self._gather(f["action"], t0, b, T + n, -1),   # starts at prev_action
            all_reward=self._gather(f["reward"], t0, b, T + n, -1),   # starts at prev_reward
            return_=reward_seq,   # n_step_return == 1: samples_return_ is samples.reward (rlpyt/replays/n_step.py:55-57)
            done=done_seq,
            done_n=done_seq,
            init_rnn_state=None,
        )
