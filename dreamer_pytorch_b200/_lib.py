"""ctypes binding of ``libdreamer_b200.so`` (C ABI declared in ``include/dreamer_b200.h``).

The library is hand-written CUDA for sm_100a.  There is no CPU fallback: if the shared object is
missing or a call fails, a ``RuntimeError`` is raised.
"""
from __future__ import annotations

import ctypes as C
import os

MAXL = 8  # DRM_MAX_LINEAR
_HERE = os.path.dirname(os.path.abspath(__file__))
# DRM_B200_LIB: developer override (e.g. the -DDRM_TIMELINE build of tools/timeline.py); same ABI, same checks.
LIB_PATH = os.environ.get("DRM_B200_LIB") or os.path.join(_HERE, "libdreamer_b200.so")

MODE_POLICY, MODE_TRANSITION, MODE_REPRESENTATION = 0, 1, 2
DIST = {"tanh_normal": 0, "one_hot": 1, "relaxed_one_hot": 2}
FP32, BF16 = 0, 1

P = C.c_void_p


class drm_mlp(C.Structure):
    _fields_ = [("n_linear", C.c_int32), ("in_dim", C.c_int32), ("hidden", C.c_int32), ("out_dim", C.c_int32),
                ("W", P * MAXL), ("b", P * MAXL)]


class drm_mlp_grad(C.Structure):
    _fields_ = [("dW", P * MAXL), ("db", P * MAXL)]


_RSSM_W = ["W_in", "b_in", "W_ih", "W_hh", "b_ih", "b_hh", "W_p1", "b_p1", "W_p2", "b_p2", "W_q1", "b_q1", "W_q2",
           "b_q2"]


class drm_rssm(C.Structure):
    _fields_ = [(k, C.c_int32) for k in ("A", "S", "D", "Hd", "E")] + [(k, P) for k in _RSSM_W]


_RSSM_G = ["dW_in", "db_in", "dW_ih", "dW_hh", "db_ih", "db_hh", "dW_p1", "db_p1", "dW_p2", "db_p2", "dW_q1", "db_q1",
           "dW_q2", "db_q2"]


class drm_rssm_grad(C.Structure):
    _fields_ = [(k, P) for k in _RSSM_G]


class drm_rollout(C.Structure):
    _fields_ = (
        [(k, C.c_int32) for k in ("mode", "steps", "n", "act_dist", "precision", "training")]
        + [(k, C.c_float) for k in ("act_mean_scale", "act_raw_init_std", "act_min_std", "act_temperature", "min_std")]
        + [("rssm", drm_rssm), ("actor", drm_mlp)]
        + [(k, P) for k in ("start_stoch", "start_deter", "action_in", "obs_embed", "eps_prior", "eps_post",
                            "noise_act", "prior_mean", "prior_std", "prior_stoch", "deter", "post_mean", "post_std",
                            "post_stoch", "action_out")]
        + [("save", P), ("save_bytes", C.c_size_t), ("workspace", P), ("workspace_bytes", C.c_size_t), ("step_flags", P), ("noise_pair", P)]
    )


PHASE_ALL, PHASE_PREPARE, PHASE_RUN = 0, 1, 2


class drm_rollout_grad(C.Structure):
    _fields_ = (
        [(k, P) for k in ("g_prior_mean", "g_prior_std", "g_prior_stoch", "g_deter", "g_post_mean", "g_post_std",
                          "g_post_stoch", "g_action_out", "d_start_stoch", "d_start_deter", "d_action_in",
                          "d_obs_embed")]
        + [("rssm", drm_rssm_grad), ("want_rssm_grad", C.c_int32), ("want_actor_grad", C.c_int32),
           ("actor", drm_mlp_grad), ("phase", C.c_int32), ("g_wait_target", C.c_int32), ("g_wait_flags", P),
           ("wgrad_stream", P)]
    )


class drm_mlp_call(C.Structure):
    _fields_ = [("mlp", drm_mlp), ("M", C.c_int32), ("k0", C.c_int32), ("k1", C.c_int32), ("precision", C.c_int32),
                ("training", C.c_int32), ("x0", P), ("ld0", C.c_int64), ("x1", P), ("ld1", C.c_int64), ("y", P),
                ("save", P), ("save_bytes", C.c_size_t), ("workspace", P), ("workspace_bytes", C.c_size_t),
                ("wait_flags", P), ("wait_target", C.c_int32), ("wait_rows", C.c_int32), ("wait_steps", C.c_int32),
                ("max_ctas", C.c_int32)]


class drm_mlp_bwd_call(C.Structure):
    _fields_ = [("fwd", drm_mlp_call), ("dy", P), ("dy_scale", P), ("dx0", P), ("dx1", P),
                ("accumulate_dx", C.c_int32), ("want_wgrad", C.c_int32), ("grad", drm_mlp_grad),
                ("phase", C.c_int32), ("reverse_tiles", C.c_int32), ("tile_done", P), ("tile_wait", P),
                ("step_rows_done", P), ("step_rows", C.c_int32), ("step_count", C.c_int32)]


class drm_returns_loss_call(C.Structure):
    _fields_ = [("H", C.c_int32), ("N", C.c_int32), ("discount", C.c_float), ("lambda_", C.c_float)] + [
        (k, P) for k in ("reward", "value", "pcont", "returns", "disc", "losses", "d_reward", "d_value", "d_pcont",
                         "d_value_pred")]


class drm_kl_call(C.Structure):
    _fields_ = [("M", C.c_int32), ("S", C.c_int32), ("free_nats", C.c_float)] + [
        (k, P) for k in ("post_mean", "post_std", "prior_mean", "prior_std", "out", "d_post_mean", "d_post_std",
                         "d_prior_mean", "d_prior_std")]


class drm_clip_adam_call(C.Structure):
    _fields_ = ([("n_tensors", C.c_int32), ("pad_", C.c_int32)]
                + [(k, C.c_float) for k in ("max_norm", "beta1", "beta2", "eps", "one_minus_beta1", "one_minus_beta2", "step_size", "bias2_sqrt")]
                + [("lr", C.c_float), ("pad2_", C.c_int32), ("beta1_d", C.c_double), ("beta2_d", C.c_double), ("step_dev", P)]
                + [(k, P * 64) for k in ("param", "grad", "exp_avg", "exp_avg_sq")]
                + [("numel", C.c_int64 * 64), ("norm_out", P), ("workspace", P), ("workspace_bytes", C.c_size_t)]
                + [("mult", C.c_int32 * 64), ("update", C.c_int32), ("pad3_", C.c_int32)])


class drm_image_logprob_call(C.Structure):
    _fields_ = [("rows", C.c_int64), ("D", C.c_int32), ("mean_bf16", C.c_int32), ("mean", P), ("obs", P), ("logp", P),
                ("g", P), ("dmean", P)]


class drm_policy_mode_call(C.Structure):
    _fields_ = [("n", C.c_int32), ("A", C.c_int32), ("samples", C.c_int32), ("mean_scale", C.c_float),
                ("raw_init_std", C.c_float), ("min_std", C.c_float), ("raw", P), ("noise", P), ("action", P), ("logp", P),
                ("index", P)]


class drm_replay_gather_call(C.Structure):
    _fields_ = [("T_cap", C.c_int64), ("B_env", C.c_int32), ("T_seq", C.c_int32), ("batch", C.c_int32),
                ("t_offset", C.c_int32), ("row_bytes", C.c_int64), ("src", P), ("dst", P), ("t0", P), ("b", P)]


MAX_NOISE_BLOCKS = 4
NOISE_NORMAL, NOISE_GUMBEL = 0, 1
NOISE_STREAMS = {"eps_prior": 0, "eps_post": 1, "noise_act": 2, "mode_noise": 3}


class drm_noise_fill_call(C.Structure):
    _fields_ = [("state", P), ("pair", P), ("pair_out", P), ("n_blocks", C.c_int32), ("pad_", C.c_int32),
                ("out", P * MAX_NOISE_BLOCKS), ("numel", C.c_int64 * MAX_NOISE_BLOCKS),
                ("stream_id", C.c_int32 * MAX_NOISE_BLOCKS), ("kind", C.c_int32 * MAX_NOISE_BLOCKS)]


class drm_bias_act_call(C.Structure):
    _fields_ = [("rows", C.c_int64), ("C", C.c_int32), ("relu", C.c_int32), ("x", P), ("bias", P), ("g", P), ("y", P),
                ("g_pre", P), ("dbias", P)]


MAX_RANKS, AR_MAX_CTAS = 8, 64


class drm_allreduce_call(C.Structure):
    _fields_ = [("world", C.c_int32), ("rank", C.c_int32), ("ctas", C.c_int32), ("scale", C.c_float), ("numel", C.c_int64),
                ("algo", C.c_int32), ("pad_", C.c_int32), ("buf", P * MAX_RANKS), ("signal", P * MAX_RANKS), ("staging", P), ("epoch", P)]


STRUCTS = {s.__name__: s for s in (drm_bias_act_call, drm_allreduce_call, drm_noise_fill_call, drm_replay_gather_call, drm_policy_mode_call, drm_image_logprob_call, drm_clip_adam_call, drm_mlp, drm_mlp_grad, drm_rssm, drm_rssm_grad, drm_rollout, drm_rollout_grad,
                                   drm_mlp_call, drm_mlp_bwd_call, drm_returns_loss_call, drm_kl_call)}

NOALLOC_SCRATCH_BYTES = 2 << 20  # DRM_NOALLOC_SCRATCH_BYTES

# every symbol include/dreamer_b200.h declares: name -> (restype, argtypes)
SYMBOLS = {
    "drm_version": (C.c_int, []),
    "drm_init": (C.c_int, []),
    "drm_num_sms": (C.c_int, []),
    "drm_last_error": (C.c_char_p, []),
    "drm_sizeof": (C.c_size_t, [C.c_char_p]),
    "drm_launch_count": (C.c_ulonglong, []),
    "drm_prof_enable": (None, [C.c_int]),
    "drm_prof_read": (C.c_int, [C.c_int, C.POINTER(C.c_double), C.POINTER(C.c_int)]),
    "drm_rollout_save_bytes": (C.c_size_t, [C.POINTER(drm_rollout)]),
    "drm_rollout_workspace_bytes": (C.c_size_t, [C.POINTER(drm_rollout)]),
    "drm_rollout_fwd": (C.c_int, [C.POINTER(drm_rollout), P]),
    "drm_rollout_tiles": (C.c_int, [C.POINTER(drm_rollout)]),
    "drm_rollout_bwd": (C.c_int, [C.POINTER(drm_rollout), C.POINTER(drm_rollout_grad), P]),
    "drm_mlp_save_bytes": (C.c_size_t, [C.POINTER(drm_mlp_call)]),
    "drm_mlp_workspace_bytes": (C.c_size_t, [C.POINTER(drm_mlp_call)]),
    "drm_mlp_fwd": (C.c_int, [C.POINTER(drm_mlp_call), P]),
    "drm_mlp_bwd": (C.c_int, [C.POINTER(drm_mlp_bwd_call), P]),
    "drm_mlp_noalloc_scratch_bytes": (C.c_size_t, [C.POINTER(drm_mlp_call)]),
    "drm_mlp_fwd_noalloc": (C.c_int, [C.POINTER(drm_mlp_call), P]),
    "drm_returns_loss": (C.c_int, [C.POINTER(drm_returns_loss_call), P]),
    "drm_clip_adam": (C.c_int, [C.POINTER(drm_clip_adam_call), P]),
    "drm_image_logprob_fwd": (C.c_int, [C.POINTER(drm_image_logprob_call), P]),
    "drm_image_logprob_bwd": (C.c_int, [C.POINTER(drm_image_logprob_call), P]),
    "drm_policy_mode": (C.c_int, [C.POINTER(drm_policy_mode_call), P]),
    "drm_replay_gather": (C.c_int, [C.POINTER(drm_replay_gather_call), P]),
    "drm_allreduce": (C.c_int, [C.POINTER(drm_allreduce_call), P]),
    "drm_bias_act_fwd": (C.c_int, [C.POINTER(drm_bias_act_call), P]),
    "drm_bias_act_bwd": (C.c_int, [C.POINTER(drm_bias_act_call), P]),
    "drm_noise_next": (C.c_int, [P, P, P]),
    "drm_noise_fill": (C.c_int, [C.POINTER(drm_noise_fill_call), P]),
    "drm_noise_words": (C.c_int, [P, C.c_int, C.c_ulonglong, C.c_int, P, P]),
    "drm_kl_fwd": (C.c_int, [C.POINTER(drm_kl_call), P]),
    "drm_kl_bwd": (C.c_int, [C.POINTER(drm_kl_call), P]),
}

_lib = None
launch_count = 0  # kernels-launching C-ABI calls made through this binding (bench bookkeeping)


def load():
    """Load the shared library once; raises RuntimeError (no fallback) when it is missing."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RuntimeError(
            f"{LIB_PATH} not found: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
            "(nvcc, sm_100a). dreamer_pytorch_b200 has no CPU or PyTorch fallback.")
    try:
        import torch  # noqa: F401  (makes torch's libcudart the one both sides share)
    except Exception:  # pragma: no cover
        pass
    lib = C.CDLL(LIB_PATH)
    for name, (res, args) in SYMBOLS.items():
        fn = getattr(lib, name)
        fn.restype = res
        fn.argtypes = args
    if lib.drm_version() != 1:
        raise RuntimeError("libdreamer_b200.so ABI version mismatch")
    for name, st in STRUCTS.items():
        got = lib.drm_sizeof(name.encode())
        if got != C.sizeof(st):
            raise RuntimeError(f"ABI mismatch for {name}: C says {got} bytes, ctypes says {C.sizeof(st)}")
    _lib = lib
    return lib


_inited = set()


def ensure_init():
    """drm_init() once per device: architecture check + eager load of every kernel (a kernel that spins on the
    counters of a producer must never be resident while another kernel is loaded lazily at its first launch)."""
    import torch

    dev = torch.cuda.current_device()
    if dev not in _inited:
        check(load().drm_init(), "drm_init")
        _inited.add(dev)


def check(rc: int, what: str):
    if rc != 0:
        msg = load().drm_last_error().decode(errors="replace")
        raise RuntimeError(f"{what} failed with drm_status {rc}: {msg}")


def ptr(t):
    """Device pointer of a contiguous fp32 CUDA tensor (or NULL for None)."""
    if t is None:
        return None
    if not t.is_cuda:
        raise RuntimeError("dreamer_pytorch_b200 ops need CUDA tensors (there is no CPU fallback)")
    import torch

    if t.dtype != torch.float32 or not t.is_contiguous():
        raise RuntimeError(f"expected a contiguous float32 tensor, got {t.dtype} contiguous={t.is_contiguous()}")
    return t.data_ptr()
