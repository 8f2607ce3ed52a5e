"""Arithmetic mode of the fused kernels.

``fp32``  fp32 FFMA kernels; parity budget 1e-4 (normwise) against the reference.
``bf16``  tcgen05 tensor-core kernels: bf16 operands, fp32 accumulate in TMEM, fp32 pointwise math
          and fp32 carried state; parity budget 1e-2.

The default comes from ``DREAMER_B200_PRECISION`` (fp32 when unset); ``set_precision`` changes it for
the process and ``precision(...)`` is a context manager.  A rollout / MLP remembers the mode it ran
forward in, so its backward always matches.
"""
import contextlib
import os

_VALID = ("fp32", "bf16")
_current = os.environ.get("DREAMER_B200_PRECISION", "fp32").lower()
if _current not in _VALID:
    raise RuntimeError(f"DREAMER_B200_PRECISION must be one of {_VALID}, got {_current!r}")


def get_precision() -> str:
    return _current


def set_precision(mode: str) -> None:
    global _current
    if mode not in _VALID:
        raise ValueError(f"precision must be one of {_VALID}, got {mode!r}")
    _current = mode


@contextlib.contextmanager
def precision(mode: str):
    old = get_precision()
    set_precision(mode)
    try:
        yield
    finally:
        set_precision(old)


# ---- convolution stack (observation encoder / decoder: library calls) -------------------------------
# "nchw"           the reference's fp32 NCHW calls, bit for bit (default under precision fp32)
# "channels_last"  fp32 tensors in NHWC memory format: cuDNN's tensor-op kernels without the NCHW<->NHWC conversion kernels
#                  around every layer (default under precision bf16; TF32 math where torch.backends.cudnn.allow_tf32 permits)
# "bf16"           channels-last under bf16 autocast: fastest; costs 1-3e-2 (normwise) on the world-model gradients that
#                  flow back through the decoder and ~6e-2 on the first encoder layers' weight gradients -- opt in
_CONV = ("auto", "nchw", "channels_last", "bf16")
_conv = os.environ.get("DREAMER_B200_CONV", "auto").lower()
if _conv not in _CONV:
    raise RuntimeError(f"DREAMER_B200_CONV must be one of {_CONV}, got {_conv!r}")


def get_conv_mode() -> str:
    if _conv != "auto":
        return _conv
    return "channels_last" if _current == "bf16" else "nchw"


def set_conv_mode(mode: str) -> None:
    global _conv
    if mode not in _CONV:
        raise ValueError(f"conv mode must be one of {_CONV}, got {mode!r}")
    _conv = mode


@contextlib.contextmanager
def conv_mode(mode: str):
    global _conv
    old = _conv
    set_conv_mode(mode)
    try:
        yield
    finally:
        _conv = old
