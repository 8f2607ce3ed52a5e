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
