"""dreamer_pytorch_b200 -- the RSSM rollout hot path of juliusfrost/dreamer-pytorch as hand-written
CUDA for sm_100a behind the reference's module API.  See DESIGN.md."""
from . import _lib, noise  # noqa: F401

__all__ = ["_lib", "noise", "functional", "dropin", "algo"]
