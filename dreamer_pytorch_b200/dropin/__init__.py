"""Host-side mirror of the reference's model interface for the RSSM hot path.

``install()`` makes these modules answer to the reference's import paths
(``dreamer.models.rnns`` etc.), so ``dreamer/algos/dreamer_algo.py`` and the rlpyt agents run
unchanged on the fused CUDA kernels.
"""
import sys
import types

from . import action, agent, dense, distribution, module, observation, rnns  # noqa: F401

_ALIASES = {
    "dreamer.models.rnns": rnns,
    "dreamer.models.action": action,
    "dreamer.models.dense": dense,
    "dreamer.models.distribution": distribution,
    "dreamer.models.agent": agent,
    "dreamer.models.observation": observation,
    "dreamer.utils.module": module,
}


def install():
    """Alias the drop-in modules under the reference's module paths.  If the reference's own
    ``dreamer`` package is importable it stays the parent package (its ``algos``, ``agents`` ...
    keep working); otherwise empty parent packages are created."""
    for parent in ("dreamer", "dreamer.models", "dreamer.utils"):
        if parent not in sys.modules:
            try:
                __import__(parent)
            except Exception:
                pkg = types.ModuleType(parent)
                pkg.__path__ = []
                sys.modules[parent] = pkg
    for name, mod in _ALIASES.items():
        sys.modules[name] = mod
        parent, _, leaf = name.rpartition(".")
        setattr(sys.modules[parent], leaf, mod)
