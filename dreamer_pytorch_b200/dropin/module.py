"""Drop-in for ``dreamer/utils/module.py`` (module.py:5-38): ``get_parameters`` and the
``FreezeParameters`` context manager.  The fused ops read ``requires_grad`` at call time
(``ctx.needs_input_grad``), so freezing a module here is what makes the backward kernels skip its
weight-gradient GEMMs."""
from typing import Iterable

from torch.nn import Module


def get_parameters(modules: Iterable[Module]):
    params = []
    for m in modules:
        params.extend(m.parameters())
    return params


class FreezeParameters:
    def __init__(self, modules: Iterable[Module]):
        self.modules = modules
        self.param_states = [p.requires_grad for p in get_parameters(self.modules)]

    def __enter__(self):
        for p in get_parameters(self.modules):
            p.requires_grad = False

    def __exit__(self, exc_type, exc_val, exc_tb):
        for p, state in zip(get_parameters(self.modules), self.param_states):
            p.requires_grad = state
