"""bf16 tensor-core (tcgen05) path against fp64 references.  Tolerance: BASELINE.json north_star,
max relative (normwise) error <= 1e-2 on the bf16 path."""
import math

import pytest
import torch

from _cases import rel_err

pytestmark = pytest.mark.gpu

BF16_TOL = 1e-2


def _mlp_case(M, k0, k1, hidden, out, layers, seed=5):
    g = torch.Generator().manual_seed(seed)
    ws, d = [], k0 + k1
    for _ in range(layers):
        ws += [torch.randn(hidden, d, generator=g) / math.sqrt(d), torch.randn(hidden, generator=g) * 0.1]
        d = hidden
    ws += [torch.randn(out, d, generator=g) / math.sqrt(d), torch.randn(out, generator=g) * 0.1]
    x0 = torch.randn(M, k0, generator=g)
    x1 = torch.randn(M, k1, generator=g) if k1 else None
    gy = torch.randn(M, out, generator=g)
    return ws, x0, x1, gy


def _mlp_ref(x0, x1, ws, layers):
    x = torch.cat([x0, x1], -1) if x1 is not None else x0
    for l in range(layers):
        x = torch.nn.functional.elu(torch.nn.functional.linear(x, ws[2 * l], ws[2 * l + 1]))
    return torch.nn.functional.linear(x, ws[-2], ws[-1])


MLP_SHAPES = [
    (777, 30, 200, 200, 1, 3),     # value head, ragged M
    (50, 30, 200, 300, 1, 3),      # reward head (hidden 300 -> two MMA N parts)
    (1, 8, 0, 32, 5, 1),           # single row, single segment
    (2500, 30, 200, 200, 12, 3),   # actor-shaped, 20 tiles
    (333, 7, 9, 36, 40, 2),        # odd segment sizes (padding paths)
    (37500, 30, 200, 200, 1, 3),   # full size: 293 tiles > 148 CTAs (multi-tile loop per CTA)
]


@pytest.mark.parametrize("M,k0,k1,hidden,out,layers", MLP_SHAPES)
def test_mlp_fwd_bf16(cuda, M, k0, k1, hidden, out, layers):
    from dreamer_pytorch_b200.functional import mlp_forward_raw

    ws, x0, x1, _ = _mlp_case(M, k0, k1, hidden, out, layers)
    ry = _mlp_ref(x0.double(), None if x1 is None else x1.double(), [w.double() for w in ws], layers)
    cw = [w.to(cuda) for w in ws]
    y, _ = mlp_forward_raw(x0.to(cuda), None if x1 is None else x1.to(cuda), cw, False, precision="bf16")
    torch.cuda.synchronize()
    err = rel_err(y, ry)
    assert err < BF16_TOL, err
    assert err > 0  # bf16 operands: an exact match would mean the tensor-core path did not run
