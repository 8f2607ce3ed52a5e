"""drm_mlp_fwd_noalloc (include/dreamer_b200.h): the fp32 inference forward of the batched MLP with its packed-weight
scratch in a static __device__ array of the library instead of a caller workspace -- for callers that cannot call
cudaMalloc.  Not part of the reference's path.

Called through the C-ABI and compared with drm_mlp_fwd (the existing entry point, caller workspace) on the same inputs:
both run the same fp32 kernels in the same order, so the outputs must agree bit for bit (asserted at the fp32 parity
tolerance of the other tests, and exactly).  Two calls in a row with DIFFERENT shapes share the static array (the second
one finds the first one's packed weights in it), then the first shape again.  Collected last on purpose
(file name).
"""
import ctypes as C

import pytest
import torch

from _cases import FP32_TOL, rel_err

pytestmark = pytest.mark.gpu


def _weights(in_dim, hidden, out_dim, layers, dev, seed):
    g = torch.Generator().manual_seed(seed)
    dims = [in_dim] + [hidden] * (layers - 1) + [out_dim]
    ws = []
    for a, b in zip(dims[:-1], dims[1:]):
        ws.append((torch.randn(b, a, generator=g) / a ** 0.5).to(dev))
        ws.append((torch.randn(b, generator=g) * 0.1).to(dev))
    return ws


def _call(x0, x1, weights):
    from dreamer_pytorch_b200 import functional as F

    c = F._build_mlp_call(x0, x1, weights, x0.shape[0], training=False, precision="fp32")
    c._keep = (x0, x1, weights)
    return c


def _existing(x0, x1, weights):
    """drm_mlp_fwd with a caller-provided workspace (torch allocates it)."""
    from dreamer_pytorch_b200 import _lib

    lib = _lib.load()
    c = _call(x0, x1, weights)
    y = torch.full((x0.shape[0], c.mlp.out_dim), float("nan"), device=x0.device)
    ws = torch.empty(lib.drm_mlp_workspace_bytes(C.byref(c)) // 4 + 64, device=x0.device)
    c.y, c.workspace, c.workspace_bytes = y.data_ptr(), ws.data_ptr(), ws.numel() * 4
    _lib.check(lib.drm_mlp_fwd(C.byref(c), C.c_void_p(torch.cuda.current_stream().cuda_stream)), "drm_mlp_fwd")
    torch.cuda.synchronize()
    return y


def _noalloc(x0, x1, weights, y):
    """drm_mlp_fwd_noalloc: workspace / save left NULL; output buffer allocated by the caller beforehand."""
    from dreamer_pytorch_b200 import _lib

    lib = _lib.load()
    c = _call(x0, x1, weights)
    c.y = y.data_ptr()
    stream = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    before = torch.cuda.memory_stats()["allocation.all.allocated"]
    rc = lib.drm_mlp_fwd_noalloc(C.byref(c), stream)
    after = torch.cuda.memory_stats()["allocation.all.allocated"]
    _lib.check(rc, "drm_mlp_fwd_noalloc")
    assert after == before  # nothing went through torch's allocator either
    torch.cuda.synchronize()
    return y


def _torch_mlp(x, weights):
    h = x.double()
    n = len(weights) // 2
    for l in range(n):
        h = h @ weights[2 * l].double().t() + weights[2 * l + 1].double()
        if l < n - 1:
            h = torch.nn.functional.elu(h)
    return h


# (rows, k0, k1, hidden, out, linears): the reward head of the reference (230 -> 3 x 300 -> 1, dense.py) on a walker batch,
# then a small ragged two-segment MLP, so the second call finds the first one's larger packed weights in the array
SHAPES = [(2500, 30, 200, 300, 1, 4), (37, 30, 6, 64, 12, 2)]


def test_noalloc_matches_existing_entry_point_two_calls_in_a_row(cuda):
    from dreamer_pytorch_b200 import _lib

    _lib.ensure_init()
    cases = []
    for i, (M, k0, k1, hid, out, nl) in enumerate(SHAPES):
        g = torch.Generator().manual_seed(10 + i)
        x0 = torch.randn(M, k0, generator=g).to(cuda)
        x1 = torch.randn(M, k1, generator=g).to(cuda) if k1 else None
        w = _weights(k0 + k1, hid, out, nl, cuda, 20 + i)
        cases.append((x0, x1, w, _existing(x0, x1, w), torch.full((M, out), float("nan"), device=cuda)))
    # back to back, same stream, no synchronisation in between; then the first shape once more
    for x0, x1, w, ref, y in cases + cases[:1]:
        y.fill_(float("nan"))
        _noalloc(x0, x1, w, y)
        assert torch.isfinite(y).all()
        assert rel_err(y, ref) < FP32_TOL
        assert torch.equal(y, ref)  # same kernels, same order of operations
        x = x0 if x1 is None else torch.cat([x0, x1], 1)
        assert rel_err(y, _torch_mlp(x, w)) < FP32_TOL  # and both are the MLP torch computes


def test_noalloc_rejects_what_does_not_fit_or_is_not_inference(cuda):
    from dreamer_pytorch_b200 import _lib

    lib = _lib.load()
    stream = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    # 1024 -> 1024 -> 1024: 2 x 4 MiB of packed weights, beyond the 2 MiB static array
    x0 = torch.randn(8, 1024, device=cuda)
    w = _weights(1024, 1024, 1024, 2, cuda, 3)
    c = _call(x0, None, w)
    y = torch.zeros(8, 1024, device=cuda)
    c.y = y.data_ptr()
    assert lib.drm_mlp_noalloc_scratch_bytes(C.byref(c)) > _lib.NOALLOC_SCRATCH_BYTES
    assert lib.drm_mlp_fwd_noalloc(C.byref(c), stream) == -4  # DRM_E_WORKSPACE
    msg = lib.drm_last_error().decode()
    assert "DRM_NOALLOC_SCRATCH_BYTES" in msg and str(_lib.NOALLOC_SCRATCH_BYTES) in msg
    assert float(y.abs().sum()) == 0.0  # nothing was launched
    # the reference's widest head fits with room to spare
    x0 = torch.randn(4, 230, device=cuda)
    w = _weights(230, 300, 1, 4, cuda, 4)
    c = _call(x0, None, w)
    assert 0 < lib.drm_mlp_noalloc_scratch_bytes(C.byref(c)) <= _lib.NOALLOC_SCRATCH_BYTES
    c.y = torch.zeros(4, 1, device=cuda).data_ptr()
    c.training = 1
    assert lib.drm_mlp_fwd_noalloc(C.byref(c), stream) == -6  # DRM_E_ARG: inference only
    c.training, c.precision = 0, _lib.BF16
    assert lib.drm_mlp_fwd_noalloc(C.byref(c), stream) == -6  # fp32 kernels only
