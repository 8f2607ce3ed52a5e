"""The C-ABI shared library loads on a box without a GPU, exports every symbol declared in
include/dreamer_b200.h, and its struct layouts agree with the ctypes binding.  No compute calls."""
import ctypes
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "dreamer_b200.h")


@pytest.fixture(scope="module")
def lib():
    from dreamer_pytorch_b200 import _lib

    if not os.path.exists(_lib.LIB_PATH):
        import __graft_entry__

        __graft_entry__.build()
    return _lib.load()


def declared_functions():
    src = open(HEADER).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(drm_[a-z_0-9]+)\s*\(", src)))


def test_every_declared_symbol_is_exported(lib):
    names = declared_functions()
    assert len(names) >= 14
    for n in names:
        assert hasattr(lib, n), f"{n} declared in the header but not exported"


def test_binding_covers_the_header():
    from dreamer_pytorch_b200 import _lib

    assert set(declared_functions()) == set(_lib.SYMBOLS)


def test_struct_sizes_match(lib):
    from dreamer_pytorch_b200 import _lib

    for name, st in _lib.STRUCTS.items():
        assert lib.drm_sizeof(name.encode()) == ctypes.sizeof(st), name
    assert lib.drm_sizeof(b"no_such_struct") == 0
    assert lib.drm_version() == 1


def test_size_queries_without_gpu(lib):
    """Pure host arithmetic: workspace / save sizes for the BASELINE.json configuration."""
    from dreamer_pytorch_b200 import _lib

    r = _lib.drm_rollout()
    r.mode, r.steps, r.n, r.training = _lib.MODE_POLICY, 15, 2500, 1
    r.rssm.A, r.rssm.S, r.rssm.D, r.rssm.Hd = 6, 30, 200, 200
    r.actor.n_linear, r.actor.in_dim, r.actor.hidden, r.actor.out_dim = 4, 230, 200, 12
    save = lib.drm_rollout_save_bytes(ctypes.byref(r))
    per_state = (3 * 200 + 12 + 200 + 4 * 200 + 200) * 4  # actor hiddens, raw, x, gates, prior hidden
    assert per_state * 15 * 2500 <= save <= per_state * 15 * 2500 + 4096
    assert lib.drm_rollout_workspace_bytes(ctypes.byref(r)) > 0


def test_errors_are_reported_not_thrown(lib):
    from dreamer_pytorch_b200 import _lib

    r = _lib.drm_rollout()  # all zero: invalid
    rc = lib.drm_rollout_fwd(ctypes.byref(r), None)
    assert rc < 0
    assert len(lib.drm_last_error()) > 0
