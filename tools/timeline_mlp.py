"""Handshake timeline of the bf16 batched-MLP kernels (CTA 0), from a -DDRM_TIMELINE build (see tools/timeline.py).
Prints raw stamp deltas: epilogue warp 0 alternates (wake, release) per phase; the MMA thread (start, commit)."""
import ctypes as C
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from dreamer_pytorch_b200 import _lib  # noqa: E402
from dreamer_pytorch_b200.functional import MlpFn  # noqa: E402
from dreamer_pytorch_b200.precision import set_precision  # noqa: E402

dev = torch.device("cuda:0")
set_precision("bf16")
torch.manual_seed(0)
M = int(sys.argv[1]) if len(sys.argv) > 1 else 148 * 128 * 4
hid = int(sys.argv[2]) if len(sys.argv) > 2 else 200
ws, d = [], 230
for _ in range(3):
    ws += [torch.randn(hid, d, device=dev) / d ** 0.5, torch.zeros(hid, device=dev)]
    d = hid
ws += [torch.randn(1, d, device=dev) / d ** 0.5, torch.zeros(1, device=dev)]
ws = [w.requires_grad_(True) for w in ws]
x0 = torch.randn(M, 30, device=dev, requires_grad=True)
x1 = torch.randn(M, 200, device=dev, requires_grad=True)
lib = _lib.load()
stamps = (C.c_longlong * (4 * 4096))()
counts = (C.c_int * 4)()
flush = torch.empty(512 << 20, dtype=torch.uint8, device=dev)


def dump(tag, per_tile):
    lib.drm_debug_timeline_mlp(stamps, counts)
    e = [stamps[i] for i in range(counts[0])]
    m = [stamps[4096 + i] for i in range(counts[1])]
    print(f"== {tag}: epilogue stamps {len(e)}, mma stamps {len(m)}")
    k = per_tile
    if len(e) > 3 * k:
        base = e[k]
        print("  epilogue (tile 1), cycles since its first stamp:", [x - base for x in e[k:2 * k + 1]])
        print("  tile period:", [e[(i + 1) * k] - e[i * k] for i in range(min(6, len(e) // k - 1))])
    if len(m) > 16:
        print("  mma (start, commit) deltas:", [m[i + 1] - m[i] for i in range(8, 24)])


for rep in range(2):
    flush.fill_(rep)
    y = MlpFn.apply(x0, x1, *ws)
    torch.cuda.synchronize()
    if rep == 1:
        dump("mlp_fwd_tc", 8)   # per tile: release(stage), then (wake, release) x 3 hidden layers, wake (output)
    else:
        lib.drm_debug_timeline_mlp(stamps, counts)
    flush.fill_(rep + 9)
    y.sum().backward()
    torch.cuda.synchronize()
    if rep == 1:
        dump("mlp_bwd_tc (+ wgrad)", 8)
    else:
        lib.drm_debug_timeline_mlp(stamps, counts)
