"""MMA-phase durations of the bf16 MLP forward (CTA 0, -DDRM_TIMELINE build) as a function of how many CTAs stream the
same weights at once: python tools/timeline_mlp_ctas.py [hidden]"""
import ctypes as C
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from dreamer_pytorch_b200 import _lib  # noqa: E402
from dreamer_pytorch_b200.functional import mlp_forward_raw  # noqa: E402

dev = torch.device("cuda:0")
torch.manual_seed(0)
hid = int(sys.argv[1]) if len(sys.argv) > 1 else 200
ws, d = [], 230
for _ in range(3):
    ws += [torch.randn(hid, d, device=dev) / d ** 0.5, torch.zeros(hid, device=dev)]
    d = hid
ws += [torch.randn(1, d, device=dev) / d ** 0.5, torch.zeros(1, device=dev)]
lib = _lib.load()
stamps = (C.c_longlong * (4 * 4096))()
counts = (C.c_int * 4)()
for ctas in (1, 4, 16, 37, 74, 148):
    M = ctas * 128 * 4
    x0 = torch.randn(M, 30, device=dev)
    x1 = torch.randn(M, 200, device=dev)
    for rep in range(2):
        mlp_forward_raw(x0, x1, ws, True, "bf16", max_ctas=ctas)
        torch.cuda.synchronize()
        lib.drm_debug_timeline_mlp(stamps, counts)
    e = [stamps[i] for i in range(counts[0])]
    m = [stamps[4096 + i] for i in range(counts[1])]
    k = 8
    w = [stamps[3 * 4096 + i] for i in range(counts[3])]
    if ctas == 1 and len(w) > 40 and len(m) > 16:
        # merged timeline of tile 1 on one clock: E = epilogue warp 0 (wake / release alternate), M = MMA warp (phase start /
        # final commit alternate), w = wait for a weight slab (begin, end)
        t0 = m[8]
        ev = [(x - t0, "M") for x in m[8:16]] + [(x - t0, "E") for x in e if m[8] - 3000 <= x <= m[16]]
        ev += [(w[i] - t0, f"w{w[i + 1] - w[i]}") for i in range(0, len(w) - 1, 2) if m[8] <= w[i] < m[16]]
        print("   " + " ".join(f"{t}:{k}" for t, k in sorted(ev)))
    print(f"ctas {ctas:4d}: tile period {[e[(i + 1) * k] - e[i * k] for i in range(min(3, len(e) // k - 1))]}"
          f"  mma (wait, run) x 4 layers of tile 1: {[m[i + 1] - m[i] for i in range(8, 16)]}")
