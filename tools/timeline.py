"""Per-phase handshake timeline of the bf16 rollout kernels (CTA 0), from a -DDRM_TIMELINE build.

    make -C dreamer_pytorch_b200/csrc OBJDIR=build_tl LIB=../libdreamer_b200_tl.so EXTRA=-DDRM_TIMELINE
    DRM_B200_LIB=dreamer_pytorch_b200/libdreamer_b200_tl.so python tools/timeline.py [n_states]

Roles: 0 = epilogue warp 0 (wait_acc return, release), 1 = MMA thread (a_ready seen, phase committed),
2 = epilogue warp 15.  Prints, per phase of one time step, MMA span, commit->epilogue wake, epilogue span,
release->MMA wake (SM cycles)."""
import ctypes as C
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from dreamer_pytorch_b200 import _lib  # noqa: E402
from dreamer_pytorch_b200.functional import RolloutCfg, RolloutFn  # noqa: E402
from dreamer_pytorch_b200.precision import set_precision  # noqa: E402

dev = torch.device("cuda:0")
set_precision("bf16")
torch.manual_seed(0)
n = int(sys.argv[1]) if len(sys.argv) > 1 else 2500
A, S, D, Hd, H = 6, 30, 200, 200, 15


def lin(o, i):
    return [torch.randn(o, i, device=dev) / i ** 0.5, torch.zeros(o, device=dev)]


w_in = lin(Hd, A + S)
rssm = [w_in[0], w_in[1], torch.randn(3 * D, Hd, device=dev) / Hd ** 0.5, torch.randn(3 * D, D, device=dev) / D ** 0.5,
        torch.zeros(3 * D, device=dev), torch.zeros(3 * D, device=dev)] + lin(Hd, D) + lin(2 * S, Hd)
actor = [w.requires_grad_(True) for w in lin(200, S + D) + lin(200, 200) + lin(200, 200) + lin(2 * A, 200)]
cfg = RolloutCfg(_lib.MODE_POLICY, H, n, (A, S, D, Hd, 0), act_dist="tanh_normal")
ss, sd = torch.randn(n, S, device=dev), torch.tanh(torch.randn(n, D, device=dev))
lib = _lib.load()
stamps = (C.c_longlong * (4 * 4096))()
counts = (C.c_int * 4)()
flush = torch.empty(512 << 20, dtype=torch.uint8, device=dev)


def dump(tag):
    lib.drm_debug_timeline(stamps, counts)
    e = [stamps[i] for i in range(counts[0])]
    m = [stamps[4096 + i] for i in range(counts[1])]
    e2 = [stamps[2 * 4096 + i] for i in range(counts[2])]
    print(f"== {tag}: epilogue stamps {len(e)}, mma stamps {len(m)}")
    # epilogue stamps: [release(start)] then per phase (wake, release); mma: per phase (start, commit)
    e, e2 = e[1:], e2[1:]
    nph = len(m) // 2 // H
    print(f"phases per step: {nph}; step cycles: {[m[2 * nph * (t + 1)] - m[2 * nph * t] for t in range(H - 1)]}")
    for t in (1, 7):
        print(f"-- step {t}:  phase | mma_span | commit->wake | epi_span(w0) | epi_span(w15) | release->mma")
        for ph in range(nph):
            k = t * nph + ph
            ms, mc = m[2 * k], m[2 * k + 1]
            ew, er = e[2 * k], e[2 * k + 1] if 2 * k + 1 < len(e) else e[2 * k]
            ew2, er2 = e2[2 * k], e2[2 * k + 1] if 2 * k + 1 < len(e2) else e2[2 * k]
            nxt = m[2 * k + 2] if 2 * k + 2 < len(m) else er
            print(f"   {ph:2d} | {mc - ms:6d} | {ew - mc:6d} | {er - ew:6d} | {er2 - ew2:6d} | {nxt - max(er, er2):6d}")


for rep in range(2):
    flush.fill_(rep)
    outs = RolloutFn.apply(cfg, ss, sd, None, None, torch.randn(H, n, S, device=dev), None, torch.randn(H, n, A, device=dev),
                           *rssm, None, None, None, None, *actor)
    torch.cuda.synchronize()
    if rep == 1:
        dump("rollout_fwd_tc")
    else:
        lib.drm_debug_timeline(stamps, counts)
    flush.fill_(rep + 7)
    (outs[2].sum() + outs[3].sum()).backward()
    torch.cuda.synchronize()
    if rep == 1:
        dump("rollout_bwd_tc")
    else:
        lib.drm_debug_timeline(stamps, counts)
