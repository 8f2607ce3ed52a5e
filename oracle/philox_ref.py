"""CPU restatement of the production noise streams -- TEST INFRASTRUCTURE ONLY (never imported by the product).

Philox4x32-10 as published (Salmon, Moraes, Dror, Shaw: "Parallel random numbers: as easy as 1, 2, 3", SC'11; Random123
1.x `philox.h`), pinned by the known-answer vectors of Random123's `kat_vectors` file, and the key layout / Box-Muller /
Gumbel mapping of `dreamer_pytorch_b200/csrc/drm_philox.cuh`.  The reference itself has no counterpart: it draws
`torch.randn` per step (rnns.py:104,155; action.py:58-65) -- what is pinned here is the generator, not a reference stream.
"""
from __future__ import annotations

import numpy as np

M0, M1 = np.uint64(0xD2511F53), np.uint64(0xCD9E8D57)
W0, W1 = 0x9E3779B9, 0xBB67AE85
MASK = np.uint64(0xFFFFFFFF)

# Random123 kat_vectors, "philox4x32 10": (counter[4], key[2]) -> output[4]
KAT = [
    ((0x00000000,) * 4, (0x00000000,) * 2, (0x6627E8D5, 0xE169C58D, 0xBC57AC4C, 0x9B00DBD8)),
    ((0xFFFFFFFF,) * 4, (0xFFFFFFFF,) * 2, (0x408F276D, 0x41C83B0E, 0xA20BC7C6, 0x6D5451FD)),
    ((0x243F6A88, 0x85A308D3, 0x13198A2E, 0x03707344), (0xA4093822, 0x299F31D0),
     (0xD16CFE09, 0x94FDCCEB, 0x5001E420, 0x24126EA1)),
]


def philox4x32_10(c, k):
    """c: uint32 array [..., 4], k: (k0, k1) -> uint32 array [..., 4]."""
    c = np.asarray(c, dtype=np.uint64)
    c0, c1, c2, c3 = (c[..., i].copy() for i in range(4))
    k0, k1 = int(k[0]), int(k[1])
    for _ in range(10):
        p0, p1 = M0 * c0, M1 * c2
        hi0, lo0, hi1, lo1 = p0 >> np.uint64(32), p0 & MASK, p1 >> np.uint64(32), p1 & MASK
        c0, c1, c2, c3 = hi1 ^ c1 ^ np.uint64(k0), lo1, hi0 ^ c3 ^ np.uint64(k1), lo0
        k0, k1 = (k0 + W0) & 0xFFFFFFFF, (k1 + W1) & 0xFFFFFFFF
    return np.stack([c0, c1, c2, c3], -1).astype(np.uint32)


def words(seed: int, offset: int, stream: int, groups) -> np.ndarray:
    """Raw words of the given group indices (element index // 4) of a stream: drm_philox.cuh `noise_words`."""
    g = np.asarray(groups, dtype=np.uint64)
    c = np.stack([g & MASK, ((g >> np.uint64(32)) & np.uint64(0x00FFFFFF)) | np.uint64(stream << 24),
                  np.full_like(g, offset & 0xFFFFFFFF), np.full_like(g, (offset >> 32) & 0xFFFFFFFF)], -1)
    return philox4x32_10(c, (seed & 0xFFFFFFFF, (seed >> 32) & 0xFFFFFFFF))


def stream_values(seed: int, offset: int, stream: int, kind: str, n: int) -> np.ndarray:
    """The first n elements of a stream (float64 arithmetic on the float32 uniforms the kernel forms)."""
    w = words(seed, offset, stream, np.arange((n + 3) // 4)).astype(np.float64)
    unit = (w * np.float32(2.3283064365386963e-10) + np.float32(2.3283064365386963e-10)).astype(np.float32).astype(np.float64)
    if kind == "gumbel":
        u = np.minimum(unit, np.float64(np.float32(0.99999994)))
        return (-np.log(-np.log(u))).reshape(-1)[:n]
    r = np.sqrt(-2.0 * np.log(unit[:, [0, 2]]))
    th = (w[:, [1, 3]].astype(np.float32) * np.float32(1.4629180792671596e-9)).astype(np.float64)
    out = np.stack([r[:, 0] * np.cos(th[:, 0]), r[:, 0] * np.sin(th[:, 0]), r[:, 1] * np.cos(th[:, 1]), r[:, 1] * np.sin(th[:, 1])], -1)
    return out.reshape(-1)[:n]
