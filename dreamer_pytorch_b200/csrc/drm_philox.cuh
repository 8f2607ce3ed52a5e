// Counter-based noise for the production path (last sentence): N(0,1) / Gumbel(0,1) values keyed on
// (seed, offset, stream, element index) with Philox4x32-10 (Salmon et al., "Parallel random numbers: as easy as 1, 2, 3",
// SC'11; the generator behind torch.cuda's and cuRAND's default engines).  A value depends on nothing but its key, so the
// fp32 rollout kernels generate the noise of an element where they would have loaded it (forward and, again, backward --
// nothing goes through HBM), and drm_noise_fill materialises the very same values for the kernels that want tensors.
//
// Key layout:  counter = { e / 4 (low 32 bits), e / 4 (high 24 bits) | stream << 24, offset low, offset high }
//              key     = { seed low, seed high }
// e = row-major element index of the noise tensor; element e takes output word pair (e & 3).  Normal: Box-Muller on
// words (0,1) -> elements 0,1 and words (2,3) -> elements 2,3.  Gumbel: -ln(-ln u), u from the element's own word.
#pragma once
#include <stdint.h>

namespace drm {

enum NoiseStream { NZ_EPS_PRIOR = 0, NZ_EPS_POST = 1, NZ_NOISE_ACT = 2, NZ_MODE = 3 };
enum NoiseKind { NZ_NORMAL = 0, NZ_GUMBEL = 1 };

struct NoiseKey {
    uint32_t seed_lo, seed_hi, off_lo, off_hi;
};

__device__ __forceinline__ NoiseKey load_noise_key(const unsigned long long* pair) {
    const unsigned long long s = pair[0], o = pair[1];
    return NoiseKey{(uint32_t)s, (uint32_t)(s >> 32), (uint32_t)o, (uint32_t)(o >> 32)};
}

__device__ __forceinline__ void philox4x32_10(uint32_t (&c)[4], uint32_t k0, uint32_t k1) {
#pragma unroll
    for (int r = 0; r < 10; ++r) {
        const uint32_t hi0 = __umulhi(0xD2511F53u, c[0]), lo0 = 0xD2511F53u * c[0];
        const uint32_t hi1 = __umulhi(0xCD9E8D57u, c[2]), lo1 = 0xCD9E8D57u * c[2];
        c[0] = hi1 ^ c[1] ^ k0;
        c[1] = lo1;
        c[2] = hi0 ^ c[3] ^ k1;
        c[3] = lo0;
        k0 += 0x9E3779B9u;
        k1 += 0xBB67AE85u;
    }
}

// The four raw words of group g (= element index / 4) of a stream.
__device__ __forceinline__ void noise_words(const NoiseKey& k, int stream, unsigned long long g, uint32_t (&w)[4]) {
    w[0] = (uint32_t)g;
    w[1] = ((uint32_t)(g >> 32) & 0x00FFFFFFu) | ((uint32_t)stream << 24);
    w[2] = k.off_lo;
    w[3] = k.off_hi;
    philox4x32_10(w, k.seed_lo, k.seed_hi);
}

// (0, 1]: never 0, so the logarithm below is finite
__device__ __forceinline__ float unit_open_low(uint32_t x) { return __fmaf_rn((float)x, 2.3283064365386963e-10f, 2.3283064365386963e-10f); }

// One element of a stream.  `sel` = element index & 3.
__device__ __forceinline__ float noise_from_words(const uint32_t (&w)[4], int sel, int kind) {
    if (kind == NZ_GUMBEL) {
        const uint32_t x = sel == 0 ? w[0] : sel == 1 ? w[1] : sel == 2 ? w[2] : w[3];
        // u in (0, 1): the upper end is excluded by clamping, like the reference-side clamp of torch.rand
        const float u = fminf(unit_open_low(x), 0.99999994f);
        return -logf(-logf(u));
    }
    const uint32_t a = sel < 2 ? w[0] : w[2], b = sel < 2 ? w[1] : w[3];
    const float r = sqrtf(-2.f * logf(unit_open_low(a)));
    const float th = __fmul_rn((float)b, 1.4629180792671596e-9f);  // 2 pi / 2^32
    float s, c;
    sincosf(th, &s, &c);
    return __fmul_rn(r, (sel & 1) ? s : c);
}

__device__ __forceinline__ float noise_element(const NoiseKey& k, int stream, int kind, unsigned long long e) {
    uint32_t w[4];
    noise_words(k, stream, e >> 2, w);
    return noise_from_words(w, (int)(e & 3), kind);
}

// What the rollout kernels call where they used to read p.eps_prior[o] etc.: the tensor when one was given (injected
// noise of the parity tests, or a materialised production tensor), the generator otherwise.
__device__ __forceinline__ float noise_at(const float* __restrict__ tensor, const NoiseKey& k, int stream, int kind,
                                          unsigned long long e) {
    return tensor ? __ldg(tensor + e) : noise_element(k, stream, kind, e);
}

}  // namespace drm
