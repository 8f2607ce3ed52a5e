// Production noise: the (seed, offset) state of the Philox streams of drm_philox.cuh lives on the device, so
// that drawing a fresh offset is a kernel (recordable into a CUDA graph: every replay draws new noise) and never a host
// synchronisation.
//   state  uint64[4] = { seed, offset, ticket, unused }     one per device, owned by the caller
//   pair   uint64[2] = { seed, offset }                     frozen copy a rollout's forward AND backward are keyed on
#include "drm_common.cuh"
#include "drm_philox.cuh"

namespace drm {

__global__ void noise_next_kernel(unsigned long long* state, unsigned long long* pair) {
    if (threadIdx.x == 0 && blockIdx.x == 0) {
        pair[0] = state[0];
        pair[1] = state[1];
        state[1] = state[1] + 1ull;
    }
}

constexpr int kFillThreads = 256;

// Each thread produces whole groups of four elements (one Philox call, two Box-Muller pairs).
__global__ void __launch_bounds__(kFillThreads) noise_fill_kernel(const drm_noise_fill_call c) {
    const unsigned long long* src = c.pair ? c.pair : c.state;
    const NoiseKey key = load_noise_key(src);
    for (int b = 0; b < c.n_blocks; ++b) {
        float* out = c.out[b];
        const unsigned long long n = (unsigned long long)c.numel[b];
        const unsigned long long groups = (n + 3) >> 2;
        const int stream = c.stream_id[b], kind = c.kind[b];
        for (unsigned long long g = (unsigned long long)blockIdx.x * kFillThreads + threadIdx.x; g < groups;
             g += (unsigned long long)gridDim.x * kFillThreads) {
            uint32_t w[4];
            noise_words(key, stream, g, w);
            float v[4];
#pragma unroll
            for (int j = 0; j < 4; ++j) v[j] = noise_from_words(w, j, kind);
            const unsigned long long e = g << 2;
            if (e + 3 < n && (((size_t)out & 15) == 0)) {
                *reinterpret_cast<float4*>(out + e) = make_float4(v[0], v[1], v[2], v[3]);
            } else {
#pragma unroll
                for (int j = 0; j < 4; ++j)
                    if (e + j < n) out[e + j] = v[j];
            }
        }
    }
    if (c.pair == nullptr) {
        // the LAST block to finish bumps the offset: every block read the key when it started, before it could finish
        __shared__ unsigned int last;
        __syncthreads();
        if (threadIdx.x == 0) {
            __threadfence();
            const unsigned long long t = atomicAdd(c.state + 2, 1ull);
            last = (t == (unsigned long long)gridDim.x - 1ull) ? 1u : 0u;
            if (last) {
                c.state[2] = 0ull;
                c.state[1] = c.state[1] + 1ull;
                if (c.pair_out) { c.pair_out[0] = key.seed_lo | ((unsigned long long)key.seed_hi << 32);
                                  c.pair_out[1] = key.off_lo | ((unsigned long long)key.off_hi << 32); }
            }
        }
    }
}

__global__ void noise_words_kernel(const unsigned long long* pair, int stream, unsigned long long g0, int n, uint32_t* out) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    uint32_t w[4];
    noise_words(load_noise_key(pair), stream, g0 + i, w);
    for (int j = 0; j < 4; ++j) out[4 * i + j] = w[j];
}

void warm_noise() { warm_kernel(noise_next_kernel); warm_kernel(noise_fill_kernel); warm_kernel(noise_words_kernel); }

}  // namespace drm

using namespace drm;

extern "C" {

int drm_noise_next(unsigned long long* state, unsigned long long* pair_out, void* stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    DRM_REQUIRE(state && pair_out, DRM_E_ARG, "null noise state");
    int rc = check_device(); if (rc) return rc;
    ProfScope prof(PROF_LOSS, stream);
    noise_next_kernel<<<1, 32, 0, stream>>>(state, pair_out);
    DRM_CUDA_OK(cudaGetLastError());
    return DRM_OK;
}

int drm_noise_fill(const drm_noise_fill_call* c, void* stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    DRM_REQUIRE(c != nullptr, DRM_E_ARG, "null call");
    DRM_REQUIRE((c->state != nullptr) != (c->pair != nullptr), DRM_E_ARG, "exactly one of state / pair");
    DRM_REQUIRE(c->n_blocks >= 1 && c->n_blocks <= DRM_MAX_NOISE_BLOCKS, DRM_E_SHAPE, "1..4 noise blocks per call");
    unsigned long long most = 0;
    for (int b = 0; b < c->n_blocks; ++b) {
        DRM_REQUIRE(c->out[b] && c->numel[b] >= 0, DRM_E_ARG, "null noise block");
        DRM_REQUIRE(c->stream_id[b] >= 0 && c->stream_id[b] < 256, DRM_E_ARG, "stream id out of range");
        DRM_REQUIRE(c->kind[b] == NZ_NORMAL || c->kind[b] == NZ_GUMBEL, DRM_E_ARG, "unknown noise kind");
        if ((unsigned long long)c->numel[b] > most) most = (unsigned long long)c->numel[b];
    }
    int rc = check_device(); if (rc) return rc;
    const unsigned long long groups = (most + 3) >> 2;
    unsigned long long blocks = (groups + kFillThreads - 1) / kFillThreads;
    const unsigned long long cap = (unsigned long long)num_sms() * 8;
    if (blocks > cap) blocks = cap;
    if (blocks < 1) blocks = 1;
    ProfScope prof(PROF_LOSS, stream);
    noise_fill_kernel<<<(unsigned)blocks, kFillThreads, 0, stream>>>(*c);
    DRM_CUDA_OK(cudaGetLastError());
    return DRM_OK;
}

int drm_noise_words(const unsigned long long* pair, int stream_id, unsigned long long group0, int n_groups,
                    unsigned int* out, void* stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    DRM_REQUIRE(pair && out && n_groups >= 0, DRM_E_ARG, "null argument");
    int rc = check_device(); if (rc) return rc;
    if (n_groups == 0) return DRM_OK;
    ProfScope prof(PROF_LOSS, stream);
    noise_words_kernel<<<(n_groups + 127) / 128, 128, 0, stream>>>(pair, stream_id, group0, n_groups, out);
    DRM_CUDA_OK(cudaGetLastError());
    return DRM_OK;
}

}  // extern "C"
