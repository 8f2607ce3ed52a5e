// One-shot gradient all-reduce over NVLink peer memory (the only exchange of the data-parallel path).
//
// The reference has no multi-GPU path; rlpyt's synchronous runners would average gradients with an NCCL all-reduce
// (torch DistributedDataParallel) after the three backwards of dreamer_algo.py:174-180.  Here the actor and value weight
// gradients of one imagination step are written by the backward kernels into ONE flat fp32 bucket that lives in
// symmetric (peer-mapped) memory, and this kernel averages it in place: every GPU reads the bucket of every peer through
// NVLink / NVSwitch P2P loads, sums in rank order (so all ranks end up with bit-identical gradients) and writes the mean
// back.  At ~1 MB the cost of a collective is hand-shake latency, not bandwidth: two flag exchanges and one pass of
// 16-byte loads instead of a ring / tree protocol.
//
//   barrier A   "my bucket is complete" (it is: stream order) -> every peer; wait for every peer
//   reduce      slice of this CTA from all peers -> local staging buffer (scaled by 1 / world)
//   barrier B   "I have finished reading your bucket" -> every peer; wait (only then may a bucket be overwritten)
//   copy        staging -> local bucket
// (2 or 3 ranks; from 4 ranks on the two-shot form below moves less data.)
//
// Flags are epoch numbers (monotonic, never reset): slot [phase][cta][source rank] of the destination rank's signal pad.
// The epoch lives on the device and is advanced by the last CTA to finish, so the call can be recorded into a CUDA graph.
#include "drm_common.cuh"

namespace drm {

constexpr int kArThreads = 512;

__device__ __forceinline__ void st_release_sys(unsigned int* p, unsigned int v) {
    asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ unsigned int ld_acquire_sys(const unsigned int* p) {
    unsigned int v;
    asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ float4 ld_peer4(const float* p) {
    float4 v;
    asm volatile("ld.volatile.global.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p) : "memory");
    return v;
}

// All ranks' CTA `blockIdx.x` meet: thread r signals rank r and waits for rank r's signal.
__device__ __forceinline__ void peer_barrier(const drm_allreduce_call& c, int phase, unsigned int epoch) {
    __syncthreads();  // everything this CTA did before (its peer reads, for barrier B) is ordered before the signal
    const int t = threadIdx.x;
    if (t < c.world) {
        const size_t slot = ((size_t)phase * gridDim.x + blockIdx.x) * DRM_MAX_RANKS;
        __threadfence_system();
        st_release_sys(reinterpret_cast<unsigned int*>(c.signal[t]) + slot + c.rank, epoch);
        const unsigned int* mine = reinterpret_cast<const unsigned int*>(c.signal[c.rank]) + slot + t;
        unsigned long long t0 = 0;
        while ((int)(ld_acquire_sys(mine) - epoch) < 0) {
            __nanosleep(40);
            unsigned long long now;
            asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(now));
            if (t0 == 0) t0 = now;
            else if (now - t0 > 60000000000ull) __trap();  // 60 s: a peer never launched its half of the collective
        }
    }
    __syncthreads();
}

__global__ void __launch_bounds__(kArThreads) allreduce_oneshot_kernel(const drm_allreduce_call c) {
    const unsigned int epoch = (unsigned int)(c.epoch[0] + 1ull);
    peer_barrier(c, 0, epoch);
    const long long n4 = c.numel >> 2;
    const float* src[DRM_MAX_RANKS];
#pragma unroll
    for (int r = 0; r < DRM_MAX_RANKS; ++r) src[r] = reinterpret_cast<const float*>(c.buf[r < c.world ? r : 0]);
    float4* stage = reinterpret_cast<float4*>(c.staging);
    for (long long i = (long long)blockIdx.x * kArThreads + threadIdx.x; i < n4; i += (long long)gridDim.x * kArThreads) {
        float4 v[DRM_MAX_RANKS];
#pragma unroll
        for (int r = 0; r < DRM_MAX_RANKS; ++r)
            if (r < c.world) v[r] = ld_peer4(src[r] + 4 * i);  // all loads in flight before the first add
        float4 a = v[0];
#pragma unroll
        for (int r = 1; r < DRM_MAX_RANKS; ++r)
            if (r < c.world) { a.x += v[r].x; a.y += v[r].y; a.z += v[r].z; a.w += v[r].w; }  // rank order on every rank
        a.x *= c.scale; a.y *= c.scale; a.z *= c.scale; a.w *= c.scale;
        stage[i] = a;
    }
    peer_barrier(c, 1, epoch);
    float4* mine = reinterpret_cast<float4*>(c.buf[c.rank]);
    for (long long i = (long long)blockIdx.x * kArThreads + threadIdx.x; i < n4; i += (long long)gridDim.x * kArThreads)
        mine[i] = stage[i];  // written and read by the same thread
    // the last CTA to finish advances the epoch (every CTA read it when it started)
    __syncthreads();
    if (threadIdx.x == 0) {
        __threadfence();
        const unsigned long long t = atomicAdd(c.epoch + 1, 1ull);
        if (t == (unsigned long long)gridDim.x - 1ull) {
            c.epoch[1] = 0ull;
            c.epoch[0] = c.epoch[0] + 1ull;
        }
    }
}

// Two-shot form for world >= 4 (one-shot makes every GPU read world x the bucket: 8 MB per rank at 8 GPUs, ~27 us of
// NVLink time for a 1 MB bucket; here a rank reads 2 x (world - 1) / world of it):
//   barrier A;  rank r reduces slice r of every peer's bucket into slice r of its OWN bucket (rank order, scaled);
//   barrier B;  rank r copies slice s of rank s's bucket (the reduced one) into its own slice s, for every s != r;
//   barrier C   (nobody overwrites a reduced slice -- the next step's gradients -- while a peer still reads it).
// Slice r of rank p's bucket is read only by rank r in the first phase and written only by rank p in the second, after
// barrier B, so no staging buffer is needed.  Every slice is reduced exactly once: results are bit-identical everywhere.
__global__ void __launch_bounds__(kArThreads) allreduce_twoshot_kernel(const drm_allreduce_call c) {
    const unsigned int epoch = (unsigned int)(c.epoch[0] + 1ull);
    peer_barrier(c, 0, epoch);
    const long long n4 = c.numel >> 2;
    const long long per = (n4 + c.world - 1) / c.world;
    const float* src[DRM_MAX_RANKS];
#pragma unroll
    for (int r = 0; r < DRM_MAX_RANKS; ++r) src[r] = reinterpret_cast<const float*>(c.buf[r < c.world ? r : 0]);
    float4* mine = reinterpret_cast<float4*>(c.buf[c.rank]);
    {
        const long long lo = per * c.rank, hi = lo + per < n4 ? lo + per : n4;
        for (long long i = lo + (long long)blockIdx.x * kArThreads + threadIdx.x; i < hi; i += (long long)gridDim.x * kArThreads) {
            float4 v[DRM_MAX_RANKS];
#pragma unroll
            for (int r = 0; r < DRM_MAX_RANKS; ++r)
                if (r < c.world) v[r] = ld_peer4(src[r] + 4 * i);
            float4 a = v[0];
#pragma unroll
            for (int r = 1; r < DRM_MAX_RANKS; ++r)
                if (r < c.world) { a.x += v[r].x; a.y += v[r].y; a.z += v[r].z; a.w += v[r].w; }
            a.x *= c.scale; a.y *= c.scale; a.z *= c.scale; a.w *= c.scale;
            mine[i] = a;
        }
    }
    __threadfence_system();
    peer_barrier(c, 1, epoch);
    for (int k = 1; k < c.world; ++k) {
        const int s = (c.rank + k) % c.world;  // staggered: the ranks do not all pull from the same peer at once
        const long long lo = per * s, hi = lo + per < n4 ? lo + per : n4;
        for (long long i = lo + (long long)blockIdx.x * kArThreads + threadIdx.x; i < hi; i += (long long)gridDim.x * kArThreads)
            mine[i] = ld_peer4(src[s] + 4 * i);
    }
    peer_barrier(c, 2, epoch);
    __syncthreads();
    if (threadIdx.x == 0) {
        __threadfence();
        const unsigned long long t = atomicAdd(c.epoch + 1, 1ull);
        if (t == (unsigned long long)gridDim.x - 1ull) {
            c.epoch[1] = 0ull;
            c.epoch[0] = c.epoch[0] + 1ull;
        }
    }
}

void warm_allreduce() { warm_kernel(allreduce_oneshot_kernel); warm_kernel(allreduce_twoshot_kernel); }

}  // namespace drm

using namespace drm;

extern "C" int drm_allreduce(const drm_allreduce_call* c, void* stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    DRM_REQUIRE(c != nullptr, DRM_E_ARG, "null call");
    DRM_REQUIRE(c->world >= 2 && c->world <= DRM_MAX_RANKS && c->rank >= 0 && c->rank < c->world, DRM_E_ARG,
                "2..8 ranks, rank inside the world");
    DRM_REQUIRE(c->numel > 0 && c->numel % 4 == 0, DRM_E_SHAPE, "bucket length must be a positive multiple of 4 floats");
    DRM_REQUIRE(c->staging && c->epoch, DRM_E_ARG, "null staging buffer / epoch counter");
    DRM_REQUIRE(c->ctas >= 1 && c->ctas <= DRM_AR_MAX_CTAS, DRM_E_ARG, "1..64 CTAs");
    for (int r = 0; r < c->world; ++r) {
        DRM_REQUIRE(c->buf[r] && c->signal[r], DRM_E_ARG, "null peer pointer");
        DRM_REQUIRE(((size_t)c->buf[r] & 15) == 0, DRM_E_ALIGN, "bucket must be 16-byte aligned");
    }
    int rc = check_device(); if (rc) return rc;
    ProfScope prof(PROF_LOSS, stream);
    DRM_REQUIRE(c->algo >= 0 && c->algo <= 2 && (c->algo != 2 || c->world >= 3), DRM_E_ARG, "bad algo");
    const bool two_shot = c->algo == 2 || (c->algo == 0 && c->world >= 4 && c->numel >= (1 << 20));
    if (two_shot) allreduce_twoshot_kernel<<<c->ctas, kArThreads, 0, stream>>>(*c);
    else allreduce_oneshot_kernel<<<c->ctas, kArThreads, 0, stream>>>(*c);
    DRM_CUDA_OK(cudaGetLastError());
    return DRM_OK;
}
