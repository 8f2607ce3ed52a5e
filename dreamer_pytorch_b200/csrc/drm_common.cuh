// Shared device/host helpers for libdreamer_b200 (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include "../../include/dreamer_b200.h"

namespace drm {

constexpr int kThreads = 256;          // 8 warps per CTA for the row-tile kernels
constexpr int kWarps = kThreads / 32;
int num_sms();                         // SMs of the current device (cudaDevAttrMultiProcessorCount, cached per device; 148 on B200)
constexpr int kMaxSmem = 227 * 1024;   // usable dynamic shared memory per CTA

__host__ __device__ inline int rup(int x, int m) { return (x + m - 1) / m * m; }
__host__ __device__ inline size_t rup_sz(size_t x, size_t m) { return (x + m - 1) / m * m; }

void set_error(const char* fmt, ...);
int check_device();  // DRM_OK or DRM_E_ARCH / DRM_E_CUDA
// Forces the CUDA module load of one kernel (lazy loading would otherwise do it at the first launch, which may have to
// wait for the device to go idle -- something a kernel spinning on device counters never lets happen).
template <class K>
inline void warm_kernel(K kernel) {
    cudaFuncAttributes a;
    (void)cudaFuncGetAttributes(&a, kernel);
}
// per translation unit: load every kernel it defines (drm_init)
void warm_util(); void warm_mlp(); void warm_mlp_tc(); void warm_wgrad_tc(); void warm_rollout_fwd(); void warm_rollout_bwd();
void warm_rollout_tc(); void warm_loss(); void warm_optim(); void warm_policy(); void warm_replay(); void warm_noise(); void warm_allreduce(); void warm_conv_glue();

// Diagnostics (drm_launch_count / drm_prof_*): every kernel launch site bumps the counter and,
// when profiling is on, is bracketed by a cudaEvent pair on the launching stream.
enum ProfSlot { PROF_PACK = 0, PROF_ROLLOUT_FWD, PROF_ROLLOUT_BWD, PROF_MLP_FWD, PROF_MLP_BWD, PROF_WGRAD,
                PROF_LOSS, PROF_N };
struct ProfScope {
    int slot; cudaStream_t stream; int idx;
    ProfScope(int slot, cudaStream_t stream);
    ~ProfScope();
};

#define DRM_CUDA_OK(expr)                                                                    \
    do {                                                                                     \
        cudaError_t _e = (expr);                                                             \
        if (_e != cudaSuccess) {                                                             \
            drm::set_error("%s:%d %s -> %s", __FILE__, __LINE__, #expr, cudaGetErrorString(_e)); \
            return DRM_E_CUDA;                                                               \
        }                                                                                    \
    } while (0)

#define DRM_REQUIRE(cond, code, msg)                                     \
    do {                                                                 \
        if (!(cond)) {                                                   \
            drm::set_error("%s:%d %s (%s)", __FILE__, __LINE__, msg, #cond); \
            return code;                                                 \
        }                                                                \
    } while (0)

// ------------------------------------------------------------------------------------------
// pointwise math (fp32, matches torch's formulas)
// ------------------------------------------------------------------------------------------
__device__ __forceinline__ float elu(float x) { return x > 0.f ? x : expm1f(x); }
// derivative of ELU expressed through its OUTPUT y: 1 for y>0, y+1 (= e^x) otherwise
__device__ __forceinline__ float elu_grad_from_out(float y) { return y > 0.f ? 1.f : y + 1.f; }
__device__ __forceinline__ float sigmoidf_(float x) { return 1.f / (1.f + expf(-x)); }
// torch softplus (beta=1, threshold=20)
__device__ __forceinline__ float softplusf_(float x) { return x > 20.f ? x : log1pf(expf(x)); }

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}

// ------------------------------------------------------------------------------------------
// Weight packing: every Linear is re-laid once per call into a zero-padded matrix whose
// reduction index is the ROW (padded to 4) and whose output index is the COLUMN (leading
// dimension padded to 32), so the row-tile GEMM below streams it with coalesced, unguarded loads.
//   transpose=1: dst[k][n] = src[(r0+n)*src_ld + c0 + k]   (forward:  y = x W^T, reduce over in-features)
//   transpose=0: dst[n][k] = src[(r0+n)*src_ld + c0 + k]   (backward: dx = dy W, reduce over out-features)
// ------------------------------------------------------------------------------------------
struct PackJob {
    const float* src;
    float* dst;
    int src_ld;
    int r0, c0;      // origin of the block inside src
    int nr, nc;      // block size: nr rows (out-features) x nc cols (in-features)
    int transpose;
    int dst_rows;    // padded reduction extent (multiple of 4)
    int dst_ld;      // padded output extent (multiple of 32)
};
constexpr int kMaxPackJobs = 40;
struct PackJobs {
    PackJob j[kMaxPackJobs];
    int n;
};

// Appends a job and returns the number of floats it occupies in the destination.
inline size_t pack_size(int red, int out) { return (size_t)rup(red, 4) * (size_t)rup(out, 32); }

struct Packer {
    PackJobs jobs;
    float* base;
    size_t off;  // floats
    Packer() : base(nullptr), off(0) { jobs.n = 0; }
    explicit Packer(float* b) : base(b), off(0) { jobs.n = 0; }
    // forward layout of block rows [r0,r0+nr) x cols [c0,c0+nc) of src: reduce over cols
    const float* fwd(const float* src, int src_ld, int r0, int nr, int c0, int nc) {
        return add(src, src_ld, r0, nr, c0, nc, 1, rup(nc, 4), rup(nr, 32));
    }
    // backward layout: reduce over rows
    const float* bwd(const float* src, int src_ld, int r0, int nr, int c0, int nc) {
        return add(src, src_ld, r0, nr, c0, nc, 0, rup(nr, 4), rup(nc, 32));
    }
    const float* add(const float* src, int src_ld, int r0, int nr, int c0, int nc, int tr, int drows, int dld) {
        float* dst = base ? base + off : nullptr;
        if (base && jobs.n < kMaxPackJobs) {
            PackJob& j = jobs.j[jobs.n];
            j.src = src; j.dst = dst; j.src_ld = src_ld; j.r0 = r0; j.c0 = c0; j.nr = nr; j.nc = nc;
            j.transpose = tr; j.dst_rows = drows; j.dst_ld = dld;
        }
        jobs.n++;
        off += (size_t)drows * dld;
        return dst;
    }
};
int launch_pack(const PackJobs& jobs, cudaStream_t stream);

// ------------------------------------------------------------------------------------------
// Row-tile GEMM primitive.  A CTA owns R rows (R compile-time) whose inputs sit in shared
// memory; each warp produces 32 output columns at a time for ALL R rows, streaming the packed
// weight matrix (L2-resident) exactly once per CTA.  Up to two K-segments accumulate into the
// same outputs (e.g. [action | stoch], [x | h]).
// ------------------------------------------------------------------------------------------
struct Seg {
    const float* W;  // packed [K4][ldw]
    const float* X;  // shared memory [R][ldx], ldx % 4 == 0, 16-byte aligned
    int K4;          // multiple of 4; 0 = unused segment
    int ldx;
};

template <int R>
__device__ __forceinline__ void seg_accumulate(float (&acc)[R], const Seg& s, int ldw, int col) {
    const float* __restrict__ wp = s.W + col;
    const float* __restrict__ xp = s.X;
    const int ldx = s.ldx;
#pragma unroll 2
    for (int k = 0; k < s.K4; k += 4) {
        const float w0 = __ldg(wp + (size_t)(k + 0) * ldw);
        const float w1 = __ldg(wp + (size_t)(k + 1) * ldw);
        const float w2 = __ldg(wp + (size_t)(k + 2) * ldw);
        const float w3 = __ldg(wp + (size_t)(k + 3) * ldw);
#pragma unroll
        for (int r = 0; r < R; ++r) {
            const float4 x = *reinterpret_cast<const float4*>(xp + r * ldx + k);
            acc[r] = fmaf(x.x, w0, acc[r]);
            acc[r] = fmaf(x.y, w1, acc[r]);
            acc[r] = fmaf(x.z, w2, acc[r]);
            acc[r] = fmaf(x.w, w3, acc[r]);
        }
    }
}

// epi(r, col, value) is called for every r < R and col < nout.
// Few rows per CTA (R <= 8; the observation rollout has 50 rows in all, 4 per CTA): the FMAs are nothing, the kernel is bound
// by the latency of its weight loads, and 8 warps (2 per scheduler) cannot hide it (ncu: 24 % issue slots used, 3.3 warps
// stalled on the long scoreboard per issue).  Such CTAs are launched with kWideThreads threads and split K over the warps
// that would otherwise idle: warp = (K part, column block); partial sums meet in a shared-memory scratch
// [warp][R][32] and are reduced, bias / activation applied, by all threads.  Same epi(r, col, value) contract.
constexpr int kWideThreads = 1024;
template <int R, class Epi>
__device__ __forceinline__ void cta_linear_splitk(const Seg& s0, const Seg& s1, int ldw, int nout, Epi epi, float* scratch) {
    const int nwarps = blockDim.x >> 5;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int CB = (nout + 31) >> 5;                       // column blocks of 32
    const int ksplit = CB >= nwarps ? 1 : nwarps / CB;     // K parts
    __syncthreads();                                       // scratch may still be read by the previous call's reduction
    for (int slot = warp; slot < CB * ksplit; slot += nwarps) {
        const int cb = slot % CB, ks = slot / CB;
        const int col = cb * 32 + lane;
        float acc[R];
#pragma unroll
        for (int r = 0; r < R; ++r) acc[r] = 0.f;
        const int Ktot = s0.K4 + s1.K4;
        int kbeg = (int)(((long long)Ktot * ks / ksplit) & ~3LL), kend = (int)(((long long)Ktot * (ks + 1) / ksplit) & ~3LL);
        if (ks == ksplit - 1) kend = Ktot;
        // [kbeg, kend) over the concatenation of the two segments
        if (kbeg < s0.K4) {
            Seg p = s0;
            const int e = kend < s0.K4 ? kend : s0.K4;
            p.W = s0.W + (size_t)kbeg * ldw; p.X = s0.X + kbeg; p.K4 = e - kbeg;
            seg_accumulate<R>(acc, p, ldw, col);
        }
        if (kend > s0.K4 && s1.K4 > 0) {
            Seg p = s1;
            const int b = kbeg > s0.K4 ? kbeg - s0.K4 : 0;
            p.W = s1.W + (size_t)b * ldw; p.X = s1.X + b; p.K4 = kend - s0.K4 - b;
            seg_accumulate<R>(acc, p, ldw, col);
        }
        float* dst = scratch + (size_t)slot * R * 32 + lane;
#pragma unroll
        for (int r = 0; r < R; ++r) dst[r * 32] = acc[r];
    }
    __syncthreads();
    for (int idx = threadIdx.x; idx < R * CB * 32; idx += blockDim.x) {
        const int r = idx / (CB * 32), c = idx - r * (CB * 32);
        if (c >= nout) continue;
        const int cb = c >> 5, ln = c & 31;
        float v = 0.f;
        for (int ks = 0; ks < ksplit; ++ks) v += scratch[((size_t)(ks * CB + cb) * R + r) * 32 + ln];
        epi(r, c, v);
    }
}
// Very few rows per CTA (R <= 4; the observation rollout runs ONE row per CTA): the 32-column form above keeps at most
// 8 x 128 B of weights in flight per warp and was bound by L2 latency (ncu: 59 % of issue slots, long-scoreboard stalls).
// Here a lane owns FOUR adjacent columns (one 16-byte load per k), a warp a 128-column block, and K is split over all the
// other warps (200 outputs: 2 column blocks x 16 K parts of ~12 k each), so that every weight byte of a layer is requested
// within the first few instructions of the call and the layer costs about one L2 round trip.  Scratch [warp][R][128].
template <int R>
__device__ __forceinline__ void seg_accumulate4(float (&acc)[R][4], const Seg& s, int ldw, int col) {
    const float* __restrict__ wp = s.W + col;
    const float* __restrict__ xp = s.X;
    const int ldx = s.ldx;
    if constexpr (R >= 12 && R <= 16) {  // (20 rows: the prefetch registers spill)
        // Mid-height tiles run long K ranges per warp (a 600-wide layer leaves no room to split K): the weights of step
        // k + 4 are requested before the FFMAs of step k, so one L2 round trip hides under 12-20 rows of math instead of
        // preceding it.
        if (s.K4 <= 0) return;
        float4 w0 = __ldg(reinterpret_cast<const float4*>(wp + (size_t)0 * ldw));
        float4 w1 = __ldg(reinterpret_cast<const float4*>(wp + (size_t)1 * ldw));
        float4 w2 = __ldg(reinterpret_cast<const float4*>(wp + (size_t)2 * ldw));
        float4 w3 = __ldg(reinterpret_cast<const float4*>(wp + (size_t)3 * ldw));
#pragma unroll 1
        for (int k = 0; k < s.K4; k += 4) {
            float4 n0 = w0, n1 = w1, n2 = w2, n3 = w3;
            if (k + 4 < s.K4) {
                n0 = __ldg(reinterpret_cast<const float4*>(wp + (size_t)(k + 4) * ldw));
                n1 = __ldg(reinterpret_cast<const float4*>(wp + (size_t)(k + 5) * ldw));
                n2 = __ldg(reinterpret_cast<const float4*>(wp + (size_t)(k + 6) * ldw));
                n3 = __ldg(reinterpret_cast<const float4*>(wp + (size_t)(k + 7) * ldw));
            }
#pragma unroll
            for (int r = 0; r < R; ++r) {
                const float4 x = *reinterpret_cast<const float4*>(xp + r * ldx + k);
                acc[r][0] = fmaf(x.w, w3.x, fmaf(x.z, w2.x, fmaf(x.y, w1.x, fmaf(x.x, w0.x, acc[r][0]))));
                acc[r][1] = fmaf(x.w, w3.y, fmaf(x.z, w2.y, fmaf(x.y, w1.y, fmaf(x.x, w0.y, acc[r][1]))));
                acc[r][2] = fmaf(x.w, w3.z, fmaf(x.z, w2.z, fmaf(x.y, w1.z, fmaf(x.x, w0.z, acc[r][2]))));
                acc[r][3] = fmaf(x.w, w3.w, fmaf(x.z, w2.w, fmaf(x.y, w1.w, fmaf(x.x, w0.w, acc[r][3]))));
            }
            w0 = n0; w1 = n1; w2 = n2; w3 = n3;
        }
        return;
    }
#pragma unroll(R <= 2 ? 3 : 1)
    for (int k = 0; k < s.K4; k += 4) {
        const float4 w0 = __ldg(reinterpret_cast<const float4*>(wp + (size_t)(k + 0) * ldw));
        const float4 w1 = __ldg(reinterpret_cast<const float4*>(wp + (size_t)(k + 1) * ldw));
        const float4 w2 = __ldg(reinterpret_cast<const float4*>(wp + (size_t)(k + 2) * ldw));
        const float4 w3 = __ldg(reinterpret_cast<const float4*>(wp + (size_t)(k + 3) * ldw));
#pragma unroll
        for (int r = 0; r < R; ++r) {
            const float4 x = *reinterpret_cast<const float4*>(xp + r * ldx + k);
            acc[r][0] = fmaf(x.w, w3.x, fmaf(x.z, w2.x, fmaf(x.y, w1.x, fmaf(x.x, w0.x, acc[r][0]))));
            acc[r][1] = fmaf(x.w, w3.y, fmaf(x.z, w2.y, fmaf(x.y, w1.y, fmaf(x.x, w0.y, acc[r][1]))));
            acc[r][2] = fmaf(x.w, w3.z, fmaf(x.z, w2.z, fmaf(x.y, w1.z, fmaf(x.x, w0.z, acc[r][2]))));
            acc[r][3] = fmaf(x.w, w3.w, fmaf(x.z, w2.w, fmaf(x.y, w1.w, fmaf(x.x, w0.w, acc[r][3]))));
        }
    }
}

// Scratch handed to cta_linear by the wide kernels: pointer + size (the K split adapts to what fits).
struct KScratch {
    float* p;
    int floats;
};
// Threads per CTA of the rollout kernels by tile height: <= 8 rows 1 024 (split K over the idle warps), 12..20 rows 512
// (same, float4 form; 128 registers per thread hold the [R][4] accumulators), more rows 256 (classic form).
__host__ __device__ constexpr int rollout_threads(int R) { return R <= 8 ? 1024 : R <= 20 ? 512 : kThreads; }

template <int R, class Epi>
__device__ __forceinline__ void cta_linear_splitk4(const Seg& s0, const Seg& s1, int ldw, int nout, Epi epi, KScratch sc) {
    const int nwarps = blockDim.x >> 5;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int CB = (nout + 127) >> 7;                    // column blocks of 128
    const int per = CB < nwarps ? CB : nwarps;           // column blocks per round (one round unless nout > 128 x warps)
    int ksplit = nwarps / per;                           // K parts: as many as there are warps ...
    const int fit = sc.p ? sc.floats / (per * R * 128) : 0;  // ... and scratch slots [K part][column block][R][128]
    if (ksplit > fit) ksplit = fit;
    if (ksplit < 2) ksplit = 1;
    const int Ktot = s0.K4 + s1.K4;
    for (int cb0 = 0; cb0 < CB; cb0 += per) {
        if (ksplit > 1) __syncthreads();                 // scratch may still be read by the previous reduction
        const int cbl = warp % per, ks = warp / per;
        const int col = (cb0 + cbl) * 128 + lane * 4;
        if (ks < ksplit && cb0 + cbl < CB) {
            float acc[R][4];
#pragma unroll
            for (int r = 0; r < R; ++r) { acc[r][0] = acc[r][1] = acc[r][2] = acc[r][3] = 0.f; }
            if (col < ldw) {                             // ldw is a multiple of 32: a float4 at col < ldw stays inside the row
                int kbeg = (int)(((long long)Ktot * ks / ksplit) & ~3LL), kend = (int)(((long long)Ktot * (ks + 1) / ksplit) & ~3LL);
                if (ks == ksplit - 1) kend = Ktot;
                if (kbeg < s0.K4) {
                    Seg p = s0;
                    const int e = kend < s0.K4 ? kend : s0.K4;
                    p.W = s0.W + (size_t)kbeg * ldw; p.X = s0.X + kbeg; p.K4 = e - kbeg;
                    seg_accumulate4<R>(acc, p, ldw, col);
                }
                if (kend > s0.K4 && s1.K4 > 0) {
                    Seg p = s1;
                    const int b = kbeg > s0.K4 ? kbeg - s0.K4 : 0;
                    p.W = s1.W + (size_t)b * ldw; p.X = s1.X + b; p.K4 = kend - s0.K4 - b;
                    seg_accumulate4<R>(acc, p, ldw, col);
                }
            }
            if (ksplit == 1) {                           // whole dot products in registers: no scratch, no barriers
#pragma unroll
                for (int r = 0; r < R; ++r) {
#pragma unroll
                    for (int j = 0; j < 4; ++j)
                        if (col + j < nout) epi(r, col + j, acc[r][j]);
                }
            } else {
                float* dst = sc.p + (size_t)(ks * per + cbl) * R * 128 + lane * 4;
#pragma unroll
                for (int r = 0; r < R; ++r)
                    *reinterpret_cast<float4*>(dst + r * 128) = make_float4(acc[r][0], acc[r][1], acc[r][2], acc[r][3]);
            }
        }
        if (ksplit == 1) continue;
        __syncthreads();
        const int ncol = (CB - cb0 < per ? CB - cb0 : per) * 128;
        for (int idx = threadIdx.x; idx < R * ncol; idx += blockDim.x) {
            const int r = idx / ncol, cl = idx - r * ncol;
            const int c = cb0 * 128 + cl;
            if (c >= nout) continue;
            const int cbl2 = cl >> 7, ln = cl & 127;
            float v = 0.f;
            for (int k2 = 0; k2 < ksplit; ++k2) v += sc.p[((size_t)(k2 * per + cbl2) * R + r) * 128 + ln];
            epi(r, c, v);
        }
    }
}
// floats of scratch a wide CTA would like to have (the launch code gives it what is left of shared memory, at most this).
// R <= 4 and 12..20: [warp][R][128] (cta_linear_splitk4); R = 8: one [R][32] block per (K part, column block) slot, at
// most 2 x warps slots (cta_linear_splitk)
__host__ __device__ inline int splitk_scratch_floats(int R) {
    return R == 8 ? 2 * (rollout_threads(R) / 32) * R * 32 : R <= 20 ? (rollout_threads(R) / 32) * R * 128 : 0;
}

template <int R, class Epi>
__device__ __forceinline__ void cta_linear(const Seg& s0, const Seg& s1, int ldw, int nout, Epi epi,
                                           KScratch scratch = KScratch{nullptr, 0}) {
    if constexpr (R <= 20) {
        if (blockDim.x > kThreads) {  // wide CTA of a rollout kernel
            if constexpr (R == 8) {
                if (scratch.p != nullptr) { cta_linear_splitk<R>(s0, s1, ldw, nout, epi, scratch.p); return; }
            } else {
                cta_linear_splitk4<R>(s0, s1, ldw, nout, epi, scratch);
                return;
            }
        }
    }
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int nwarps = blockDim.x >> 5;
    for (int c0 = warp * 32; c0 < nout; c0 += nwarps * 32) {
        const int col = c0 + lane;  // < ldw always (ldw is nout rounded up to 32)
        float acc[R];
#pragma unroll
        for (int r = 0; r < R; ++r) acc[r] = 0.f;
        seg_accumulate<R>(acc, s0, ldw, col);
        if (s1.K4 > 0) seg_accumulate<R>(acc, s1, ldw, col);
        if (col < nout) {
#pragma unroll
            for (int r = 0; r < R; ++r) epi(r, col, acc[r]);
        }
    }
}

// Narrow outputs (nout < 32): one (row, col) dot product per warp iteration, lanes split K.
// Here W is the ORIGINAL row-major [nout, ldw_orig] matrix block (k contiguous), not packed.
template <class Epi>
__device__ __forceinline__ void cta_linear_narrow(const float* __restrict__ W, int w_ld, int K,
                                                  const float* X, int ldx, int nrows, int nout, Epi epi) {
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    for (int idx = warp; idx < nrows * nout; idx += (int)(blockDim.x >> 5)) {
        const int r = idx / nout, c = idx - r * nout;
        const float* wr = W + (size_t)c * w_ld;
        const float* xr = X + r * ldx;
        float s = 0.f;
        for (int k = lane; k < K; k += 32) s = fmaf(xr[k], __ldg(wr + k), s);
        s = warp_sum(s);
        if (lane == 0) epi(r, c, s);
    }
}

// Rows-per-CTA dispatch: instantiate kernels for these tile heights only.
#define DRM_DISPATCH_R(Rv, ...)                                \
    switch (Rv) {                                              \
        case 1:  { constexpr int R = 1;  __VA_ARGS__; } break; \
        case 2:  { constexpr int R = 2;  __VA_ARGS__; } break; \
        case 4:  { constexpr int R = 4;  __VA_ARGS__; } break; \
        case 8:  { constexpr int R = 8;  __VA_ARGS__; } break; \
        case 12: { constexpr int R = 12; __VA_ARGS__; } break; \
        case 16: { constexpr int R = 16; __VA_ARGS__; } break; \
        case 20: { constexpr int R = 20; __VA_ARGS__; } break; \
        case 24: { constexpr int R = 24; __VA_ARGS__; } break; \
        case 32: { constexpr int R = 32; __VA_ARGS__; } break; \
        default: drm::set_error("bad tile height %d", Rv); return DRM_E_SHAPE; \
    }

// Picks the tile height: the smallest supported R that covers ceil(n / num_sms()) rows (so one
// wave of CTAs finishes the job), capped by what fits in shared memory.
inline int pick_rows(int n, size_t smem_bytes_per_row, size_t smem_fixed) {
    static const int opts[] = {1, 2, 4, 8, 12, 16, 20, 24, 32};
    const int sms = num_sms();
    int want = (n + sms - 1) / sms;
    int best = 0;
    for (int i = 0; i < 9; ++i) {
        int R = opts[i];
        if (smem_fixed + (size_t)R * smem_bytes_per_row > (size_t)kMaxSmem) break;
        best = R;
        if (R >= want) break;
    }
    return best;  // 0: does not fit at all
}

}  // namespace drm
