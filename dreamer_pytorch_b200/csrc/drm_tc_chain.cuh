// Warp-specialised "GEMM chain" engine shared by the bf16 tensor-core kernels (batched MLP forward /
// backward, imagination rollout forward / backward).
//
// A CTA owns a tile of 128 rows.  Per repetition (one row tile of an MLP, or one time step of a rollout)
// it executes a fixed PROGRAM of phases; a phase is a group of tcgen05 GEMM ops
//     D[128 x n] (TMEM columns d_col..) (+)= A[128 x K] (shared, chunk-major bf16) * W[n x K]^T
// followed by an epilogue that reads the accumulators and (typically) writes the next A operand.
//
// Roles:   warps 0..EPI_WARPS-1  epilogue (kernel specific code)
//          warp  EPI_WARPS       weight producer: streams the pre-packed bf16 weight images of the whole
//                                program, in consumption order, through a ring of shared-memory stages
//                                with 1-D bulk copies (TMA engine) -- weights stay L2 resident
//          warp  EPI_WARPS+1     MMA issuer (one elected thread)
// Barriers: full[s]/empty[s] per stage (producer <-> MMA), acc_full (MMA -> epilogue, one per phase),
//          a_ready (epilogue -> MMA: "A operands of the next phase are in shared memory and the
//          accumulators of the previous phase have been consumed").
#pragma once
#include "drm_common.cuh"
#include "drm_tc.cuh"

namespace drm {
namespace tc {

constexpr int kTileRows = 128;
constexpr int kEpiWarps = 16;  // 4 per scheduler: the epilogues are latency bound, not issue bound
constexpr int kEpiThreads = kEpiWarps * 32;
constexpr int kChainThreads = kEpiThreads + 64;
constexpr int kMaxOps = 24;
constexpr int kMaxPhases = 12;
constexpr int kChunkBytes = kTileRows * 16;  // one 8-wide K chunk of a 128-row A operand
constexpr int kTmemCols = 512;

struct TcOp {
    uint32_t a_off[2];     // byte offset (from the dynamic smem base) of up to two A segments
    uint16_t a_chunks[2];  // 8-wide K chunks in each segment; a_chunks[0] even, sum even
    uint16_t n;            // MMA N: multiple of 16, 16..256
    uint16_t d_col;        // first TMEM column of the accumulator
    uint16_t accumulate;   // 1: accumulate onto what the columns already hold
    uint16_t phase_end;    // 1: last op of its phase
};

struct TcProgram {
    int n_ops;
    int n_phases;
    int n_stages;
    int stage_bytes;
    uint32_t stage_off;  // byte offset of stage 0 in dynamic smem (1024-aligned)
    uint32_t a_alt;      // added to every A offset on odd repetitions (double-buffered operand of the batched MLP)
    TcOp op[kMaxOps];
};

// chunks of one weight slab for an op with MMA N = n (even, >= 2)
__host__ __device__ inline int slab_chunks(int stage_bytes, int n) {
    int kc = stage_bytes / (n * 16);
    kc &= ~1;
    return kc < 2 ? 2 : kc;
}
// bytes of the whole program's weight stream (one repetition)
inline size_t program_weight_bytes(const TcProgram& pg) {
    size_t b = 0;
    for (int i = 0; i < pg.n_ops; ++i) b += (size_t)(pg.op[i].a_chunks[0] + pg.op[i].a_chunks[1]) * pg.op[i].n * 16;
    return b;
}

// Debug timeline (-DDRM_TIMELINE): CTA 0 records SM-clock stamps of its handshakes (one copy per translation unit).
#ifdef DRM_TIMELINE
static __device__ long long g_tl[4][4096];
static __device__ int g_tl_n[4];
__device__ __forceinline__ void tl_mark(int role) {
    if (blockIdx.x == 0) { int i = g_tl_n[role]; if (i < 4096) { g_tl[role][i] = clock64(); g_tl_n[role] = i + 1; } }
}
#define DRM_TL(role) tl_mark(role)
#else
#define DRM_TL(role)
#endif

struct ChainBars {
    uint64_t full[8];
    uint64_t empty[8];
    uint64_t acc_full;
    uint64_t a_ready;
    uint32_t tmem_slot;
    uint32_t pad;
};

// Called by every thread at kernel start.  Returns the TMEM base address.
__device__ __forceinline__ uint32_t chain_setup(ChainBars* bars, const TcProgram& pg) {
    const int warp = threadIdx.x >> 5;
    if (threadIdx.x == 0) {
        for (int s = 0; s < pg.n_stages; ++s) {
            mbar_init(&bars->full[s], 1);
            mbar_init(&bars->empty[s], 1);
        }
        mbar_init(&bars->acc_full, 1);
        mbar_init(&bars->a_ready, kEpiWarps);
        fence_barrier_init();
    }
    if (warp == kEpiWarps) tmem_alloc(&bars->tmem_slot, kTmemCols);
    tc_fence_before_sync();
    __syncthreads();
    tc_fence_after_sync();
    return bars->tmem_slot;
}

__device__ __forceinline__ void chain_teardown(uint32_t tmem_base) {
    tc_fence_before_sync();
    __syncthreads();
    if ((threadIdx.x >> 5) == kEpiWarps) tmem_dealloc(tmem_base, kTmemCols);
}

// Producer warp: stream the weight images of `reps` repetitions of the program.
__device__ __forceinline__ void chain_producer(const TcProgram& pg, ChainBars* bars, uint8_t* smem,
                                               const uint8_t* __restrict__ wstream, int reps) {
    if ((threadIdx.x & 31) != 0) return;
    int stage = 0;
    uint32_t phase = 0;
    for (int rep = 0; rep < reps; ++rep) {
        const uint8_t* src = wstream;
        for (int i = 0; i < pg.n_ops; ++i) {
            const int n = pg.op[i].n;
            int chunks = pg.op[i].a_chunks[0] + pg.op[i].a_chunks[1];
            const int kc = slab_chunks(pg.stage_bytes, n);
            while (chunks > 0) {
                const int c = chunks < kc ? chunks : kc;
                const uint32_t bytes = (uint32_t)c * n * 16;
                mbar_wait(&bars->empty[stage], phase ^ 1);
                mbar_expect_tx(&bars->full[stage], bytes);
#ifdef DRM_BULK_PIECES
                {
                    uint8_t* dstp = smem + pg.stage_off + (size_t)stage * pg.stage_bytes;
                    const uint32_t piece = ((bytes / DRM_BULK_PIECES) + 15u) & ~15u;
                    for (uint32_t o = 0; o < bytes; o += piece)
                        bulk_g2s(dstp + o, src + o, min(piece, bytes - o), &bars->full[stage]);
                }
#else
                bulk_g2s(smem + pg.stage_off + (size_t)stage * pg.stage_bytes, src, bytes, &bars->full[stage]);
#endif
                src += bytes;
                chunks -= c;
                if (++stage == pg.n_stages) { stage = 0; phase ^= 1; }
            }
        }
    }
}

// MMA warp: issue the ops of `reps` repetitions.  One thread issues ~150 tcgen05.mma per rollout step, so the
// loop is kept to a handful of instructions per MMA: the 64-bit shared-memory descriptors are built once per op
// (constant high word; the low word = address | LBO advances by plain 32-bit adds) and a weight slab is issued
// by a branch-light inner loop.
//
// TR = 64 ("half tiles"): a CTA owns 64 rows.  tcgen05.mma with M = 64 fills only lanes 0..15 of each TMEM lane quadrant,
// which would idle half of every epilogue warp, so each op is issued as TWO M = 64 MMAs per K step: columns [0, n/2) of
// the weight slab accumulate into lanes 0..15, columns [n/2, n) into lanes 16..31 of the SAME TMEM columns.  An epilogue
// warp then holds 16 rows x 2 column halves in its 32 lanes and the thread-per-(row, chunk) epilogues run unchanged.
// The tensor core does the work of a 128-row tile for 64 rows -- it is not what bounds these kernels.
template <int TR = 128>
__device__ __forceinline__ void chain_mma(const TcProgram& pg, ChainBars* bars, uint8_t* smem, uint32_t tmem_base, int reps) {
    constexpr bool kHalf = TR == 64;
    constexpr int kChunkBytes = TR * 16;
    // The whole warp runs the (warp-uniform) control flow so that descriptors and counters live in uniform
    // registers; only the tcgen05 instructions themselves are issued by one elected lane.
    tmem_base = __shfl_sync(0xffffffffu, tmem_base, 0);
    const uint32_t smem_base = smem_u32(smem);
    const uint32_t desc_hi = (128u >> 4) | (1u << 14);  // SBO = 128 B, descriptor version 1 (bits 32..47)
    const uint32_t a_lo_flags = (uint32_t)(kChunkBytes >> 4) << 16;  // LBO of an A operand
    const uint32_t stage0 = smem_base + pg.stage_off;
    const int n_stages = pg.n_stages, stage_bytes = pg.stage_bytes, n_ops = pg.n_ops;
    int stage = 0;
    uint32_t phase = 0, ready_phase = 0;
    bool phase_start = true;
    for (int rep = 0; rep < reps; ++rep) {
        const uint32_t a_base = smem_base + ((rep & 1) ? pg.a_alt : 0u);
        for (int i = 0; i < n_ops; ++i) {
            const TcOp op = pg.op[i];
            if (phase_start) {
                mbar_wait(&bars->a_ready, ready_phase);
                ready_phase ^= 1;
                tc_fence_after_sync();
                phase_start = false;
                if ((threadIdx.x & 31) == 0) DRM_TL(1);
            }
            const uint32_t n = op.n;
            const int total = op.a_chunks[0] + op.a_chunks[1];
            const int seg0 = op.a_chunks[0];
            const int kc = slab_chunks(stage_bytes, (int)n);
            const uint32_t idesc = make_idesc_bf16(TR, kHalf ? (int)(n >> 1) : (int)n);
            const uint32_t b_flags = n << 16;                      // LBO = n * 16 B, in 16-byte units
            const uint32_t b_step = 2u * n;                        // two K chunks of the weight slab, 16-byte units
            uint32_t a_lo = (((a_base + op.a_off[0]) & 0x3FFFFu) >> 4) | a_lo_flags;
            const uint32_t a_lo1 = (((a_base + op.a_off[1]) & 0x3FFFFu) >> 4) | a_lo_flags;
            const uint32_t d_addr = tmem_base + op.d_col;
            uint32_t acc = op.accumulate;
            int c = 0;
            while (c < total) {
                const int cnt = min(kc, total - c);
                if ((threadIdx.x & 31) == 0) DRM_TL(3);
                mbar_wait(&bars->full[stage], phase);
                if ((threadIdx.x & 31) == 0) DRM_TL(3);
                tc_fence_after_sync();
                uint32_t b_lo = (((stage0 + (uint32_t)stage * stage_bytes) & 0x3FFFFu) >> 4) | b_flags;
#pragma unroll 1
                for (int k = 0; k < cnt; k += 2) {
                    if (c == seg0) a_lo = a_lo1;
                    const uint64_t adesc = ((uint64_t)desc_hi << 32) | a_lo;
                    const uint64_t bdesc = ((uint64_t)desc_hi << 32) | b_lo;
                    if (elect_one()) umma_bf16(d_addr, adesc, bdesc, idesc, acc);
                    if (kHalf) {  // second column half of the slab -> lanes 16..31
                        const uint64_t bdesc2 = ((uint64_t)desc_hi << 32) | (b_lo + (n >> 1));
                        if (elect_one()) umma_bf16(d_addr + (16u << 16), adesc, bdesc2, idesc, acc);
                    }
                    acc = 1;
                    a_lo += (2 * kChunkBytes) >> 4;
                    b_lo += b_step;
                    c += 2;
                }
                if (elect_one()) umma_commit(&bars->empty[stage]);  // slab consumed once these MMAs retire
                if (++stage == n_stages) { stage = 0; phase ^= 1; }
            }
            if (op.phase_end) {
                if (elect_one()) umma_commit(&bars->acc_full);
                phase_start = true;
                if ((threadIdx.x & 31) == 0) DRM_TL(1);
            }
        }
    }
}

// Epilogue-side handshakes.
struct EpiSync {
    ChainBars* bars;
    uint32_t acc_phase;
    __device__ __forceinline__ explicit EpiSync(ChainBars* b) : bars(b), acc_phase(0) {}
    // wait for the accumulators of the next phase
    __device__ __forceinline__ void wait_acc() {
        mbar_wait(&bars->acc_full, acc_phase);
        acc_phase ^= 1;
        tc_fence_after_sync();
        if (threadIdx.x == 0) DRM_TL(0);
        if (threadIdx.x == kEpiThreads - 32) DRM_TL(2);
    }
    // A operands written (generic proxy) + accumulators consumed -> let the MMA warp go on
    __device__ __forceinline__ void release() {
        fence_proxy_async_smem();
        tc_fence_before_sync();
        __syncwarp();
        if (threadIdx.x == 0) DRM_TL(0);
        if (threadIdx.x == kEpiThreads - 32) DRM_TL(2);
        if ((threadIdx.x & 31) == 0) mbar_arrive(&bars->a_ready);
    }
};

// Geometry of an epilogue thread for TR-row tiles.  TR = 128: row = TMEM lane, all chunks.  TR = 64: row = lane % 16 of
// the quadrant, and the upper 16 lanes of a warp see the second column half of every op (chunk index + n/16).
template <int TR>
struct EpiGeo {
    int row, half;
    __device__ __forceinline__ EpiGeo() {
        const int w4 = (threadIdx.x >> 5) & 3, ln = threadIdx.x & 31;
        if (TR == 64) { row = 16 * w4 + (ln & 15); half = ln >> 4; }
        else { row = 32 * w4 + ln; half = 0; }
    }
    // number of accumulator chunks (8 TMEM columns) an op with `nchunks` output chunks occupies
    __device__ __forceinline__ static int tmem_chunks(int nchunks) { return TR == 64 ? nchunks >> 1 : nchunks; }
    // global chunk index of TMEM chunk 0 as seen by this thread
    __device__ __forceinline__ int goff(int nchunks) const { return TR == 64 ? half * (nchunks >> 1) : 0; }
};

// Row / column ownership of an epilogue thread: TMEM lane quadrant = warp % 4, column group = warp / 4.
__device__ __forceinline__ int epi_row() { return 32 * ((threadIdx.x >> 5) & 3) + (threadIdx.x & 31); }
__device__ __forceinline__ int epi_group() { return (threadIdx.x >> 5) >> 2; }
constexpr int kEpiGroups = kEpiWarps / 4;
__device__ __forceinline__ uint32_t epi_taddr(uint32_t tmem_base, int col) {
    return tmem_base + ((uint32_t)(32 * ((threadIdx.x >> 5) & 3)) << 16) + (uint32_t)col;
}

// Walks this thread's accumulator chunks [cbeg, cend) of the region starting at TMEM column col0, four chunks
// (32 columns) per tcgen05.wait::ld so that the TMEM loads and the pointwise math of a batch overlap.
// f(chunk, v[8]) is called for every chunk.
// [cbeg, cend) are TMEM chunk indices; f receives the GLOBAL chunk index c + goff (goff != 0 only for half tiles).
template <class F>
__device__ __forceinline__ void epi_for_chunks(uint32_t tmem_base, int col0, int cbeg, int cend, F&& f, int goff = 0) {
    for (int c = cbeg; c < cend; c += 2) {
        float v[2][8];
        const bool two = c + 1 < cend;
        tmem_ld8(epi_taddr(tmem_base, col0 + c * 8), v[0]);
        if (two) tmem_ld8(epi_taddr(tmem_base, col0 + c * 8 + 8), v[1]);
        tmem_ld_wait();
        f(c + goff, v[0]);
        if (two) f(c + 1 + goff, v[1]);
    }
}
// Same for two accumulator regions read together (two chunks of each per wait).
template <class F>
__device__ __forceinline__ void epi_for_chunks2(uint32_t tmem_base, int colA, int colB, int cbeg, int cend, F&& f, int goff = 0) {
    for (int c = cbeg; c < cend; ++c) {
        float a[8], b[8];
        tmem_ld8(epi_taddr(tmem_base, colA + c * 8), a);
        tmem_ld8(epi_taddr(tmem_base, colB + c * 8), b);
        tmem_ld_wait();
        f(c + goff, a, b);
    }
}
// ELU-backward style epilogue: walks accumulator chunks like epi_for_chunks, but each chunk also needs a saved
// activation chunk from global memory (16 bytes per thread).  Those loads do not depend on the accumulators, so the
// first pair is issued BEFORE the accumulator wait and pair k+1 before pair k is processed.
// ld(c) -> uint4, f(c, v[8], uint4 saved); after_wait() runs once, right after the accumulator wait.
template <class L, class F, class W>
__device__ __forceinline__ void epi_for_chunks_saved(EpiSync& es, uint32_t tmem_base, int col0, int cbeg, int cend, L&& ld, F&& f,
                                                     W&& after_wait, int goff = 0) {
    const uint4 zero = make_uint4(0u, 0u, 0u, 0u);
    uint4 h0 = cbeg < cend ? ld(cbeg + goff) : zero;
    uint4 h1 = cbeg + 1 < cend ? ld(cbeg + 1 + goff) : zero;
    es.wait_acc();
    after_wait();
    for (int c = cbeg; c < cend; c += 2) {
        float v[2][8];
        const bool two = c + 1 < cend;
        tmem_ld8(epi_taddr(tmem_base, col0 + c * 8), v[0]);
        if (two) tmem_ld8(epi_taddr(tmem_base, col0 + c * 8 + 8), v[1]);
        const uint4 n0 = c + 2 < cend ? ld(c + 2 + goff) : zero;
        const uint4 n1 = c + 3 < cend ? ld(c + 3 + goff) : zero;
        tmem_ld_wait();
        f(c + goff, v[0], h0);
        if (two) f(c + 1 + goff, v[1], h1);
        h0 = n0; h1 = n1;
    }
}
template <class L, class F>
__device__ __forceinline__ void epi_for_chunks_saved(EpiSync& es, uint32_t tmem_base, int col0, int cbeg, int cend, L&& ld, F&& f) {
    epi_for_chunks_saved(es, tmem_base, col0, cbeg, cend, ld, f, [] {}, 0);
}

// 8 consecutive fp32 values from a 32-byte aligned shared-memory vector (bias blocks are zero padded)
__device__ __forceinline__ void lds8(const float* src, float (&b)[8]) {
    const float4 lo = *reinterpret_cast<const float4*>(src);
    const float4 hi = *reinterpret_cast<const float4*>(src + 4);
    b[0] = lo.x; b[1] = lo.y; b[2] = lo.z; b[3] = lo.w; b[4] = hi.x; b[5] = hi.y; b[6] = hi.z; b[7] = hi.w;
}
// v[0..8) = row[k .. k+8) of an fp32 row with `kvalid` real columns (zero beyond), widest aligned loads possible.
__device__ __forceinline__ void ldg_row8(const float* __restrict__ row, int k, int kvalid, float (&v)[8]) {
    const float* src = row + k;
    if (k + 8 <= kvalid) {
        const uintptr_t a = reinterpret_cast<uintptr_t>(src);
        if ((a & 15) == 0) {
            const float4 lo = __ldg(reinterpret_cast<const float4*>(src));
            const float4 hi = __ldg(reinterpret_cast<const float4*>(src) + 1);
            v[0] = lo.x; v[1] = lo.y; v[2] = lo.z; v[3] = lo.w; v[4] = hi.x; v[5] = hi.y; v[6] = hi.z; v[7] = hi.w;
            return;
        }
        if ((a & 7) == 0) {
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                const float2 q = __ldg(reinterpret_cast<const float2*>(src) + i);
                v[2 * i] = q.x; v[2 * i + 1] = q.y;
            }
            return;
        }
    }
#pragma unroll
    for (int i = 0; i < 8; ++i) v[i] = (k + i < kvalid) ? __ldg(src + i) : 0.f;
}

// Coherent (L2) loads for rows that ANOTHER kernel running beside this one may still be writing when this kernel
// starts (producer / consumer gating through device counters): `ld.global.nc` (__ldg) requires the data to be read-only
// for the whole lifetime of the kernel, which does not hold there.  .cg caches in L2 only, the point of coherence.
__device__ __forceinline__ float4 ldcg4(const float4* p) {
    float4 v;
    asm volatile("ld.global.cg.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ float2 ldcg2(const float2* p) {
    float2 v;
    asm volatile("ld.global.cg.v2.f32 {%0,%1}, [%2];" : "=f"(v.x), "=f"(v.y) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ float ldcg1(const float* p) {
    float v;
    asm volatile("ld.global.cg.f32 %0, [%1];" : "=f"(v) : "l"(p) : "memory");
    return v;
}
// ldg_row8 with coherent loads
__device__ __forceinline__ void ldcg_row8(const float* row, int k, int kvalid, float (&v)[8]) {
    const float* src = row + k;
    if (k + 8 <= kvalid) {
        const uintptr_t a = reinterpret_cast<uintptr_t>(src);
        if ((a & 15) == 0) {
            const float4 lo = ldcg4(reinterpret_cast<const float4*>(src));
            const float4 hi = ldcg4(reinterpret_cast<const float4*>(src) + 1);
            v[0] = lo.x; v[1] = lo.y; v[2] = lo.z; v[3] = lo.w; v[4] = hi.x; v[5] = hi.y; v[6] = hi.z; v[7] = hi.w;
            return;
        }
    }
#pragma unroll
    for (int i = 0; i < 8; ++i) v[i] = (k + i < kvalid) ? ldcg1(src + i) : 0.f;
}

// row[k .. k+8) (=|+=) v[0..8) for the columns below kvalid, widest aligned accesses possible.
__device__ __forceinline__ void stg_row8(float* row, int k, int kvalid, const float (&v)[8], bool accumulate) {
    float* dst = row + k;
    const uintptr_t a = reinterpret_cast<uintptr_t>(dst);
    if (k + 8 <= kvalid && (a & 15) == 0) {
        float4 lo = make_float4(v[0], v[1], v[2], v[3]), hi = make_float4(v[4], v[5], v[6], v[7]);
        if (accumulate) {
            const float4 o0 = *reinterpret_cast<const float4*>(dst), o1 = *reinterpret_cast<const float4*>(dst + 4);
            lo.x += o0.x; lo.y += o0.y; lo.z += o0.z; lo.w += o0.w; hi.x += o1.x; hi.y += o1.y; hi.z += o1.z; hi.w += o1.w;
        }
        *reinterpret_cast<float4*>(dst) = lo;
        *reinterpret_cast<float4*>(dst + 4) = hi;
        return;
    }
    if ((a & 7) == 0) {
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            if (k + 2 * i + 1 < kvalid) {
                float2 q = make_float2(v[2 * i], v[2 * i + 1]);
                if (accumulate) { const float2 o = *reinterpret_cast<const float2*>(dst + 2 * i); q.x += o.x; q.y += o.y; }
                *reinterpret_cast<float2*>(dst + 2 * i) = q;
            } else if (k + 2 * i < kvalid) {
                dst[2 * i] = accumulate ? dst[2 * i] + v[2 * i] : v[2 * i];
            }
        }
        return;
    }
#pragma unroll
    for (int i = 0; i < 8; ++i) if (k + i < kvalid) dst[i] = accumulate ? dst[i] + v[i] : v[i];
}

// Cooperative fill of a zero-padded bias block in shared memory: dst[i] = i < n ? a[i] (+ b[i]) : 0 for i < npad.
__device__ __forceinline__ void fill_bias(float* dst, const float* a, const float* b, int n, int npad) {
    for (int i = threadIdx.x; i < npad; i += blockDim.x) dst[i] = i < n ? (a ? a[i] : 0.f) + (b ? b[i] : 0.f) : 0.f;
}

// ---------------------------------------------------------------------------------------------- fast pointwise math
// (bf16 path only: the operands of the next GEMM are rounded to 8 bits of mantissa anyway)
__device__ __forceinline__ float elu_fast(float x) { return fmaxf(x, 0.f) + (__expf(fminf(x, 0.f)) - 1.f); }
// ELU(v + b) over one 8-column chunk.  The epilogues are bound by instruction issue (4 warps per scheduler), so this is
// written for instruction count: packed fp32 pairs (add / mul .f32x2), exp2 with flush-to-zero (no denormal fix-up code),
// and ELU(x) = max(x, exp(min(x, 0)) - 1), which holds because exp(x) - 1 >= x everywhere.  About 5 issue slots per
// element instead of 10.
__device__ __forceinline__ void elu_bias8(float (&v)[8], const float (&b)[8]) {
#pragma unroll
    for (int i = 0; i < 8; i += 2) {
        float x0, x1, t0, t1;
        asm("{\n\t.reg .b64 a, c, x;\n\t"
            "mov.b64 a, {%2, %3};\n\tmov.b64 c, {%4, %5};\n\t"
            "add.rn.f32x2 x, a, c;\n\t"
            "mov.b64 {%0, %1}, x;\n\t}"
            : "=f"(x0), "=f"(x1) : "f"(v[i]), "f"(v[i + 1]), "f"(b[i]), "f"(b[i + 1]));
        const float m0 = fminf(x0, 0.f), m1 = fminf(x1, 0.f);
        asm("{\n\t.reg .b64 a, c, s;\n\t.reg .f32 s0, s1, e0, e1;\n\t"
            "mov.b64 a, {%2, %3};\n\tmov.b64 c, {%4, %4};\n\t"
            "mul.rn.f32x2 s, a, c;\n\t"
            "mov.b64 {s0, s1}, s;\n\t"
            "ex2.approx.ftz.f32 e0, s0;\n\tex2.approx.ftz.f32 e1, s1;\n\t"
            "mov.b64 a, {e0, e1};\n\tmov.b64 c, {%5, %5};\n\t"
            "add.rn.f32x2 s, a, c;\n\t"
            "mov.b64 {%0, %1}, s;\n\t}"
            : "=f"(t0), "=f"(t1) : "f"(m0), "f"(m1), "f"(1.4426950408889634f), "f"(-1.f));
        v[i] = fmaxf(x0, t0);
        v[i + 1] = fmaxf(x1, t1);
    }
}
// ELU backward over one chunk: v *= ELU'(pre) with the saved bf16 activation h = ELU(pre): ELU' = 1 for h > 0, h + 1
// otherwise, i.e. min(h, 0) + 1 -- no compare / select; packed fp32 pairs; rows beyond the batch give zeros.
__device__ __forceinline__ void elu_bwd8(float (&v)[8], const uint4 hq, bool live) {
    if (!live) {
#pragma unroll
        for (int i = 0; i < 8; ++i) v[i] = 0.f;
        return;
    }
    const uint32_t w[4] = {hq.x, hq.y, hq.z, hq.w};
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const float m0 = fminf(__uint_as_float(w[i] << 16), 0.f), m1 = fminf(__uint_as_float(w[i] & 0xFFFF0000u), 0.f);
        asm("{\n\t.reg .b64 a, c, f, x;\n\t"
            "mov.b64 a, {%2, %3};\n\tmov.b64 c, {%4, %4};\n\t"
            "add.rn.f32x2 f, a, c;\n\t"
            "mov.b64 x, {%0, %1};\n\t"
            "mul.rn.f32x2 x, x, f;\n\t"
            "mov.b64 {%0, %1}, x;\n\t}"
            : "+f"(v[2 * i]), "+f"(v[2 * i + 1]) : "f"(m0), "f"(m1), "f"(1.f));
    }
}
__device__ __forceinline__ float tanh_fast(float x) {
    float y;
    asm("tanh.approx.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}
__device__ __forceinline__ float sigmoid_fast(float x) { return fmaf(0.5f, tanh_fast(0.5f * x), 0.5f); }
__device__ __forceinline__ float softplus_fast(float x) { return x > 20.f ? x : __logf(1.f + __expf(x)); }

// ---------------------------------------------------------------------------------------------- TCM images
// "Tiled chunk-major" global image of a [rows x F] bf16 matrix: [tile of 128 rows][chunk of 8 features][128][8].
// It is byte-for-byte the shared-memory A-operand image of each row tile, so the epilogue threads store
// it fully coalesced (32 lanes = 32 consecutive rows = 512 contiguous bytes), the backward kernels read it
// back the same way, and the weight-gradient kernel pulls whole tiles with 1-D bulk copies and uses them
// directly as MN-major tcgen05 operands (features = M/N, rows = K).  Rows past the end are zero.
__host__ __device__ inline size_t tcm_bytes(size_t tiles, int chunks) { return tiles * (size_t)chunks * kChunkBytes; }
__device__ __forceinline__ __nv_bfloat16* tcm_at(__nv_bfloat16* base, int chunks_per_tile, int tile, int chunk, int row) {
    return base + (((size_t)tile * chunks_per_tile + chunk) * kTileRows + row) * 8;
}
__device__ __forceinline__ const __nv_bfloat16* tcm_at(const __nv_bfloat16* base, int chunks_per_tile, int tile, int chunk, int row) {
    return base + (((size_t)tile * chunks_per_tile + chunk) * kTileRows + row) * 8;
}
__device__ __forceinline__ void ld_chunk_bf16(const __nv_bfloat16* src, float (&v)[8]) {
    const uint4 q = *reinterpret_cast<const uint4*>(src);
    const uint32_t w[4] = {q.x, q.y, q.z, q.w};
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        v[2 * i] = __uint_as_float(w[i] << 16);
        v[2 * i + 1] = __uint_as_float(w[i] & 0xFFFF0000u);
    }
}

// Weight-gradient GEMMs of a group of Linear layers in ONE launch (drm_wgrad_tc.cu):
//   dW_l[n][col(f)] (+)= sum_rows dY_l[row][n] * X_l[row][f],   db_l[n] (+)= sum_rows dY_l[row][n]
// dY_l and the (up to two) feature segments of X_l are TCM images over the same row tiles.
struct WgSeg {
    const __nv_bfloat16* base;  // TCM image
    int cpt;                    // chunks per tile of that image
    int chunk0, nchunks;        // chunk range used
    int col0, kvalid;           // dW column of the segment's first feature; number of real features
};
struct WgLayer {
    const __nv_bfloat16* dy;    // TCM image, dy_cpt chunks per tile
    int dy_cpt, nout;
    WgSeg seg[2];
    int nseg;
    float* dW; int ld_dw;
    float* db;                  // may be null
    int n_tiles_n;              // 128-feature tiles of nout
    int n_parts, part_chunks;   // X feature chunks (incl. the ones chunk) split into parts of <= 32 chunks
    int splits;                 // row-tile splits
    int cta0;                   // first CTA of this layer
};
constexpr int kMaxWgLayers = 10;
struct WgParams {
    int n_layers, row_tiles, total_ctas;
    int tile_rows;  // rows per TCM tile of the images (0 = 128)
    // optional producer gating: row tile i belongs to step i / tiles_per_step and exists once wait_flags[step] >= wait_target;
    // reverse = the producer runs from the last step to the first
    const int* wait_flags; int wait_target, tiles_per_step, reverse;
    WgLayer L[kMaxWgLayers];
};
// fills n_tiles_n / parts / splits / cta0 / total_ctas from the rest; slots > 0 bounds the grid (CTAs resident at once)
void wgrad_tc_plan(WgParams& w, int slots = 0);
int launch_wgrad_tc(const WgParams& w, cudaStream_t stream);

// ---------------------------------------------------------------------------------------------- weight images
// One job packs a [n_pad x 8*k_chunks] bf16 image (chunk-major) from an fp32 matrix:
//   image(n, k) = src[(n0 + n) * ld + col(k)]            (transposed = 0: N indexes src rows)
//               = src[row(k) * ld + (n0 + n)]            (transposed = 1: N indexes src columns)
// where k runs through up to three K segments; a segment maps K range [kstart, kstart + len) to source
// indices src0 .. src0 + len - 1.  Everything else is zero.
struct PackTcJob {
    const float* src;
    __nv_bfloat16* dst;
    int ld, transposed;
    int n0, n_valid, n_pad;
    int n2_start, n2_valid, n2_src0;  // optional second N segment: image rows [n2_start, +n2_valid) <- src index n2_src0 + i
    int k_chunks;
    int nseg;
    int seg_kstart[3], seg_len[3], seg_src0[3];
};
constexpr int kMaxPackTcJobs = 28;
constexpr int kMaxZeroJobs = 16;
struct PackTcJobs {
    PackTcJob j[kMaxPackTcJobs];
    int n = 0;
    // gradient sinks to zero-fill in the same launch (they are accumulated with atomics by the weight-gradient kernel)
    float* z[kMaxZeroJobs];
    int zn[kMaxZeroJobs];
    int nz = 0;
    void zero(float* p, size_t count) { if (nz < kMaxZeroJobs) { z[nz] = p; zn[nz] = (int)count; } ++nz; }
};
int launch_pack_tc(const PackTcJobs& jobs, cudaStream_t stream);

}  // namespace tc
}  // namespace drm
