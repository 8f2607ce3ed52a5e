// Weight-gradient GEMMs on the tensor cores (bf16 operands, fp32 accumulate in TMEM):
//     dW_l[n][col(f)] += sum_rows dY_l[row][n] * X_l[row][f]        db_l[n] += sum_rows dY_l[row][n]
// for a whole group of Linear layers in one launch.  Both operands are TCM images (drm_tc_chain.cuh):
// a 128-row tile of dY (16 feature chunks) and of X (<= 32 feature chunks) is pulled into shared memory
// with 1-D bulk copies and used as-is as MN-major tcgen05 operands (features = M / N, rows = K = 128
// per stage, 8 MMAs).  The bias gradient falls out of the same GEMM through a constant "ones" feature
// chunk appended to X.  A CTA owns (layer, 128-feature tile of dY, part of X's features, row-tile split),
// accumulates over its row tiles in TMEM and adds its partial sum to dW / db with fp32 atomics.
#include "drm_tc_chain.cuh"

namespace drm {
namespace tc {

constexpr int kWgThreads = 192;          // warps 0-3 epilogue, 4 producer, 5 MMA
constexpr int kWgStages = 2;
constexpr int kWgAChunks = 16;           // 128 out-features
constexpr int kWgBChunks = 32;           // <= 256 in-features per part
// one stage holds a tile of dY (16 chunks) and of X (32 chunks): 96 KB with 128-row tiles, 48 KB with 64-row tiles --
// the half-tile images therefore leave room for two CTAs per SM (2 x 256 TMEM columns)
__host__ __device__ inline int wg_stage_bytes(int tr) { return (kWgAChunks + kWgBChunks) * tr * 16; }
__host__ __device__ inline int wg_smem(int tr) { return 1024 + kWgStages * wg_stage_bytes(tr); }
constexpr int kWgTmemCols = 256;

struct WgBars {
    uint64_t full[kWgStages];
    uint64_t empty[kWgStages];
    uint64_t acc_full;
    uint32_t tmem_slot;
};

__device__ __forceinline__ void red_add_v4(float* p, float a, float b, float c, float d) {
    asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(p), "f"(a), "f"(b), "f"(c), "f"(d) : "memory");
}
__device__ __forceinline__ void red_add_v2(float* p, float a, float b) {
    asm volatile("red.global.add.v2.f32 [%0], {%1, %2};" ::"l"(p), "f"(a), "f"(b) : "memory");
}

__host__ __device__ inline uint32_t make_idesc_bf16_mn(int M, int N) { return make_idesc_bf16(M, N) | (1u << 15) | (1u << 16); }

__global__ void __launch_bounds__(kWgThreads, 2) wgrad_tc_kernel(const __grid_constant__ WgParams w) {
    extern __shared__ __align__(1024) uint8_t smem[];
    WgBars* bars = reinterpret_cast<WgBars*>(smem);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    // rows per TCM tile of the images (128, or 64 for the half-tile rollouts) and bytes of one chunk of such a tile;
    // the shared-memory regions keep their 128-row capacity
    const int tr = w.tile_rows ? w.tile_rows : kTileRows;
    const int cb = tr * 16;
    const int stage_bytes = wg_stage_bytes(tr), a_bytes = kWgAChunks * cb;
    // ---- which job is mine
    int li = 0;
    for (int i = 1; i < w.n_layers; ++i) if ((int)blockIdx.x >= w.L[i].cta0) li = i;
    const WgLayer& L = w.L[li];
    int local = blockIdx.x - L.cta0;
    const int split = local % L.splits; local /= L.splits;
    const int part = local % L.n_parts;
    const int ntile = local / L.n_parts;
    const int nch0 = L.seg[0].nchunks, nch1 = L.nseg > 1 ? L.seg[1].nchunks : 0;
    const int ft = nch0 + nch1 + (L.db ? 1 : 0);            // logical feature chunks of X (+ ones)
    const int f_begin = part * L.part_chunks;
    const int f_end = min(ft, f_begin + L.part_chunks);
    const int npc = f_end - f_begin;                         // chunks of this part
    const int n_mma = ((npc + 1) & ~1) * 8;
    const int my_tiles = split < w.row_tiles ? (w.row_tiles - split + L.splits - 1) / L.splits : 0;

    if (threadIdx.x == 0) {
        for (int s = 0; s < kWgStages; ++s) { mbar_init(&bars->full[s], 1); mbar_init(&bars->empty[s], 1); }
        mbar_init(&bars->acc_full, 1);
        fence_barrier_init();
    }
    if (warp == 4) tmem_alloc(&bars->tmem_slot, kWgTmemCols);
    // constant chunks of the B operand: the ones chunk (bias gradient) and, if npc is odd, a zero pad chunk
    if (warp < 4) {
        const int ones_pos = (L.db && f_end == ft) ? npc - 1 : -1;
        for (int s = 0; s < kWgStages; ++s) {
            uint8_t* B = smem + 1024 + s * stage_bytes + a_bytes;
            for (int r = threadIdx.x; r < tr; r += 128) {
                if (ones_pos >= 0) {
                    uint4 q = make_uint4(0x00003F80u, 0u, 0u, 0u);  // bf16 1.0 in element 0
                    *reinterpret_cast<uint4*>(B + ((size_t)ones_pos * tr + r) * 16) = q;
                }
                if (npc & 1) *reinterpret_cast<uint4*>(B + ((size_t)npc * tr + r) * 16) = make_uint4(0u, 0u, 0u, 0u);
            }
        }
        fence_proxy_async_smem();
    }
    tc_fence_before_sync();
    __syncthreads();
    tc_fence_after_sync();
    const uint32_t tmem_base = bars->tmem_slot;
    const int a_chunks = min(kWgAChunks, L.dy_cpt - ntile * kWgAChunks);

    if (warp == 4) {
        if (lane == 0) {
            // data chunk ranges of the two segments inside this part
            const int s0b = max(f_begin, 0), s0e = min(f_end, nch0);
            const int s1b = max(f_begin, nch0), s1e = min(f_end, nch0 + nch1);
            uint32_t bytes = (uint32_t)a_chunks * cb;
            if (s0e > s0b) bytes += (uint32_t)(s0e - s0b) * cb;
            if (s1e > s1b) bytes += (uint32_t)(s1e - s1b) * cb;
            int stage = 0; uint32_t phase = 0;
            for (int it = 0; it < my_tiles; ++it) {
                const int tile_i = split + it * L.splits;
                const int tile = w.reverse ? w.row_tiles - 1 - tile_i : tile_i;
                if (w.wait_flags) {  // the producer of the images runs beside this kernel
                    const int* f = w.wait_flags + tile / w.tiles_per_step;
                    wait_counter_ge(f, w.wait_target, 200);
                    asm volatile("fence.proxy.async;" ::: "memory");  // generic-proxy writes of the producer -> bulk copies
                }
                uint8_t* A = smem + 1024 + stage * stage_bytes;
                uint8_t* B = A + a_bytes;
                mbar_wait(&bars->empty[stage], phase ^ 1);
                mbar_expect_tx(&bars->full[stage], bytes);
                auto img = [&](const __nv_bfloat16* base, int cpt, int chunk) {  // chunk of TCM tile `tile` (tr rows per tile)
                    return base + ((size_t)tile * cpt + chunk) * tr * 8;
                };
                bulk_g2s(A, img(L.dy, L.dy_cpt, ntile * kWgAChunks), (uint32_t)a_chunks * cb, &bars->full[stage]);
                if (s0e > s0b)
                    bulk_g2s(B + (size_t)(s0b - f_begin) * cb, img(L.seg[0].base, L.seg[0].cpt, L.seg[0].chunk0 + s0b),
                             (uint32_t)(s0e - s0b) * cb, &bars->full[stage]);
                if (s1e > s1b)
                    bulk_g2s(B + (size_t)(s1b - f_begin) * cb, img(L.seg[1].base, L.seg[1].cpt, L.seg[1].chunk0 + (s1b - nch0)),
                             (uint32_t)(s1e - s1b) * cb, &bars->full[stage]);
                if (++stage == kWgStages) { stage = 0; phase ^= 1; }
            }
        }
    } else if (warp == 5) {
        if (lane == 0) {
            const uint32_t idesc = make_idesc_bf16_mn(128, n_mma);  // M = 128 output features of this n-tile
            int stage = 0; uint32_t phase = 0;
            for (int it = 0; it < my_tiles; ++it) {
                const uint32_t A = smem_u32(smem + 1024 + stage * stage_bytes);
                const uint32_t B = A + a_bytes;
                mbar_wait(&bars->full[stage], phase);
                tc_fence_after_sync();
                for (int k = 0; k < tr / 16; ++k)  // K = rows of the tile, 16 per instruction
                    umma_bf16(tmem_base, make_desc(A + k * 256, 128, cb), make_desc(B + k * 256, 128, cb), idesc, it > 0 || k > 0);
                umma_commit(&bars->empty[stage]);
                if (++stage == kWgStages) { stage = 0; phase ^= 1; }
            }
            umma_commit(&bars->acc_full);
        }
    } else if (my_tiles > 0) {
        mbar_wait(&bars->acc_full, 0);
        tc_fence_after_sync();
        const int gn = ntile * 128 + warp * 32 + lane;
        const uint32_t taddr = tmem_base + ((uint32_t)(warp * 32) << 16);
        for (int j = 0; j < npc; ++j) {
            float v[8];
            tmem_ld8(taddr + j * 8, v);
            tmem_ld_wait();
            if (gn >= L.nout) continue;
            const int f = f_begin + j;
            if (f < nch0 + nch1) {
                const WgSeg& sg = f < nch0 ? L.seg[0] : L.seg[1];
                const int fc = f < nch0 ? f : f - nch0;
                float* dst = L.dW + (size_t)gn * L.ld_dw + sg.col0 + fc * 8;
                // vector reductions (one L2 operation per 8 / 16 bytes): the partial sums of ~40 row splits land on the
                // same addresses, and the L2 atomic rate -- not the GEMM -- bounds this kernel
                const bool full = fc * 8 + 8 <= sg.kvalid;
                if (full && (reinterpret_cast<uintptr_t>(dst) & 15) == 0) {
                    red_add_v4(dst, v[0], v[1], v[2], v[3]);
                    red_add_v4(dst + 4, v[4], v[5], v[6], v[7]);
                } else if (full && (reinterpret_cast<uintptr_t>(dst) & 7) == 0) {
#pragma unroll
                    for (int i = 0; i < 8; i += 2) red_add_v2(dst + i, v[i], v[i + 1]);
                } else {
#pragma unroll
                    for (int i = 0; i < 8; ++i) if (fc * 8 + i < sg.kvalid) atomicAdd(dst + i, v[i]);
                }
            } else {
                atomicAdd(L.db + gn, v[0]);
            }
        }
    }
    tc_fence_before_sync();
    __syncthreads();
    if (warp == 4) tmem_dealloc(tmem_base, kWgTmemCols);
}

void wgrad_tc_plan(WgParams& w, int slots) {
    int base = 0;
    for (int i = 0; i < w.n_layers; ++i) {
        WgLayer& L = w.L[i];
        L.n_tiles_n = (L.nout + 127) / 128;
        const int ft = L.seg[0].nchunks + (L.nseg > 1 ? L.seg[1].nchunks : 0) + (L.db ? 1 : 0);
        L.n_parts = (ft + kWgBChunks - 1) / kWgBChunks;
        L.part_chunks = (ft + L.n_parts - 1) / L.n_parts;
        base += L.n_tiles_n * L.n_parts;
    }
    // row-tile splits: at most two CTAs per SM in total (one wave of co-resident pairs with 64-row tiles, two full waves
    // with 128-row tiles -- never a third, nearly empty one), but keep >= 4 row tiles per CTA
    // (measured at 2 500 x 15: flat between one and two CTAs per SM, slower outside)
    if (slots <= 0) slots = 2 * num_sms();
    int splits = base > 0 ? slots / base : 1;
    const int max_splits = (w.row_tiles + 3) / 4;
    if (splits > max_splits) splits = max_splits;
    if (splits < 1) splits = 1;
    int cta = 0;
    for (int i = 0; i < w.n_layers; ++i) {
        WgLayer& L = w.L[i];
        L.splits = splits;
        L.cta0 = cta;
        cta += L.n_tiles_n * L.n_parts * splits;
    }
    w.total_ctas = cta;
}

int launch_wgrad_tc(const WgParams& w, cudaStream_t stream) {
    if (w.n_layers == 0 || w.total_ctas == 0 || w.row_tiles == 0) return DRM_OK;
    DRM_REQUIRE(w.n_layers <= kMaxWgLayers, DRM_E_SHAPE, "too many layers in one weight-gradient launch");
    const int smem = wg_smem(w.tile_rows ? w.tile_rows : kTileRows);
    DRM_CUDA_OK(cudaFuncSetAttribute(wgrad_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    ProfScope prof(PROF_WGRAD, stream);
    wgrad_tc_kernel<<<w.total_ctas, kWgThreads, smem, stream>>>(w);
    DRM_CUDA_OK(cudaGetLastError());
    return DRM_OK;
}

}  // namespace tc
void warm_wgrad_tc() { warm_kernel(tc::wgrad_tc_kernel); }

}  // namespace drm
