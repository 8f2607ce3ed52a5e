// bf16 tensor-core (tcgen05) path of the batched MLP: DenseModel / ActionDecoder feed-forward stacks
// (reference: dreamer/models/dense.py:28-43, dreamer/models/action.py:37-52).  Operands are bf16, the
// accumulators fp32 in TMEM, bias / ELU / everything pointwise fp32.  See drm_tc_chain.cuh for the engine.
#include <string.h>
#include "drm_mlp.cuh"
#include "drm_tc_chain.cuh"

namespace drm {
namespace tc {

// ------------------------------------------------------------------------------------------------
// weight image packing
// ------------------------------------------------------------------------------------------------
__global__ void pack_tc_kernel(PackTcJobs jobs) {
    if ((int)blockIdx.y == jobs.n) {  // the extra grid row zero-fills the gradient sinks
        for (int q = 0; q < jobs.nz; ++q) {
            float* z = jobs.z[q];
            for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < jobs.zn[q]; i += gridDim.x * blockDim.x) z[i] = 0.f;
        }
        return;
    }
    const PackTcJob j = jobs.j[blockIdx.y];
    const int total = j.k_chunks * j.n_pad * 8;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
        const int kk = i & 7;
        const int n = (i >> 3) % j.n_pad;
        const int chunk = (i >> 3) / j.n_pad;
        const int k = chunk * 8 + kk;
        float v = 0.f;
        int ns = -1;
        if (n < j.n_valid) ns = j.n0 + n;
        else if (n >= j.n2_start && n < j.n2_start + j.n2_valid) ns = j.n2_src0 + (n - j.n2_start);
        if (ns >= 0) {
            for (int s = 0; s < j.nseg; ++s) {
                const int d = k - j.seg_kstart[s];
                if (d >= 0 && d < j.seg_len[s]) {
                    const int ks = j.seg_src0[s] + d;
                    v = j.transposed ? j.src[(size_t)ks * j.ld + ns] : j.src[(size_t)ns * j.ld + ks];
                }
            }
        }
        j.dst[i] = __float2bfloat16_rn(v);
    }
}

int launch_pack_tc(const PackTcJobs& jobs, cudaStream_t stream) {
    if (jobs.n == 0 && jobs.nz == 0) return DRM_OK;
    DRM_REQUIRE(jobs.n <= kMaxPackTcJobs, DRM_E_SHAPE, "too many tensor-core pack jobs");
    DRM_REQUIRE(jobs.nz <= kMaxZeroJobs, DRM_E_SHAPE, "too many zero-fill jobs");
    dim3 grid(24, jobs.n + (jobs.nz > 0 ? 1 : 0));
    ProfScope prof(PROF_PACK, stream);
    pack_tc_kernel<<<grid, 256, 0, stream>>>(jobs);
    DRM_CUDA_OK(cudaGetLastError());
    return DRM_OK;
}

// ------------------------------------------------------------------------------------------------
// device-side description
// ------------------------------------------------------------------------------------------------
struct MlpTcDev {
    int M, k0, k1, c0, c1, cin;  // input segments: widths and 8-wide chunk counts (cin = c0 + c1 rounded up to even)
    int n_linear, hidden, ch, out_dim, cout;  // ch / cout: chunks of a hidden layer / of the output (even)
    const float* x0; int64_t ld0;
    const float* x1; int64_t ld1;
    float* y;
    const float* b[DRM_MAX_LINEAR];
    const uint8_t* wstream;
    __nv_bfloat16* sx;                    // TCM image of the input, cin chunks per tile (or nullptr)
    __nv_bfloat16* sh[DRM_MAX_LINEAR];    // TCM images of the hidden activations, ch chunks per tile
    uint32_t act_off;                     // shared-memory offset of the activation buffer
    uint32_t bias_off; int bstride;       // zero-padded fp32 bias blocks in shared memory: [n_linear][bstride]
    // backward only
    const float* dy; const float* dy_scale;
    float* dx0; float* dx1; int accumulate_dx;
    __nv_bfloat16* dys[DRM_MAX_LINEAR];   // TCM images of dY_l (ch chunks; cout for the last layer) or nullptr
    float* db_last;                       // bias gradient of the last Linear, reduced in fp32 from the exact dY
    const int* wait_flags; int wait_target, wait_rows, wait_steps;  // producer gating of the input rows (forward)
    // backward hand-over between kernels on different streams (drm_mlp_bwd_call)
    int reverse_tiles; int* tile_done; const int* tile_wait; int* step_rows_done; int step_rows, step_count;
    TcProgram pg;
};


// ------------------------------------------------------------------------------------------------
// forward
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(kChainThreads, 1) mlp_fwd_tc_kernel(const __grid_constant__ MlpTcDev p) {
    extern __shared__ __align__(1024) uint8_t smem[];
    ChainBars* bars = reinterpret_cast<ChainBars*>(smem);
    float* sbias = reinterpret_cast<float*>(smem + p.bias_off);
    for (int l = 0; l < p.n_linear; ++l)
        fill_bias(sbias + l * p.bstride, p.b[l], nullptr, l == p.n_linear - 1 ? p.out_dim : p.hidden, p.bstride);
    const uint32_t tmem_base = chain_setup(bars, p.pg);
    const int warp = __shfl_sync(0xffffffffu, threadIdx.x >> 5, 0);  // warp-uniform for the compiler
    const int tiles = (p.M + kTileRows - 1) / kTileRows;
    const int reps = blockIdx.x < tiles ? (tiles - blockIdx.x + gridDim.x - 1) / gridDim.x : 0;
    if (warp == kEpiWarps) {
        chain_producer(p.pg, bars, smem, p.wstream, reps);
    } else if (warp == kEpiWarps + 1) {
        chain_mma(p.pg, bars, smem, tmem_base, reps);
    } else {
        EpiSync es(bars);
        uint8_t* act = smem + p.act_off;
        const int row = epi_row(), grp = epi_group();
        const int last = p.n_linear - 1;
        // coalesced staging needs vector-aligned rows: x1 as float4 (k1 % 8 == 0 keeps its chunks whole), x0 as float2
        const bool fast_stage = p.x1 && p.k1 > 0 && (p.k1 & 7) == 0 && (p.ld1 & 3) == 0 &&
                                (reinterpret_cast<uintptr_t>(p.x1) & 15) == 0 &&
                                p.c0 <= 4 && (p.c0 == 0 || (p.x0 && (p.k0 & 1) == 0 && (p.ld0 & 1) == 0 && (reinterpret_cast<uintptr_t>(p.x0) & 7) == 0));
        // ---- staging of one input tile: fp32 global -> bf16 chunk-major shared [+ TCM copy for the wgrad], in parts
        // so that the NEXT tile can be staged, slice by slice, behind the releases of the current one (the epilogue
        // warps would otherwise idle while the tensor core runs).  Loads go 8 rows x 64 (x1) / 32 (x0) contiguous
        // bytes per warp instruction, kU in flight per thread: the staging is latency bound.
        const int wq = threadIdx.x >> 5, ln = threadIdx.x & 31;
        constexpr int kU = 7;
        const int q4 = p.k1 >> 2, qblocks1 = (q4 + 3) >> 2, nblocks1 = (kTileRows / 8) * qblocks1;
        const int ublocks1 = (nblocks1 + kEpiWarps - 1) / kEpiWarps;  // x1 blocks per warp
        const int n_parts = fast_stage ? (ublocks1 + kU - 1) / kU + 1 : 1;
        // stage_issue() puts this thread's loads of one part in flight (registers), stage_commit() converts and stores.
        float4 sv4[kU];
        auto blk1 = [&](int part, int u, int& r, int& q) {  // x1 block slot -> (row, float4 column) of this lane
            const int b = wq + (part * kU + u) * kEpiWarps;
            const int rb = b / qblocks1, qb = b - rb * qblocks1;
            r = rb * 8 + (ln >> 2); q = qb * 4 + (ln & 3);
            return b < nblocks1 && q < q4;
        };
        auto blk0 = [&](int u, int& r, int& q) {            // x0 block slot -> (row, float2 column)
            const int b = wq + u * kEpiWarps;
            const int rb = b / max(p.c0, 1), qb = b - rb * max(p.c0, 1);
            r = rb * 8 + (ln >> 2); q = qb * 4 + (ln & 3);
            return b < (kTileRows / 8) * p.c0;
        };
        auto stage_issue = [&](int tile, int part) {
            if (!fast_stage) return;
            const int row0 = tile * kTileRows;
            if (part < n_parts - 1) {
#pragma unroll
                for (int u = 0; u < kU; ++u) {
                    int r, q;
                    sv4[u] = make_float4(0.f, 0.f, 0.f, 0.f);
                    if (blk1(part, u, r, q) && row0 + r < p.M)
                        sv4[u] = ldcg4(reinterpret_cast<const float4*>(p.x1 + (size_t)(row0 + r) * p.ld1) + q);  // rows may come from a kernel running beside this one
                }
            } else {
#pragma unroll
                for (int u = 0; u < 4; ++u) {  // requires c0 <= 4 * kEpiWarps / 16 blocks per warp = 4 (k0 <= 32)
                    int r, q;
                    sv4[u] = make_float4(0.f, 0.f, 0.f, 0.f);
                    if (blk0(u, r, q) && 2 * q < p.k0 && row0 + r < p.M) {
                        const float2 w = ldcg2(reinterpret_cast<const float2*>(p.x0 + (size_t)(row0 + r) * p.ld0) + q);
                        sv4[u].x = w.x; sv4[u].y = w.y;
                    }
                }
            }
        };
        auto stage_commit = [&](int tile, uint8_t* dst, int part) {
            const int row0 = tile * kTileRows;
            if (!fast_stage) {
                for (int idx = threadIdx.x; idx < p.cin * kTileRows; idx += kEpiThreads) {
                    const int chunk = idx / kTileRows, r = idx - chunk * kTileRows;
                    const int grow = row0 + r;
                    float v[8];
#pragma unroll
                    for (int i = 0; i < 8; ++i) v[i] = 0.f;
                    if (grow < p.M) {
                        if (chunk < p.c0) ldcg_row8(p.x0 + (size_t)grow * p.ld0, chunk * 8, p.k0, v);
                        else if (chunk < p.c0 + p.c1) ldcg_row8(p.x1 + (size_t)grow * p.ld1, (chunk - p.c0) * 8, p.k1, v);
                    }
                    st_chunk_bf16(dst + (size_t)idx * 16, v);
                    if (p.sx) st_chunk_bf16(p.sx + ((size_t)tile * p.cin * kTileRows + idx) * 8, v);
                }
                return;
            }
            uint8_t* sx_t = p.sx ? reinterpret_cast<uint8_t*>(p.sx + (size_t)tile * p.cin * kTileRows * 8) : nullptr;
            if (part < n_parts - 1) {
#pragma unroll
                for (int u = 0; u < kU; ++u) {
                    int r, q;
                    if (blk1(part, u, r, q)) {
                        const uint2 w = make_uint2(pack_bf16x2(sv4[u].x, sv4[u].y), pack_bf16x2(sv4[u].z, sv4[u].w));
                        const size_t off = ((size_t)(p.c0 + (q >> 1)) * kTileRows + r) * 16 + (q & 1) * 8;
                        *reinterpret_cast<uint2*>(dst + off) = w;
                        if (sx_t) *reinterpret_cast<uint2*>(sx_t + off) = w;
                    }
                }
                return;
            }
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                int r, q;
                if (blk0(u, r, q)) {
                    const uint32_t w = pack_bf16x2(sv4[u].x, sv4[u].y);
                    const size_t off = ((size_t)(q >> 2) * kTileRows + r) * 16 + (q & 3) * 4;
                    *reinterpret_cast<uint32_t*>(dst + off) = w;
                    if (sx_t) *reinterpret_cast<uint32_t*>(sx_t + off) = w;
                }
            }
            for (int idx = (p.c0 + p.c1) * kTileRows + threadIdx.x; idx < p.cin * kTileRows; idx += kEpiThreads) {
                *reinterpret_cast<uint4*>(dst + (size_t)idx * 16) = make_uint4(0u, 0u, 0u, 0u);
                if (sx_t) *reinterpret_cast<uint4*>(sx_t + (size_t)idx * 16) = make_uint4(0u, 0u, 0u, 0u);
            }
        };
        auto stage_part = [&](int tile, uint8_t* dst, int part) { stage_issue(tile, part); stage_commit(tile, dst, part); };
        // Producer gating (drm_mlp_call.wait_flags): the rows of step t exist once wait_flags[t] >= wait_target.  Lane 0
        // of each warp polls with an acquire load before the warp touches a tile it has not checked yet.
        int ready_tile = -1;
        auto wait_tile = [&](int tile) {
            if (!p.wait_flags || tile <= ready_tile) return;
            if (ln == 0) {
                const int last_row = min(p.M, (tile + 1) * kTileRows) - 1;
                const int t_hi = min(p.wait_steps - 1, last_row / p.wait_rows);
                const int* f = p.wait_flags + t_hi;
                wait_counter_ge(f, p.wait_target, 200);
            }
            __syncwarp();
            ready_tile = tile;
        };
        auto l2_pull = [&](int tile) {  // one thread: the rows of `tile` into L2
            if (tile < tiles && threadIdx.x == 0) {
                const int nrows = min(kTileRows, p.M - tile * kTileRows);
                if (p.x0) l2_prefetch(p.x0 + (size_t)tile * kTileRows * p.ld0, ((size_t)(nrows - 1) * p.ld0 + p.k0) * 4);
                if (p.x1) l2_prefetch(p.x1 + (size_t)tile * kTileRows * p.ld1, ((size_t)(nrows - 1) * p.ld1 + p.k1) * 4);
            }
        };
        const bool dbl = p.pg.a_alt != 0;  // second operand buffer available
        if ((int)blockIdx.x < tiles) {
            wait_tile(blockIdx.x);
            if (!p.wait_flags) l2_pull(blockIdx.x + gridDim.x);
            for (int part = 0; part < n_parts; ++part) stage_part(blockIdx.x, act, part);
            es.release();
        }
        int it = 0;
        for (int tile = blockIdx.x; tile < tiles; tile += gridDim.x, ++it) {
            const int row0 = tile * kTileRows;
            uint8_t* act_cur = act + ((it & 1) ? p.pg.a_alt : 0u);
            uint8_t* act_nxt = act + ((it & 1) ? 0u : p.pg.a_alt);
            const int nxt = tile + gridDim.x;
            const bool has_next = nxt < tiles;
            int parts_done = 0;
            if (!p.wait_flags) l2_pull(nxt + gridDim.x);
            const int grow = row0 + row;
            const bool live = grow < p.M;
            for (int l = 0; l <= last; ++l) {
                es.wait_acc();
                const float* bias = sbias + l * p.bstride;
                if (l < last) {
                    // a slice of the next tile's input: loads in flight across this epilogue, stored after its release
                    const bool slice = dbl && has_next && parts_done < n_parts;
                    if (slice) { wait_tile(nxt); stage_issue(nxt, parts_done); }
                    // hidden layer: h = ELU(acc + b) -> bf16 A operand of the next layer (in place) [+ TCM save]
                    const int per = (p.ch + kEpiGroups - 1) / kEpiGroups;
                    const int cbeg = grp * per, cend = min(p.ch, cbeg + per);
                    __nv_bfloat16* sv = p.sh[l];
                    epi_for_chunks(tmem_base, 0, cbeg, cend, [&](int c, float (&v)[8]) {
                        float b[8];
                        lds8(bias + c * 8, b);
                        elu_bias8(v, b);
                        const uint4 q = pack_bf16x8(v);
                        *reinterpret_cast<uint4*>(act_cur + ((size_t)c * kTileRows + row) * 16) = q;
                        if (sv) *reinterpret_cast<uint4*>(tcm_at(sv, p.ch, tile, c, row)) = q;
                    });
                    es.release();
                    if (slice) stage_commit(nxt, act_nxt, parts_done++);
                } else {
                    // output layer: y = acc + b (fp32, row-major)
                    const int oc = (p.out_dim + 7) / 8;
                    for (int c = grp; c < oc; c += kEpiGroups) {
                        float v[8];
                        tmem_ld8(epi_taddr(tmem_base, c * 8), v);
                        tmem_ld_wait();
                        if (live) {
#pragma unroll
                            for (int i = 0; i < 8; ++i) {
                                const int col = c * 8 + i;
                                if (col < p.out_dim) p.y[(size_t)grow * p.out_dim + col] = v[i] + bias[col];
                            }
                        }
                    }
                }
            }
            if (has_next) {
                // whatever is left of the next tile's input (everything when there is no second buffer: the operand
                // buffer is free now that the output GEMM has retired); release() also covers the TMEM loads above
                uint8_t* dst = dbl ? act_nxt : act;
                wait_tile(nxt);
                while (parts_done < n_parts) stage_part(nxt, dst, parts_done++);
                es.release();
            }
        }
    }
    chain_teardown(tmem_base);
}

// ------------------------------------------------------------------------------------------------
// backward: data gradients.  dH_{l-1} = (dH_l . W_l) * ELU'(h_{l-1}), dX = dH_0 . W_0; every dH_l is also
// written to global memory as a TCM image for the weight-gradient GEMMs.
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(kChainThreads, 1) mlp_bwd_tc_kernel(const __grid_constant__ MlpTcDev p) {
    extern __shared__ __align__(1024) uint8_t smem[];
    ChainBars* bars = reinterpret_cast<ChainBars*>(smem);
    const uint32_t tmem_base = chain_setup(bars, p.pg);
    const int warp = __shfl_sync(0xffffffffu, threadIdx.x >> 5, 0);  // warp-uniform for the compiler
    const int tiles = (p.M + kTileRows - 1) / kTileRows;
    const int reps = blockIdx.x < tiles ? (tiles - blockIdx.x + gridDim.x - 1) / gridDim.x : 0;
    if (warp == kEpiWarps) {
        chain_producer(p.pg, bars, smem, p.wstream, reps);
    } else if (warp == kEpiWarps + 1) {
        chain_mma(p.pg, bars, smem, tmem_base, reps);
    } else {
        EpiSync es(bars);
        uint8_t* act = smem + p.act_off;
        const int row = epi_row(), grp = epi_group();
        const int last = p.n_linear - 1;
        const float scale = p.dy_scale ? __ldg(p.dy_scale) : 1.f;
        const bool want_dx = p.dx0 != nullptr || p.dx1 != nullptr;
        for (int ti = blockIdx.x; ti < tiles; ti += gridDim.x) {
            const int tile = p.reverse_tiles ? tiles - 1 - ti : ti;
            const int row0 = tile * kTileRows;
            // ---- stage dY of the last Linear (scaled) as the first A operand [+ TCM copy for its wgrad]
            for (int idx = threadIdx.x; idx < p.cout * kTileRows; idx += kEpiThreads) {
                const int chunk = idx / kTileRows, r = idx - chunk * kTileRows;
                const int grow = row0 + r;
                float v[8];
#pragma unroll
                for (int i = 0; i < 8; ++i) {
                    const int col = chunk * 8 + i;
                    v[i] = (grow < p.M && col < p.out_dim) ? p.dy[(size_t)grow * p.out_dim + col] * scale : 0.f;
                }
                st_chunk_bf16(act + (size_t)idx * 16, v);
                if (p.dys[last]) st_chunk_bf16(p.dys[last] + ((size_t)tile * p.cout * kTileRows + idx) * 8, v);
                if (p.db_last && chunk * 8 < p.out_dim) {
                    // a warp holds 32 consecutive rows of this chunk: fp32 column sums, one atomic per column
#pragma unroll
                    for (int i = 0; i < 8; ++i) {
                        const float s = warp_sum(v[i]);
                        if ((threadIdx.x & 31) == 0 && chunk * 8 + i < p.out_dim) atomicAdd(p.db_last + chunk * 8 + i, s);
                    }
                }
            }
            es.release();
            const int grow = row0 + row;
            const bool live = grow < p.M;
            // the next tile's saved activations are pulled into L2 while this tile is processed
            {
                const int nti = ti + gridDim.x, nxt = p.reverse_tiles ? tiles - 1 - nti : nti;
                if (nti < tiles && threadIdx.x == 0)
                    for (int l = 0; l < last; ++l) l2_prefetch(tcm_at(p.sh[l], p.ch, nxt, 0, 0), (size_t)p.ch * kChunkBytes);
            }
            for (int l = last; l >= 1; --l) {
                const int per = (p.ch + kEpiGroups - 1) / kEpiGroups;
                const int cbeg = grp * per, cend = min(p.ch, cbeg + per);
                const __nv_bfloat16* hs = p.sh[l - 1];
                __nv_bfloat16* dsv = p.dys[l - 1];
                epi_for_chunks_saved(es, tmem_base, 0, cbeg, cend,
                    [&](int c) { return *reinterpret_cast<const uint4*>(tcm_at(hs, p.ch, tile, c, row)); },
                    [&](int c, float (&v)[8], const uint4 hq) {
                        elu_bwd8(v, hq, live);
                        const uint4 q = pack_bf16x8(v);
                        *reinterpret_cast<uint4*>(act + ((size_t)c * kTileRows + row) * 16) = q;
                        if (dsv) *reinterpret_cast<uint4*>(tcm_at(dsv, p.ch, tile, c, row)) = q;
                    });
                if (l > 1 || want_dx) es.release();
            }
            if (want_dx) {
                es.wait_acc();
                if (p.tile_wait) {  // the previous head's rows of this tile (we accumulate into them)
                    if ((threadIdx.x & 31) == 0) {
                        wait_counter_ge(p.tile_wait + tile, 1, 100);
                    }
                    __syncwarp();
                }
                // accumulator columns: [x0: c0 chunks][x1: c1 chunks] (the dX weight image is laid out per segment)
                for (int c = grp; c < p.c0 + p.c1; c += kEpiGroups) {
                    float v[8];
                    tmem_ld8(epi_taddr(tmem_base, c * 8), v);
                    tmem_ld_wait();
                    if (!live) continue;
                    if (c < p.c0) { if (p.dx0) stg_row8(p.dx0 + (size_t)grow * p.ld0, c * 8, p.k0, v, p.accumulate_dx != 0); }
                    else if (p.dx1) stg_row8(p.dx1 + (size_t)grow * p.ld1, (c - p.c0) * 8, p.k1, v, p.accumulate_dx != 0);
                }
                if (p.tile_done || p.step_rows_done) {
                    named_bar<kEpiThreads, 1>();  // every thread's dx stores ...
                    if (threadIdx.x == 0) {
                        __threadfence();          // ... are released by one cumulative gpu-scope fence before the signal
                        if (p.tile_done) atomicExch(p.tile_done + tile, 1);
                        if (p.step_rows_done) {
                            const int r_end = min(p.M, row0 + kTileRows);
                            for (int r = row0; r < r_end;) {
                                const int s_ = min(r / p.step_rows, p.step_count - 1);
                                const int e = s_ == p.step_count - 1 ? r_end : min(r_end, (s_ + 1) * p.step_rows);
                                atomicAdd(p.step_rows_done + s_, e - r);
                                r = e;
                            }
                        }
                    }
                }
            }
        }
    }
    chain_teardown(tmem_base);
}

// ------------------------------------------------------------------------------------------------
// host side
// ------------------------------------------------------------------------------------------------
static int n_parts(int npad, int part[2]) {
    if (npad <= 256) { part[0] = npad; part[1] = 0; return 1; }
    part[0] = rup(npad / 2, 16); part[1] = npad - part[0];
    return 2;
}

struct MlpTcPlan {
    MlpTcDev dev;
    PackTcJobs jobs;
    size_t wbytes;
    size_t smem_bytes;
    int tiles;
};

int mlp_tc_supported(const drm_mlp_call* c) {
    const drm_mlp& m = c->mlp;
    const int cin = rup((c->k0 + 7) / 8 + (c->k1 + 7) / 8, 2);
    const int ch = rup(((m.n_linear > 1 ? m.hidden : 0) + 7) / 8, 2);
    return ch * 8 <= 512 && rup(m.out_dim, 16) <= 512 && cin * 8 <= 512;
}

static int plan_dims(const drm_mlp_call* c, MlpTcPlan& pl) {
    const drm_mlp& m = c->mlp;
    MlpTcDev& d = pl.dev;
    memset(&d, 0, sizeof(d));
    d.M = c->M; d.k0 = c->k0; d.k1 = c->k1;
    d.c0 = (c->k0 + 7) / 8; d.c1 = (c->k1 + 7) / 8; d.cin = rup(d.c0 + d.c1, 2);
    d.n_linear = m.n_linear; d.hidden = m.n_linear > 1 ? m.hidden : 0; d.ch = rup((d.hidden + 7) / 8, 2);
    d.out_dim = m.out_dim; d.cout = rup((m.out_dim + 7) / 8, 2);
    pl.tiles = (c->M + kTileRows - 1) / kTileRows;
    DRM_REQUIRE(d.ch * 8 <= 512 && rup(m.out_dim, 16) <= 512 && d.cin * 8 <= 512, DRM_E_SHAPE,
                "layer wider than 512 is not supported by the bf16 path");
    pl.dev.pg.n_ops = 0;
    pl.jobs.n = 0;
    pl.wbytes = 0;
    return DRM_OK;
}

// Appends one GEMM phase  D[128 x npad] = A(act buffer, kchunks) * image^T  and its weight-image jobs.
static int add_phase(MlpTcPlan& pl, uint8_t* wbase, int kchunks, int npad, const float* src, int ld, int transposed,
                     int n_valid_total, int nseg, const int* seg_kstart, const int* seg_len, const int* seg_src0) {
    TcProgram& pg = pl.dev.pg;
    int part[2];
    const int np = n_parts(npad, part);
    int n0 = 0;
    for (int q = 0; q < np; ++q) {
        DRM_REQUIRE(pg.n_ops < kMaxOps && pl.jobs.n < kMaxPackTcJobs, DRM_E_SHAPE, "MLP too deep for the bf16 path");
        TcOp& op = pg.op[pg.n_ops++];
        op.a_off[0] = 0; op.a_off[1] = 0;  // patched once act_off is known
        op.a_chunks[0] = (uint16_t)kchunks; op.a_chunks[1] = 0;
        op.n = (uint16_t)part[q]; op.d_col = (uint16_t)n0; op.accumulate = 0;
        op.phase_end = (q == np - 1) ? 1 : 0;
        PackTcJob& j = pl.jobs.j[pl.jobs.n++];
        j.src = src; j.dst = wbase ? reinterpret_cast<__nv_bfloat16*>(wbase + pl.wbytes) : nullptr;
        j.ld = ld; j.transposed = transposed;
        int nv = n_valid_total - n0; if (nv < 0) nv = 0; if (nv > part[q]) nv = part[q];
        j.n0 = n0; j.n_valid = nv; j.n_pad = part[q];
        j.n2_start = 0; j.n2_valid = 0; j.n2_src0 = 0;
        j.k_chunks = kchunks; j.nseg = nseg;
        for (int s = 0; s < nseg; ++s) { j.seg_kstart[s] = seg_kstart[s]; j.seg_len[s] = seg_len[s]; j.seg_src0[s] = seg_src0[s]; }
        pl.wbytes += (size_t)kchunks * part[q] * 16;
        n0 += part[q];
    }
    pg.n_phases++;
    return DRM_OK;
}

static int finish_plan(MlpTcPlan& pl, int act_chunks, bool double_act = false) {
    MlpTcDev& d = pl.dev;
    TcProgram& pg = d.pg;
    d.bias_off = 256;
    int widest = d.ch * 8;
    if (rup(d.out_dim, 16) > widest) widest = rup(d.out_dim, 16);
    d.bstride = widest;
    d.act_off = (uint32_t)rup_sz(d.bias_off + (size_t)d.n_linear * d.bstride * sizeof(float), 1024);
    pg.a_alt = 0;
    pg.stage_bytes = 32 * 1024;
    if (double_act && d.act_off + 2 * (size_t)act_chunks * kChunkBytes + 2 * pg.stage_bytes <= kMaxSmem)
        pg.a_alt = (uint32_t)act_chunks * kChunkBytes;  // second operand buffer: the next tile is staged under this one
    pg.stage_off = (uint32_t)rup_sz(d.act_off + (size_t)act_chunks * kChunkBytes + pg.a_alt, 1024);
    int stages = (int)(((size_t)kMaxSmem - pg.stage_off) / pg.stage_bytes);
    if (stages > 4) stages = 4;
    DRM_REQUIRE(stages >= 2, DRM_E_SHAPE, "MLP too wide for the bf16 path's shared memory");
    pg.n_stages = stages;
    pl.smem_bytes = pg.stage_off + (size_t)stages * pg.stage_bytes;
    for (int i = 0; i < pg.n_ops; ++i) pg.op[i].a_off[0] = d.act_off;
    return DRM_OK;
}

static int plan_fwd(const drm_mlp_call* c, uint8_t* wbase, MlpTcPlan& pl) {
    const drm_mlp& m = c->mlp;
    int rc = plan_dims(c, pl); if (rc) return rc;
    MlpTcDev& d = pl.dev;
    d.pg.n_phases = 0;
    for (int l = 0; l < m.n_linear; ++l) {
        const bool lastl = l == m.n_linear - 1;
        const int nout = lastl ? m.out_dim : m.hidden;
        const int npad = lastl ? rup(nout, 16) : d.ch * 8;
        if (l == 0) {
            const int ks[2] = {0, d.c0 * 8}, len[2] = {c->k0, c->k1}, s0[2] = {0, c->k0};
            rc = add_phase(pl, wbase, d.cin, npad, m.W[0], m.in_dim, 0, nout, c->k1 > 0 ? 2 : 1, ks, len, s0);
        } else {
            const int ks[1] = {0}, len[1] = {m.hidden}, s0[1] = {0};
            rc = add_phase(pl, wbase, d.ch, npad, m.W[l], m.hidden, 0, nout, 1, ks, len, s0);
        }
        if (rc) return rc;
    }
    return finish_plan(pl, d.cin > d.ch ? d.cin : d.ch, true);
}

static int plan_bwd(const drm_mlp_call* c, bool want_dx, uint8_t* wbase, MlpTcPlan& pl) {
    const drm_mlp& m = c->mlp;
    int rc = plan_dims(c, pl); if (rc) return rc;
    MlpTcDev& d = pl.dev;
    d.pg.n_phases = 0;
    const int last = m.n_linear - 1;
    for (int l = last; l >= 1; --l) {
        const int nout = l == last ? m.out_dim : m.hidden;
        const int ks[1] = {0}, len[1] = {nout}, s0[1] = {0};
        rc = add_phase(pl, wbase, l == last ? d.cout : d.ch, d.ch * 8, m.W[l], m.hidden, 1, m.hidden, 1, ks, len, s0);
        if (rc) return rc;
    }
    if (want_dx) {
        const int rows0 = m.n_linear == 1 ? m.out_dim : m.hidden;
        const int ks[1] = {0}, len[1] = {rows0}, s0[1] = {0};
        rc = add_phase(pl, wbase, m.n_linear == 1 ? d.cout : d.ch, d.cin * 8, m.W[0], m.in_dim, 1, c->k0, 1, ks, len, s0);
        if (rc) return rc;
        // second N segment: x1's columns start at accumulator column c0 * 8
        {
            int q0 = pl.jobs.n - (d.cin * 8 > 256 ? 2 : 1);
            int n0 = 0;
            for (int q = q0; q < pl.jobs.n; ++q) {
                PackTcJob& j = pl.jobs.j[q];
                // image row n (global accumulator column n0 + n) <- source column: x0 part [0,k0) at cols [0,k0),
                // x1 part [k0,k0+k1) at cols [c0*8, c0*8+k1)
                const int lo = n0, hi = n0 + j.n_pad;
                int v0 = c->k0 - lo; if (v0 < 0) v0 = 0; if (v0 > j.n_pad) v0 = j.n_pad;
                j.n0 = lo; j.n_valid = v0;                       // columns of x0 inside this part
                int s1 = d.c0 * 8 > lo ? d.c0 * 8 : lo;            // first x1 accumulator column inside this part
                int e1 = d.c0 * 8 + c->k1 < hi ? d.c0 * 8 + c->k1 : hi;
                if (e1 > s1) { j.n2_start = s1 - lo; j.n2_valid = e1 - s1; j.n2_src0 = c->k0 + (s1 - d.c0 * 8); }
                n0 += j.n_pad;
            }
        }
    }
    return finish_plan(pl, d.cout > d.ch ? d.cout : d.ch);
}

// save buffer of the bf16 path: [input image: tiles x cin][hidden images: (n_linear-1) x tiles x ch] chunks
size_t mlp_tc_save_bytes(const drm_mlp_call* c) {
    MlpTcPlan pl;
    if (plan_dims(c, pl) != DRM_OK) return 0;
    return tcm_bytes(pl.tiles, pl.dev.cin) + (size_t)(c->mlp.n_linear - 1) * tcm_bytes(pl.tiles, pl.dev.ch) + 256;
}

static size_t dys_bytes(const MlpTcPlan& pl, int n_linear) {
    return (size_t)(n_linear - 1) * tcm_bytes(pl.tiles, pl.dev.ch) + tcm_bytes(pl.tiles, pl.dev.cout);
}

size_t mlp_tc_workspace_bytes(const drm_mlp_call* c) {
    MlpTcPlan pf, pb;
    if (plan_fwd(c, nullptr, pf) != DRM_OK || plan_bwd(c, true, nullptr, pb) != DRM_OK) return 0;
    const size_t f = rup_sz(pf.wbytes, 256);
    const size_t b = rup_sz(pb.wbytes, 256) + dys_bytes(pb, c->mlp.n_linear);
    return (f > b ? f : b) + 512;
}

static void carve_save(const drm_mlp_call* c, const MlpTcPlan& pl, MlpTcDev& d) {
    uint8_t* s = reinterpret_cast<uint8_t*>(rup_sz((size_t)c->save, 256));
    d.sx = reinterpret_cast<__nv_bfloat16*>(s);
    s += tcm_bytes(pl.tiles, d.cin);
    for (int l = 0; l < c->mlp.n_linear - 1; ++l) {
        d.sh[l] = reinterpret_cast<__nv_bfloat16*>(s);
        s += tcm_bytes(pl.tiles, d.ch);
    }
}

int mlp_tc_fwd(const drm_mlp_call* c, cudaStream_t stream) {
    const drm_mlp& m = c->mlp;
    MlpTcPlan pl;
    uint8_t* ws = reinterpret_cast<uint8_t*>(rup_sz((size_t)c->workspace, 256));
    int rc = plan_fwd(c, ws, pl); if (rc) return rc;
    DRM_REQUIRE(c->workspace_bytes >= mlp_tc_workspace_bytes(c), DRM_E_WORKSPACE, "workspace too small (bf16 path)");
    MlpTcDev& d = pl.dev;
    d.x0 = c->x0; d.ld0 = c->ld0; d.x1 = c->x1; d.ld1 = c->ld1; d.y = c->y;
    for (int l = 0; l < m.n_linear; ++l) d.b[l] = m.b[l];
    d.wstream = ws;
    if (c->training) {
        DRM_REQUIRE(c->save && c->save_bytes >= mlp_tc_save_bytes(c), DRM_E_WORKSPACE, "save buffer too small (bf16 path)");
        carve_save(c, pl, d);
    }
    rc = launch_pack_tc(pl.jobs, stream); if (rc) return rc;
    int cap = num_sms();
    if (c->max_ctas > 0 && c->max_ctas < cap) cap = c->max_ctas;
    const int grid = pl.tiles < cap ? pl.tiles : cap;
    d.wait_flags = c->wait_flags; d.wait_target = c->wait_target; d.wait_rows = c->wait_rows; d.wait_steps = c->wait_steps;
    if (d.wait_flags) DRM_REQUIRE(d.wait_rows > 0 && d.wait_steps > 0, DRM_E_ARG, "wait_flags need wait_rows and wait_steps");
    DRM_CUDA_OK(cudaFuncSetAttribute(mlp_fwd_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)pl.smem_bytes));
    ProfScope prof(PROF_MLP_FWD, stream);
    mlp_fwd_tc_kernel<<<grid, kChainThreads, pl.smem_bytes, stream>>>(d);
    DRM_CUDA_OK(cudaGetLastError());
    return DRM_OK;
}

int mlp_tc_bwd(const drm_mlp_bwd_call* bc, cudaStream_t stream) {
    const drm_mlp_call* c = &bc->fwd;
    const drm_mlp& m = c->mlp;
    const int last = m.n_linear - 1;
    const bool want_dx = bc->dx0 || bc->dx1;
    MlpTcPlan pl;
    uint8_t* ws = reinterpret_cast<uint8_t*>(rup_sz((size_t)c->workspace, 256));
    int rc = plan_bwd(c, want_dx, ws, pl); if (rc) return rc;
    DRM_REQUIRE(c->workspace_bytes >= mlp_tc_workspace_bytes(c), DRM_E_WORKSPACE, "workspace too small (bf16 path)");
    DRM_REQUIRE(c->save && c->save_bytes >= mlp_tc_save_bytes(c), DRM_E_WORKSPACE, "save buffer missing (bf16 path)");
    MlpTcDev& d = pl.dev;
    d.ld0 = c->ld0; d.ld1 = c->ld1;
    d.wstream = ws;
    carve_save(c, pl, d);
    d.dy = bc->dy; d.dy_scale = bc->dy_scale; d.dx0 = bc->dx0; d.dx1 = bc->dx1; d.accumulate_dx = bc->accumulate_dx;
    d.reverse_tiles = bc->reverse_tiles; d.tile_done = bc->tile_done; d.tile_wait = bc->tile_wait;
    d.step_rows_done = bc->step_rows_done; d.step_rows = bc->step_rows; d.step_count = bc->step_count;
    DRM_REQUIRE(!d.step_rows_done || (d.step_rows > 0 && d.step_count > 0), DRM_E_ARG, "step_rows_done needs step_rows / step_count");
    DRM_REQUIRE(!(d.tile_done || d.tile_wait || d.step_rows_done) || want_dx, DRM_E_ARG, "hand-over counters need dx0 / dx1");
    if (bc->want_wgrad) {
        d.db_last = bc->grad.db[last];
        uint8_t* s = ws + rup_sz(pl.wbytes, 256);
        for (int l = 0; l <= last; ++l) {
            d.dys[l] = reinterpret_cast<__nv_bfloat16*>(s);
            s += tcm_bytes(pl.tiles, l == last ? d.cout : d.ch);
        }
    }
    if (!want_dx && !bc->want_wgrad) return DRM_OK;
    if (bc->phase != DRM_PHASE_RUN) {
        if (bc->want_wgrad) {
            for (int l = 0; l <= last; ++l) {
                const int nout = l == last ? m.out_dim : m.hidden;
                const int nin = l == 0 ? m.in_dim : m.hidden;
                pl.jobs.zero(bc->grad.dW[l], (size_t)nout * nin);
                pl.jobs.zero(bc->grad.db[l], (size_t)nout);
            }
        }
        rc = launch_pack_tc(pl.jobs, stream); if (rc) return rc;
        if (bc->phase == DRM_PHASE_PREPARE) return DRM_OK;
    }
    {
        int grid = pl.tiles < num_sms() ? pl.tiles : num_sms();
        if (c->max_ctas > 0 && c->max_ctas < grid) grid = c->max_ctas;
        DRM_CUDA_OK(cudaFuncSetAttribute(mlp_bwd_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)pl.smem_bytes));
        ProfScope prof(PROF_MLP_BWD, stream);
        mlp_bwd_tc_kernel<<<grid, kChainThreads, pl.smem_bytes, stream>>>(d);
        DRM_CUDA_OK(cudaGetLastError());
    }
    if (bc->want_wgrad) {
        WgParams w;
        memset(&w, 0, sizeof(w));
        w.n_layers = m.n_linear; w.row_tiles = pl.tiles;
        for (int l = 0; l <= last; ++l) {
            WgLayer& L = w.L[l];
            L.dy = d.dys[l]; L.dy_cpt = l == last ? d.cout : d.ch;
            L.nout = l == last ? m.out_dim : m.hidden;
            if (l == 0) {
                L.nseg = c->k1 > 0 ? 2 : 1;
                L.seg[0] = WgSeg{d.sx, d.cin, 0, d.c0, 0, c->k0};
                L.seg[1] = WgSeg{d.sx, d.cin, d.c0, d.c1, c->k0, c->k1};
                L.ld_dw = m.in_dim;
            } else {
                L.nseg = 1;
                L.seg[0] = WgSeg{d.sh[l - 1], d.ch, 0, (m.hidden + 7) / 8, 0, m.hidden};
                L.ld_dw = m.hidden;
            }
            L.dW = bc->grad.dW[l]; L.db = l == last ? nullptr : bc->grad.db[l];
        }
        wgrad_tc_plan(w);
        rc = launch_wgrad_tc(w, stream); if (rc) return rc;
    }
    return DRM_OK;
}

}  // namespace tc
void warm_mlp_tc() { warm_kernel(tc::pack_tc_kernel); warm_kernel(tc::mlp_fwd_tc_kernel); warm_kernel(tc::mlp_bwd_tc_kernel); }

}  // namespace drm

#ifdef DRM_TIMELINE
// Debug only: handshake stamps CTA 0 of the batched-MLP kernels recorded (see tools/timeline.py).
extern "C" int drm_debug_timeline_mlp(long long* stamps, int* counts) {
    cudaDeviceSynchronize();
    cudaMemcpyFromSymbol(stamps, drm::tc::g_tl, sizeof(long long) * 4 * 4096);
    cudaMemcpyFromSymbol(counts, drm::tc::g_tl_n, sizeof(int) * 4);
    int zero[4] = {0, 0, 0, 0};
    cudaMemcpyToSymbol(drm::tc::g_tl_n, zero, sizeof(zero));
    return 0;
}
#endif
