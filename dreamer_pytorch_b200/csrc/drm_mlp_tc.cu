// bf16 tensor-core (tcgen05) path of the batched MLP: DenseModel / ActionDecoder feed-forward stacks
// (reference: dreamer/models/dense.py:28-43, dreamer/models/action.py:37-52).  Operands are bf16, the
// accumulators fp32 in TMEM, bias / ELU / everything pointwise fp32.  See drm_tc_chain.cuh for the engine.
#include "drm_mlp.cuh"
#include "drm_tc_chain.cuh"

namespace drm {
namespace tc {

// ------------------------------------------------------------------------------------------------
// weight image packing
// ------------------------------------------------------------------------------------------------
__global__ void pack_tc_kernel(PackTcJobs jobs) {
    const PackTcJob j = jobs.j[blockIdx.y];
    const int total = j.k_chunks * j.n_pad * 8;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
        const int kk = i & 7;
        const int n = (i >> 3) % j.n_pad;
        const int chunk = (i >> 3) / j.n_pad;
        const int k = chunk * 8 + kk;
        float v = 0.f;
        if (n < j.n_valid) {
            for (int s = 0; s < j.nseg; ++s) {
                const int d = k - j.seg_kstart[s];
                if (d >= 0 && d < j.seg_len[s]) {
                    const int ks = j.seg_src0[s] + d;
                    v = j.transposed ? j.src[(size_t)ks * j.ld + j.n0 + n] : j.src[(size_t)(j.n0 + n) * j.ld + ks];
                }
            }
        }
        j.dst[i] = __float2bfloat16_rn(v);
    }
}

int launch_pack_tc(const PackTcJobs& jobs, cudaStream_t stream) {
    if (jobs.n == 0) return DRM_OK;
    DRM_REQUIRE(jobs.n <= kMaxPackTcJobs, DRM_E_SHAPE, "too many tensor-core pack jobs");
    dim3 grid(24, jobs.n);
    ProfScope prof(PROF_PACK, stream);
    pack_tc_kernel<<<grid, 256, 0, stream>>>(jobs);
    DRM_CUDA_OK(cudaGetLastError());
    return DRM_OK;
}

// ------------------------------------------------------------------------------------------------
// forward
// ------------------------------------------------------------------------------------------------
struct MlpTcDev {
    int M, k0, k1, c0, c1, cin;  // input segments: widths and 8-wide chunk counts (cin = c0 + c1 rounded up to even)
    int n_linear, hidden, ch, out_dim;
    const float* x0; int64_t ld0;
    const float* x1; int64_t ld1;
    float* y;
    const float* b[DRM_MAX_LINEAR];
    const uint8_t* wstream;
    __nv_bfloat16* save_h;  // [n_linear-1][M][ldh] or nullptr
    int ldh;
    uint32_t act_off;       // shared-memory offset of the activation buffer
    // backward only
    const float* dy; const float* dy_scale;
    float* dx0; float* dx1; int accumulate_dx;
    __nv_bfloat16* dysave[DRM_MAX_LINEAR];  // [M][ld_l] bf16 or nullptr
    int ld_dys[DRM_MAX_LINEAR];
    TcProgram pg;
};

__device__ __forceinline__ float elu_fast(float x) { return x > 0.f ? x : __expf(x) - 1.f; }

__global__ void __launch_bounds__(kChainThreads, 1) mlp_fwd_tc_kernel(const __grid_constant__ MlpTcDev p) {
    extern __shared__ __align__(1024) uint8_t smem[];
    ChainBars* bars = reinterpret_cast<ChainBars*>(smem);
    const uint32_t tmem_base = chain_setup(bars, p.pg);
    const int warp = threadIdx.x >> 5;
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
        for (int tile = blockIdx.x; tile < tiles; tile += gridDim.x) {
            const int row0 = tile * kTileRows;
            // ---- stage the input tile: fp32 global -> bf16 chunk-major shared
            for (int idx = threadIdx.x; idx < p.cin * kTileRows; idx += kEpiThreads) {
                const int chunk = idx / kTileRows, r = idx - chunk * kTileRows;
                const int grow = row0 + r;
                float v[8];
#pragma unroll
                for (int i = 0; i < 8; ++i) v[i] = 0.f;
                if (grow < p.M) {
                    if (chunk < p.c0) {
                        const float* src = p.x0 + (size_t)grow * p.ld0 + chunk * 8;
#pragma unroll
                        for (int i = 0; i < 8; ++i) if (chunk * 8 + i < p.k0) v[i] = __ldg(src + i);
                    } else if (chunk < p.c0 + p.c1) {
                        const int cc = chunk - p.c0;
                        const float* src = p.x1 + (size_t)grow * p.ld1 + cc * 8;
#pragma unroll
                        for (int i = 0; i < 8; ++i) if (cc * 8 + i < p.k1) v[i] = __ldg(src + i);
                    }
                }
                st_chunk_bf16(act + (size_t)idx * 16, v);
            }
            es.release();
            for (int l = 0; l <= last; ++l) {
                es.wait_acc();
                const float* bias = p.b[l];
                if (l < last) {
                    // hidden layer: h = ELU(acc + b) -> bf16 A operand of the next layer (in place) [+ save]
                    const int per = (p.ch + kEpiGroups - 1) / kEpiGroups;
                    const int cbeg = grp * per, cend = min(p.ch, cbeg + per);
                    __nv_bfloat16* sv = p.save_h ? p.save_h + (size_t)l * p.M * p.ldh : nullptr;
                    const int grow = row0 + row;
                    for (int c = cbeg; c < cend; c += 2) {
                        float v0[8], v1[8];
                        tmem_ld8(epi_taddr(tmem_base, c * 8), v0);
                        const bool two = c + 1 < cend;
                        if (two) tmem_ld8(epi_taddr(tmem_base, c * 8 + 8), v1);
                        tmem_ld_wait();
#pragma unroll
                        for (int i = 0; i < 8; ++i) {
                            const int col = c * 8 + i;
                            v0[i] = col < p.hidden ? elu_fast(v0[i] + __ldg(bias + col)) : 0.f;
                        }
                        st_chunk_bf16(act + ((size_t)c * kTileRows + row) * 16, v0);
                        if (sv && grow < p.M) st_chunk_bf16(sv + (size_t)grow * p.ldh + c * 8, v0);
                        if (two) {
#pragma unroll
                            for (int i = 0; i < 8; ++i) {
                                const int col = c * 8 + 8 + i;
                                v1[i] = col < p.hidden ? elu_fast(v1[i] + __ldg(bias + col)) : 0.f;
                            }
                            st_chunk_bf16(act + ((size_t)(c + 1) * kTileRows + row) * 16, v1);
                            if (sv && grow < p.M) st_chunk_bf16(sv + (size_t)grow * p.ldh + c * 8 + 8, v1);
                        }
                    }
                    es.release();
                } else {
                    // output layer: y = acc + b (fp32, row-major)
                    const int oc = (p.out_dim + 7) / 8;
                    const int grow = row0 + row;
                    for (int c = grp; c < oc; c += kEpiGroups) {
                        float v[8];
                        tmem_ld8(epi_taddr(tmem_base, c * 8), v);
                        tmem_ld_wait();
                        if (grow < p.M) {
#pragma unroll
                            for (int i = 0; i < 8; ++i) {
                                const int col = c * 8 + i;
                                if (col < p.out_dim) p.y[(size_t)grow * p.out_dim + col] = v[i] + (bias ? __ldg(bias + col) : 0.f);
                            }
                        }
                    }
                    // the next tile's input staging is followed by release(), which also covers these loads
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
};

// Lays out the forward program, its weight images (at wbase, may be null for sizing) and shared memory.
static int plan_fwd(const drm_mlp_call* c, uint8_t* wbase, MlpTcPlan& pl) {
    const drm_mlp& m = c->mlp;
    MlpTcDev& d = pl.dev;
    memset(&d, 0, sizeof(d));
    d.M = c->M; d.k0 = c->k0; d.k1 = c->k1;
    d.c0 = (c->k0 + 7) / 8; d.c1 = (c->k1 + 7) / 8; d.cin = rup(d.c0 + d.c1, 2);
    d.n_linear = m.n_linear; d.hidden = m.n_linear > 1 ? m.hidden : 0; d.ch = rup((d.hidden + 7) / 8, 2);
    d.out_dim = m.out_dim;
    d.ldh = rup(d.hidden > 0 ? d.hidden : 8, 8);
    DRM_REQUIRE(d.ch * 8 <= 512 && rup(m.out_dim, 16) <= 512, DRM_E_SHAPE, "layer wider than 512 is not supported by the bf16 path");
    TcProgram& pg = pl.dev.pg;
    pg.n_ops = 0; pg.n_phases = m.n_linear;
    pl.jobs.n = 0;
    size_t woff = 0;
    for (int l = 0; l < m.n_linear; ++l) {
        const bool lastl = l == m.n_linear - 1;
        const int nout = lastl ? m.out_dim : m.hidden;
        const int npad = lastl ? rup(nout, 16) : d.ch * 8;
        const int kchunks = l == 0 ? d.cin : d.ch;
        int part[2];
        const int np = n_parts(npad, part);
        int n0 = 0;
        for (int q = 0; q < np; ++q) {
            DRM_REQUIRE(pg.n_ops < kMaxOps && pl.jobs.n < kMaxPackTcJobs, DRM_E_SHAPE, "MLP too deep for the bf16 path");
            TcOp& op = pg.op[pg.n_ops++];
            op.a_off[0] = 0; op.a_off[1] = 0;  // patched below once act_off is known
            op.a_chunks[0] = (uint16_t)kchunks; op.a_chunks[1] = 0;
            op.n = (uint16_t)part[q]; op.d_col = (uint16_t)n0; op.accumulate = 0;
            op.phase_end = (q == np - 1) ? 1 : 0;
            PackTcJob& j = pl.jobs.j[pl.jobs.n++];
            j.src = m.W[l]; j.dst = wbase ? reinterpret_cast<__nv_bfloat16*>(wbase + woff) : nullptr;
            j.ld = l == 0 ? m.in_dim : m.hidden; j.transposed = 0;
            j.n0 = n0; j.n_valid = nout - n0 < part[q] ? (nout - n0 > 0 ? nout - n0 : 0) : part[q]; j.n_pad = part[q];
            j.k_chunks = kchunks;
            if (l == 0) {
                j.nseg = c->k1 > 0 ? 2 : 1;
                j.seg_kstart[0] = 0; j.seg_len[0] = c->k0; j.seg_src0[0] = 0;
                j.seg_kstart[1] = d.c0 * 8; j.seg_len[1] = c->k1; j.seg_src0[1] = c->k0;
            } else {
                j.nseg = 1; j.seg_kstart[0] = 0; j.seg_len[0] = m.hidden; j.seg_src0[0] = 0;
            }
            woff += (size_t)kchunks * part[q] * 16;
            n0 += part[q];
        }
    }
    pl.wbytes = woff;
    // shared memory: [bars 256 B][pad to 1024][activation buffer][stages]
    const int act_chunks = d.cin > d.ch ? d.cin : d.ch;
    d.act_off = 1024;
    const size_t act_bytes = (size_t)act_chunks * kChunkBytes;
    pg.stage_off = (uint32_t)rup_sz(d.act_off + act_bytes, 1024);
    pg.stage_bytes = 32 * 1024;
    int stages = (int)(((size_t)kMaxSmem - pg.stage_off) / pg.stage_bytes);
    if (stages > 4) stages = 4;
    DRM_REQUIRE(stages >= 2, DRM_E_SHAPE, "MLP too wide for the bf16 path's shared memory");
    pg.n_stages = stages;
    pl.smem_bytes = pg.stage_off + (size_t)stages * pg.stage_bytes;
    for (int i = 0; i < pg.n_ops; ++i) pg.op[i].a_off[0] = d.act_off;
    return DRM_OK;
}

size_t mlp_tc_fwd_workspace_bytes(const drm_mlp_call* c) {
    MlpTcPlan pl;
    if (plan_fwd(c, nullptr, pl) != DRM_OK) return 0;
    return rup_sz(pl.wbytes, 256) + 256;
}

int mlp_tc_fwd(const drm_mlp_call* c, cudaStream_t stream) {
    const drm_mlp& m = c->mlp;
    MlpTcPlan pl;
    uint8_t* ws = reinterpret_cast<uint8_t*>(rup_sz((size_t)c->workspace, 256));
    int rc = plan_fwd(c, ws, pl); if (rc) return rc;
    DRM_REQUIRE(c->workspace_bytes >= mlp_tc_fwd_workspace_bytes(c), DRM_E_WORKSPACE, "workspace too small (bf16 path)");
    MlpTcDev& d = pl.dev;
    d.x0 = c->x0; d.ld0 = c->ld0; d.x1 = c->x1; d.ld1 = c->ld1; d.y = c->y;
    for (int l = 0; l < m.n_linear; ++l) d.b[l] = m.b[l];
    d.wstream = ws;
    const bool save = c->training && m.n_linear > 1;
    d.save_h = save ? reinterpret_cast<__nv_bfloat16*>(c->save) : nullptr;
    rc = launch_pack_tc(pl.jobs, stream); if (rc) return rc;
    const int tiles = (c->M + kTileRows - 1) / kTileRows;
    const int grid = tiles < kNumSM ? tiles : kNumSM;
    DRM_CUDA_OK(cudaFuncSetAttribute(mlp_fwd_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)pl.smem_bytes));
    ProfScope prof(PROF_MLP_FWD, stream);
    mlp_fwd_tc_kernel<<<grid, kChainThreads, pl.smem_bytes, stream>>>(d);
    DRM_CUDA_OK(cudaGetLastError());
    return DRM_OK;
}

}  // namespace tc
}  // namespace drm
