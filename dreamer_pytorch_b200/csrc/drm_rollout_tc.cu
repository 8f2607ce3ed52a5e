// bf16 tensor-core (tcgen05) imagination rollout: RSSMRollout.rollout_policy (rnns.py:224-244) with
// AgentModel.policy (agent.py:101-117), ActionDecoder.forward (action.py:51-66) and RSSMTransition.forward
// (rnns.py:98-107) fused into ONE persistent kernel per direction.  A CTA owns a tile of 128 start states for
// all H steps; the recurrent state lives in shared memory as the bf16 A operand of the next GEMM (the fp32
// master copy of `deter` is the kernel's own output, re-read by the thread that wrote it); weights stream
// from L2 through the bulk-copy ring of drm_tc_chain.cuh; accumulators live in TMEM.
//
// Shared-memory operands (chunk-major, 128 rows):
//   ST  = [stoch: cs chunks][deter: cdp chunks (incl. zero pad)][action: ca chunks]
//   HB  = one hidden-activation buffer (actor hidden / rnn input x / prior hidden), rewritten in place
// Phases per step (an = Linear layers of the actor):
//   0..an-2  actor hidden layers          A = ST[stoch|deter] or HB      -> HB
//   an-1     actor output + action sample A = HB                        -> ST[action], action_out
//   an       x = ELU(W_in [a|stoch] + b)  A = ST[stoch], ST[action]      -> HB
//   an+1     r (cols 0..), h_n (cols Dp..) A = HB|ST[deter], ST[deter]   -> TMEM cols 0.. := r * h_n
//   an+2     i_n accumulated onto r*h_n, z A = HB, HB|ST[deter]          -> deter' -> ST[deter], deter out
//   an+3     prior hidden                 A = ST[deter]                  -> HB
//   an+4     prior mean | std, sample     A = HB                         -> ST[stoch], prior_* out
#include <stdlib.h>
#include <string.h>
#include "drm_rollout.cuh"
#include "drm_tc_chain.cuh"

namespace drm {
namespace tc {

constexpr int kMaxPerD = 8;  // deter chunks per epilogue thread (cdp <= 32)

struct RollTcDev {
    int T, n, ntiles, tr, act_dist, training;  // tr: rows per tile (64 or 128)
    int A, S, D, Hd, ah, an, Aout;
    int cs, cd, cdp, ca, chh, cha;  // chunk counts: stoch, deter (exact / even), action, Hd (even), actor hidden (even)
    int Sp;                         // column of the first std accumulator in the prior's last Linear (S rounded up to 8)
    float mean_scale, raw_init_std, act_min_std, temperature, min_std;
    const float *b_in, *b_ih, *b_hh, *b_p1, *b_p2;
    const float* ba[DRM_MAX_LINEAR];
    const float *start_stoch, *start_deter, *eps_prior, *noise_act;
    float *prior_mean, *prior_std, *prior_stoch, *deter, *action_out;
    // saves (TCM images indexed by tile-step ts = t * ntiles + tile, except raw)
    __nv_bfloat16* sv_feat;                 // [stoch|deter] fed into step t            (cs + cdp chunks)
    __nv_bfloat16* sv_ah[DRM_MAX_LINEAR];   // actor hidden activations                  (cha)
    __nv_bfloat16* sv_x;                    // ELU(rnn input)                            (chh)
    __nv_bfloat16 *sv_r, *sv_z, *sv_n, *sv_hn;  // GRU gates and W_hn h + b_hn           (cdp)
    __nv_bfloat16* sv_hid;                  // prior hidden                              (chh)
    __nv_bfloat16* sv_lo;                   // bf16(deter - bf16(deter)) of the deter fed into step t  (cdp);
                                            // not training: one slot per CTA (lo_per_cta = 1)
    int lo_per_cta;
    float* sv_raw;                          // actor output before the distribution, fp32 [T*n][Aout]
    const uint8_t* wstream;
    int* step_flags;                        // optional: += 1 per row tile once every output of step t is in global memory
    uint32_t st_off, hb_off, bias_off;
    // backward
    const float *g_prior_mean, *g_prior_std, *g_prior_stoch, *g_deter, *g_action_out;
    int* bwd_flags;                         // optional: += 1 per row tile once the dY images of step t are in global memory
    const int* g_wait_flags; int g_wait_target;  // optional: g_* rows of step t exist once g_wait_flags[t] >= g_wait_target
    __nv_bfloat16* dys_a[DRM_MAX_LINEAR];   // dY of every actor Linear, TCM by tile-step (cha chunks; 2 for the last)
    __nv_bfloat16* stash;                   // per-CTA scratch: [grid][2][cdp chunks] (d i_n, d h_n between two phases)
    float* db_a_last;                       // fp32 bias gradient of the actor's last Linear
    int want_actor;                         // 0: the actor phases are not part of the program
    uint32_t ga_off, gb_off, g2_off;        // backward operand buffers (GA doubles as the hidden buffer)
    TcProgram pg;
};

// ------------------------------------------------------------------------------------------------
// forward kernel
// ------------------------------------------------------------------------------------------------
// zero-padded fp32 bias blocks in shared memory (floats from bias_off)
struct BiasLayout {
    int ba0, ba_last, in, r, hn, n, z, p1, p2m, p2s, total;
};
__host__ __device__ inline BiasLayout bias_layout(int an, int Ap, int Hp, int Dp, int Sp) {
    BiasLayout b;
    b.ba0 = 0; b.ba_last = (an - 1) * Ap; b.in = b.ba_last + 16; b.r = b.in + Hp; b.hn = b.r + Dp; b.n = b.hn + Dp;
    b.z = b.n + Dp; b.p1 = b.z + Dp; b.p2m = b.p1 + Hp; b.p2s = b.p2m + Sp; b.total = b.p2s + Sp;
    return b;
}

template <int TR>
__global__ void __launch_bounds__(kChainThreads, 1) rollout_fwd_tc_kernel(const __grid_constant__ RollTcDev p) {
    // tile geometry of this instantiation (shadows the 128-row constants of drm_tc_chain.cuh)
    constexpr int kTileRows = TR;
    constexpr int kChunkBytes = TR * 16;
    constexpr bool kHalf = TR == 64;
    auto tcm_at = [](auto* base, int chunks_per_tile, int tile, int chunk, int row) {
        return base + (((size_t)tile * chunks_per_tile + chunk) * TR + row) * 8;
    };
    extern __shared__ __align__(1024) uint8_t smem[];
    ChainBars* bars = reinterpret_cast<ChainBars*>(smem);
    const int S = p.S, D = p.D, A = p.A;
    const int Dp = p.cdp * 8, Hp = p.chh * 8, Ap = p.cha * 8;
    float* sbias = reinterpret_cast<float*>(smem + p.bias_off);
    const BiasLayout bl = bias_layout(p.an, Ap, Hp, Dp, p.Sp);
    for (int l = 0; l < p.an - 1; ++l) fill_bias(sbias + bl.ba0 + l * Ap, p.ba[l], nullptr, p.ah, Ap);
    fill_bias(sbias + bl.ba_last, p.ba[p.an - 1], nullptr, p.Aout, 16);
    fill_bias(sbias + bl.in, p.b_in, nullptr, p.Hd, Hp);
    fill_bias(sbias + bl.r, p.b_ih, p.b_hh, D, Dp);
    fill_bias(sbias + bl.hn, p.b_hh + 2 * D, nullptr, D, Dp);
    fill_bias(sbias + bl.n, p.b_ih + 2 * D, nullptr, D, Dp);
    fill_bias(sbias + bl.z, p.b_ih + D, p.b_hh + D, D, Dp);
    fill_bias(sbias + bl.p1, p.b_p1, nullptr, p.Hd, Hp);
    fill_bias(sbias + bl.p2m, p.b_p2, nullptr, S, p.Sp);
    fill_bias(sbias + bl.p2s, p.b_p2 + S, nullptr, S, p.Sp);
    const uint32_t tmem_base = chain_setup(bars, p.pg);
    const int warp = __shfl_sync(0xffffffffu, threadIdx.x >> 5, 0);  // warp-uniform for the compiler
    const int my_tiles = (int)blockIdx.x < p.ntiles ? (p.ntiles - blockIdx.x + gridDim.x - 1) / gridDim.x : 0;
    const int reps = my_tiles * p.T;
    if (warp == kEpiWarps) {
        chain_producer(p.pg, bars, smem, p.wstream, reps);
    } else if (warp == kEpiWarps + 1) {
        chain_mma<TR>(p.pg, bars, smem, tmem_base, reps);
    } else {
        EpiSync es(bars);
        uint8_t* st = smem + p.st_off;
        uint8_t* hb = smem + p.hb_off;
        const EpiGeo<TR> geo;
        const int row = geo.row, grp = epi_group();
        const bool lead = geo.half == 0;  // half tiles: lanes 16..31 duplicate the rows of lanes 0..15
        const int cfeat = p.cs + p.cdp;
#ifdef DRM_NO_SAVE  // timing experiment only: the backward pass then reads garbage
        const bool save = false;
#else
        const bool save = p.training != 0;
#endif
        const size_t lo_stride = (size_t)p.cdp * kTileRows * 8;
        const int tid = threadIdx.x, lane = threadIdx.x & 31;
        float* hbf = reinterpret_cast<float*>(hb);
        // this thread's chunk ranges for the three operand widths
        // (ranges are TMEM chunk indices; x_go is this thread's offset to global chunk indices, see EpiGeo)
        auto rng = [&](int nchunks, int& beg, int& end, int& go) {
            const int nct = geo.tmem_chunks(nchunks), per = (nct + kEpiGroups - 1) / kEpiGroups;
            beg = min(nct, grp * per); end = min(nct, beg + per); go = geo.goff(nchunks);
        };
        int a_beg, a_end, a_go, h_beg, h_end, h_go, d_beg, d_end, d_go;
        rng(p.cha, a_beg, a_end, a_go);
        rng(p.chh, h_beg, h_end, h_go);
        rng(p.cdp, d_beg, d_end, d_go);
        const int DpT = kHalf ? Dp / 2 : Dp;  // TMEM columns of a D-wide accumulator
        // prior head: mean chunks; half tiles keep mean in lanes 0..15 and std in lanes 16..31 of the same columns
        const int per_s = (p.cs + kEpiGroups - 1) / kEpiGroups, s_beg = min(p.cs, grp * per_s), s_end = min(p.cs, s_beg + per_s);
        for (int tile = blockIdx.x; tile < p.ntiles; tile += gridDim.x) {
            const int row0 = tile * kTileRows;
            const int grow = row0 + row;
            const bool live = grow < p.n;
            // ---- start state -> ST (and the feat image of step 0); action chunks zero
            __nv_bfloat16* lo0 = p.sv_lo + (size_t)(p.lo_per_cta ? (int)blockIdx.x : tile) * lo_stride;
            for (int idx = threadIdx.x; idx < (cfeat + p.ca) * kTileRows; idx += kEpiThreads) {
                const int chunk = idx / kTileRows, r = idx - chunk * kTileRows;
                const int gr = row0 + r;
                float v[8];
#pragma unroll
                for (int i = 0; i < 8; ++i) v[i] = 0.f;
                if (gr < p.n) {
                    if (chunk < p.cs) ldg_row8(p.start_stoch + (size_t)gr * S, chunk * 8, S, v);
                    else if (chunk < p.cs + p.cd) ldg_row8(p.start_deter + (size_t)gr * D, (chunk - p.cs) * 8, D, v);
                }
                uint4 hi, lo;
                split_hi_lo(v, hi, lo);
                *reinterpret_cast<uint4*>(st + (size_t)idx * 16) = hi;
                if (save && chunk < cfeat) *reinterpret_cast<uint4*>(p.sv_feat + ((size_t)tile * cfeat * kTileRows + idx) * 8) = hi;
                if (chunk >= p.cs && chunk < cfeat)
                    *reinterpret_cast<uint4*>(lo0 + ((size_t)(chunk - p.cs) * kTileRows + r) * 8) = lo;
            }
            named_bar<kEpiThreads, 1>();  // the lo image is re-read by other threads (ownership differs from this loop's)
            es.release();
            for (int t = 0; t < p.T; ++t) {
                const int ts = t * p.ntiles + tile;
                // ------------------------------------------------------------ actor hidden layers
                for (int l = 0; l < p.an - 1; ++l) {
                    es.wait_acc();
                    const float* bias = sbias + bl.ba0 + l * Ap;
                    __nv_bfloat16* sv = save ? p.sv_ah[l] : nullptr;
                    epi_for_chunks(tmem_base, 0, a_beg, a_end, [&](int c, float (&v)[8]) {
                        float b[8];
                        lds8(bias + c * 8, b);
                        elu_bias8(v, b);
                        st_chunk_bf16(hb + ((size_t)c * kTileRows + row) * 16, v);
                        if (sv) st_chunk_bf16(tcm_at(sv, p.cha, ts, c, row), v);
                    }, a_go);
                    es.release();
                }
                // ------------------------------------------------------------ actor output + action (agent.py:101-117)
                // Global I/O of this small phase goes through shared memory (the hidden buffer is dead once the
                // actor's last GEMM has retired): the 128 threads of column group 0 read the tile's noise block with
                // coalesced loads before the wait, and write action / raw blocks back coalesced after the release.
                const int arows = min(kTileRows, p.n - row0);
                float* sN = hbf;                       // [128][A] noise
                float* sA = hbf + kTileRows * 16;      // [rows][A] action
                float* sR = hbf + 2 * kTileRows * 16;  // [rows][Aout] pre-distribution actor output
                constexpr int kG0 = 128;               // threads of column group 0
                float nzr[16];
                if (grp == 0) {
                    const float* nz_t = p.noise_act + ((size_t)t * p.n + row0) * A;
#pragma unroll
                    for (int k = 0; k < 16; ++k) {
                        const int i = tid + k * kG0;
                        nzr[k] = (k < A && i < arows * A) ? __ldg(nz_t + i) : 0.f;
                    }
                }
                es.wait_acc();
                if (grp == 0) {
#pragma unroll
                    for (int k = 0; k < 16; ++k) if (k < A && tid + k * kG0 < kTileRows * A) sN[tid + k * kG0] = nzr[k];
                    named_bar<kG0, 2>();
                    float raw[16];
                    {
                        float v0[8], v1[8], b0[8], b1[8];
                        tmem_ld8(epi_taddr(tmem_base, 0), v0);
                        if (!kHalf) tmem_ld8(epi_taddr(tmem_base, 8), v1);
                        lds8(sbias + bl.ba_last, b0);
                        lds8(sbias + bl.ba_last + 8, b1);
                        tmem_ld_wait();
                        if (kHalf) {
                            // columns 0..7 live in lanes 0..15, columns 8..15 in lanes 16..31 of the same TMEM columns
#pragma unroll
                            for (int i = 0; i < 8; ++i) {
                                const float o = __shfl_xor_sync(0xffffffffu, v0[i], 16);
                                v1[i] = lead ? o : v0[i];
                                v0[i] = lead ? v0[i] : o;
                            }
                        }
#pragma unroll
                        for (int i = 0; i < 8; ++i) { raw[i] = v0[i] + b0[i]; raw[8 + i] = v1[i] + b1[i]; }
                    }
                    float act[16];
#pragma unroll
                    for (int j = 0; j < 16; ++j) act[j] = 0.f;
                    if (live && lead) {
                        const float* nz = sN + row * A;
                        if (p.act_dist == DRM_DIST_TANH_NORMAL) {
#pragma unroll
                            for (int j = 0; j < 8; ++j) {
                                if (j < A) {
                                    const float rm = raw[j], rs = raw[A + j];
                                    const float mean = p.mean_scale * tanh_fast(rm / p.mean_scale);
                                    const float std = softplus_fast(rs + p.raw_init_std) + p.act_min_std;
                                    act[j] = tanh_fast(fmaf(std, nz[j], mean));
                                }
                            }
                        } else {
                            float mx = -INFINITY;
#pragma unroll
                            for (int j = 0; j < 16; ++j) if (j < A) mx = fmaxf(mx, raw[j]);
                            if (p.act_dist == DRM_DIST_ONE_HOT) {
                                float den = 0.f, bv = -INFINITY;
                                int best = 0;
#pragma unroll
                                for (int j = 0; j < 16; ++j) {
                                    if (j < A) {
                                        den += __expf(raw[j] - mx);
                                        const float s = raw[j] + nz[j];
                                        if (s > bv) { bv = s; best = j; }
                                    }
                                }
#pragma unroll
                                for (int j = 0; j < 16; ++j) {
                                    if (j < A) {
                                        const float pj = __expf(raw[j] - mx) / den;
                                        act[j] = ((j == best ? 1.f : 0.f) + pj) - pj;  // straight-through value
                                    }
                                }
                            } else {  // relaxed one-hot: softmax((logits + g) / temperature)
                                float smx = -INFINITY, den = 0.f;
#pragma unroll
                                for (int j = 0; j < 16; ++j) if (j < A) smx = fmaxf(smx, (raw[j] + nz[j]) / p.temperature);
#pragma unroll
                                for (int j = 0; j < 16; ++j) {
                                    if (j < A) { act[j] = __expf((raw[j] + nz[j]) / p.temperature - smx); den += act[j]; }
                                }
#pragma unroll
                                for (int j = 0; j < 16; ++j) if (j < A) act[j] /= den;
                            }
                        }
                    }
                    if (lead) {
#pragma unroll
                        for (int j = 0; j < 16; ++j) {
                            if (j < A) sA[row * A + j] = act[j];
                            if (j < p.Aout) sR[row * p.Aout + j] = raw[j];
                        }
                        uint8_t* sa = st + (size_t)cfeat * kChunkBytes;
                        for (int c = 0; c < p.ca; ++c) {
                            float v[8];
#pragma unroll
                            for (int i = 0; i < 8; ++i) v[i] = c == 0 ? act[i] : act[8 + i];
                            st_chunk_bf16(sa + ((size_t)c * kTileRows + row) * 16, v);
                        }
                    }
                }
                es.release();
                if (grp == 0) {
                    named_bar<kG0, 2>();
                    float* ao = p.action_out + ((size_t)t * p.n + row0) * A;
                    for (int i = tid; i < arows * A; i += kG0) ao[i] = sA[i];
                    if (save) {
                        float* ro = p.sv_raw + ((size_t)t * p.n + row0) * p.Aout;
                        for (int i = tid; i < arows * p.Aout; i += kG0) ro[i] = sR[i];
                    }
                }
                named_bar<kEpiThreads, 1>();  // the hidden buffer is written again by the next epilogue
                // ------------------------------------------------------------ x = ELU(W_in [a|stoch] + b)   (rnns.py:99-101)
                es.wait_acc();
                {
                    const float* bias = sbias + bl.in;
                    __nv_bfloat16* sv = save ? p.sv_x : nullptr;
                    epi_for_chunks(tmem_base, 0, h_beg, h_end, [&](int c, float (&v)[8]) {
                        float b[8];
                        lds8(bias + c * 8, b);
                        elu_bias8(v, b);
                        st_chunk_bf16(hb + ((size_t)c * kTileRows + row) * 16, v);
                        if (sv) st_chunk_bf16(tcm_at(sv, p.chh, ts, c, row), v);
                    }, h_go);
                }
                es.release();
                // ------------------------------------------------------------ GRU, part 1: r and h_n  (torch GRUCell, rnns.py:102)
                es.wait_acc();
                epi_for_chunks2(tmem_base, 0, DpT, d_beg, d_end, [&](int c, float (&ar)[8], float (&ahn)[8]) {
                    float br[8], bh[8], rr[8], hn[8];
                    lds8(sbias + bl.r + c * 8, br);
                    lds8(sbias + bl.hn + c * 8, bh);
#pragma unroll
                    for (int i = 0; i < 8; ++i) {
                        rr[i] = sigmoid_fast(ar[i] + br[i]);  // pad columns: r = 0.5, h_n = 0
                        hn[i] = ahn[i] + bh[i];
                        ar[i] = rr[i] * hn[i];
                    }
                    tmem_st8(epi_taddr(tmem_base, (c - d_go) * 8), ar);  // i_n is accumulated on top of r * h_n by the next phase
                    if (save) {
                        st_chunk_bf16(tcm_at(p.sv_r, p.cdp, ts, c, row), rr);
                        st_chunk_bf16(tcm_at(p.sv_hn, p.cdp, ts, c, row), hn);
                    }
                }, d_go);
                tmem_st_wait();
                es.release();
                // ------------------------------------------------------------ GRU, part 2: n, z, h'
                // h (fed into this step) = hi + lo: hi is the bf16 operand in ST, lo its bf16 remainder in the side
                // image (written by this very thread one step ago) -- fetched before the wait, coalesced.
                uint4 lo_in[kMaxPerD];
                {
                    const __nv_bfloat16* lo_cur = p.sv_lo + (size_t)(p.lo_per_cta ? (int)blockIdx.x : ts) * lo_stride;
#pragma unroll
                    for (int i = 0; i < kMaxPerD; ++i) {
                        const int c = d_beg + i + d_go;
                        lo_in[i] = make_uint4(0u, 0u, 0u, 0u);
                        if (d_beg + i < d_end) lo_in[i] = *reinterpret_cast<const uint4*>(lo_cur + ((size_t)c * kTileRows + row) * 8);
                    }
                }
                es.wait_acc();
                {
                    const bool feat_next = save && t + 1 < p.T;
                    const bool lo_next_on = p.lo_per_cta || t + 1 < p.T;
                    __nv_bfloat16* lo_nxt = p.sv_lo + (size_t)(p.lo_per_cta ? (int)blockIdx.x : ts + p.ntiles) * lo_stride;
#pragma unroll
                    for (int i = 0; i < kMaxPerD; ++i) {
                        const int ct = d_beg + i, c = ct + d_go;  // TMEM chunk, global chunk
                        if (ct < d_end) {
                            float an_[8], az[8], bn[8], bz[8], zz[8], nn[8], hv[8], hp[8], hl[8];
                            tmem_ld8(epi_taddr(tmem_base, ct * 8), an_);
                            tmem_ld8(epi_taddr(tmem_base, DpT + ct * 8), az);
                            uint8_t* st_c = st + ((size_t)(p.cs + c) * kTileRows + row) * 16;
                            unpack_bf16x8(*reinterpret_cast<const uint4*>(st_c), hp);
                            unpack_bf16x8(lo_in[i], hl);
                            lds8(sbias + bl.n + c * 8, bn);
                            lds8(sbias + bl.z + c * 8, bz);
                            tmem_ld_wait();
#pragma unroll
                            for (int k = 0; k < 8; ++k) {
                                const float h = hp[k] + hl[k];
                                nn[k] = tanh_fast(an_[k] + bn[k]);     // pad columns: n = 0, z = 0.5, h' = 0.5 * 0
                                zz[k] = sigmoid_fast(az[k] + bz[k]);
                                hv[k] = live ? fmaf(zz[k], h - nn[k], nn[k]) : 0.f;  // (1-z) n + z h
                            }
                            uint4 hi, lo;
                            split_hi_lo(hv, hi, lo);
                            *reinterpret_cast<uint4*>(st_c) = hi;
                            *reinterpret_cast<uint4*>(hb + ((size_t)c * kTileRows + row) * 16) = lo;
                            if (lo_next_on) *reinterpret_cast<uint4*>(lo_nxt + ((size_t)c * kTileRows + row) * 8) = lo;
                            if (save) {
                                st_chunk_bf16(tcm_at(p.sv_z, p.cdp, ts, c, row), zz);
                                st_chunk_bf16(tcm_at(p.sv_n, p.cdp, ts, c, row), nn);
                            }
                            if (feat_next) *reinterpret_cast<uint4*>(tcm_at(p.sv_feat, cfeat, ts + p.ntiles, p.cs + c, row)) = hi;
                        }
                    }
                }
                es.release();
                // fp32 deter rows of this tile-step, rebuilt from hi (ST) + lo (hidden buffer, dead until the next
                // epilogue) after the release: 8 rows x 64 contiguous bytes per warp instruction
                named_bar<kEpiThreads, 1>();
                {
                    const int q4 = D >> 2, qblocks = (q4 + 3) >> 2, nblocks = (kTileRows / 8) * qblocks;
                    const uint8_t* st_d = st + (size_t)p.cs * kChunkBytes;
                    float* out_t = p.deter + ((size_t)t * p.n + row0) * D;
                    for (int b = tid >> 5; b < nblocks; b += kEpiWarps) {
                        const int rb = b / qblocks, qb = b - rb * qblocks;
                        const int r = rb * 8 + (lane >> 2), q = qb * 4 + (lane & 3);
                        if (q < q4 && row0 + r < p.n) {
                            const uint32_t off = (uint32_t)(((q >> 1) * kTileRows + r) * 16 + (q & 1) * 8);
                            const uint2 h2 = *reinterpret_cast<const uint2*>(st_d + off);
                            const uint2 l2 = *reinterpret_cast<const uint2*>(hb + off);
                            float4 o;
                            o.x = __uint_as_float(h2.x << 16) + __uint_as_float(l2.x << 16);
                            o.y = __uint_as_float(h2.x & 0xFFFF0000u) + __uint_as_float(l2.x & 0xFFFF0000u);
                            o.z = __uint_as_float(h2.y << 16) + __uint_as_float(l2.y << 16);
                            o.w = __uint_as_float(h2.y & 0xFFFF0000u) + __uint_as_float(l2.y & 0xFFFF0000u);
                            *reinterpret_cast<float4*>(out_t + (size_t)r * D + q * 4) = o;
                        }
                    }
                }
                named_bar<kEpiThreads, 1>();
                // ------------------------------------------------------------ prior hidden  (rnns.py:103)
                // the tile's eps block for the prior sample is fetched here (coalesced, 8 registers), two phases early
                const int pcount = min(kTileRows, p.n - row0) * S;
                float epsr[8];
                {
                    const float* eps_t = p.eps_prior + ((size_t)t * p.n + row0) * S;
#pragma unroll
                    for (int k = 0; k < 8; ++k) {
                        const int i = tid + k * kEpiThreads;
                        epsr[k] = i < pcount ? __ldg(eps_t + i) : 0.f;
                    }
                }
                es.wait_acc();
                {
                    const float* bias = sbias + bl.p1;
                    __nv_bfloat16* sv = save ? p.sv_hid : nullptr;
                    epi_for_chunks(tmem_base, 0, h_beg, h_end, [&](int c, float (&v)[8]) {
                        float b[8];
                        lds8(bias + c * 8, b);
                        elu_bias8(v, b);
                        st_chunk_bf16(hb + ((size_t)c * kTileRows + row) * 16, v);
                        if (sv) st_chunk_bf16(tcm_at(sv, p.chh, ts, c, row), v);
                    }, h_go);
                }
                es.release();
                // ------------------------------------------------------------ prior mean | std, sample  (rnns.py:103-106)
                // Outputs are staged in the (now dead) hidden buffer as [128][S] blocks and leave with coalesced
                // stores after the release; eps arrives the same way.
                es.wait_acc();
                {
                    float* sE = hbf;                      // eps in, stoch out
                    float* sM = hbf + kTileRows * S;
                    float* sD = hbf + 2 * kTileRows * S;
#pragma unroll
                    for (int k = 0; k < 8; ++k) {
                        const int i = tid + k * kEpiThreads;
                        if (i < kTileRows * S) sE[i] = epsr[k];
                    }
                    named_bar<kEpiThreads, 1>();
                    const bool feat_next = save && t + 1 < p.T;
                    auto prior_chunk = [&](int c, float (&am)[8], float (&as_)[8]) {
                        float bm[8], bs[8], sv[8];
                        lds8(sbias + bl.p2m + c * 8, bm);
                        lds8(sbias + bl.p2s + c * 8, bs);
#pragma unroll
                        for (int i = 0; i < 8; ++i) {
                            const int j = c * 8 + i;
                            float st_ = 0.f;
                            if (j < S && lead) {
                                const float mean = am[i] + bm[i];
                                const float std = softplus_fast(as_[i] + bs[i]) + p.min_std;
                                const int o = row * S + j;
                                st_ = live ? fmaf(std, sE[o], mean) : 0.f;
                                sM[o] = mean; sD[o] = std; sE[o] = st_;
                            }
                            sv[i] = st_;
                        }
                        if (lead) {
                            st_chunk_bf16(st + ((size_t)c * kTileRows + row) * 16, sv);
                            if (feat_next) st_chunk_bf16(tcm_at(p.sv_feat, cfeat, ts + p.ntiles, c, row), sv);
                        }
                    };
                    if (kHalf) {
                        // mean chunk c sits in lanes 0..15, std chunk c in lanes 16..31 of the same TMEM columns
                        for (int c = s_beg; c < s_end; ++c) {
                            float mine[8], am[8], as_[8];
                            tmem_ld8(epi_taddr(tmem_base, c * 8), mine);
                            tmem_ld_wait();
#pragma unroll
                            for (int i = 0; i < 8; ++i) {
                                const float o = __shfl_xor_sync(0xffffffffu, mine[i], 16);
                                am[i] = lead ? mine[i] : o;
                                as_[i] = lead ? o : mine[i];
                            }
                            prior_chunk(c, am, as_);
                        }
                    } else {
                        epi_for_chunks2(tmem_base, 0, p.Sp, s_beg, s_end, prior_chunk);
                    }
                }
                if (t + 1 < p.T) es.release();
                named_bar<kEpiThreads, 1>();
                {
                    const size_t o0 = ((size_t)t * p.n + row0) * S;
                    const float* sE = hbf;
                    const float* sM = hbf + kTileRows * S;
                    const float* sD = hbf + 2 * kTileRows * S;
                    for (int i = tid; i < pcount; i += kEpiThreads) {
                        p.prior_mean[o0 + i] = sM[i];
                        p.prior_std[o0 + i] = sD[i];
                        p.prior_stoch[o0 + i] = sE[i];
                    }
                }
                named_bar<kEpiThreads, 1>();  // the hidden buffer is written again by the next epilogue
                // release: the barrier orders every epilogue thread's stores of step t (deter, action, prior) before this
                // thread's gpu-scope fence, which is cumulative (the CUTLASS barrier pattern: sync, then one thread releases)
                if (p.step_flags && tid == 0) { __threadfence(); atomicAdd(p.step_flags + t, 1); }
            }
        }
    }
    chain_teardown(tmem_base);
}

// ------------------------------------------------------------------------------------------------
// backward kernel (reverse time).  Gradients flow state -> state through the transition and into the
// actions; the actor's input is detached (rnns.py:236-238), the world model is frozen
// (dreamer_algo.py:293), so only data gradients of the transition and data + weight gradients of the
// actor are produced.  The carried gradients live in TMEM between steps:
//   R_H  cols [0, colC)            scratch accumulator of the hidden-sized GEMMs
//   R_C  cols [colC, colC + Dp)    d deter: prior-path and GRU contributions are ACCUMULATED here by the MMAs
//   R_AS cols [colAS, ...)         d [stoch | action] of the rnn input Linear; the stoch part is next step's carry
// Phases per step:
//   1  d prior hidden   = d[mean|std] . W_p2          A = G2              -> R_H
//   2  d deter         += dhid . W_p1                 A = GA (as HB)      -> R_C (accumulate)
//   3  dx  = dpr.W_ir + dpz.W_iz ; d deter_prev += dpr.W_hr + dpz.W_hz    A = GA, GB
//   4  dx += din.W_in ; d deter_prev += dhn.W_hn                          A = GA, GB (reloaded from the stash)
//   5  d[stoch|action]  = dx_pre . W_in               A = GA (as HB)      -> R_AS
//   6  d actor hidden   = d raw . Wa[last]            A = G2              -> R_H
//   7.. d actor hidden  = dh . Wa[l]                  A = GA (as HB)      -> R_H
// ------------------------------------------------------------------------------------------------
template <int TR>
__global__ void __launch_bounds__(kChainThreads, 1) rollout_bwd_tc_kernel(const __grid_constant__ RollTcDev p) {
    constexpr int kTileRows = TR;  // tile geometry of this instantiation (see rollout_fwd_tc_kernel)
    constexpr int kChunkBytes = TR * 16;
    constexpr bool kHalf = TR == 64;
    auto tcm_at = [](auto* base, int chunks_per_tile, int tile, int chunk, int row) {
        return base + (((size_t)tile * chunks_per_tile + chunk) * TR + row) * 8;
    };
    extern __shared__ __align__(1024) uint8_t smem[];
    ChainBars* bars = reinterpret_cast<ChainBars*>(smem);
    const uint32_t tmem_base = chain_setup(bars, p.pg);
    const int warp = __shfl_sync(0xffffffffu, threadIdx.x >> 5, 0);  // warp-uniform for the compiler
    const int my_tiles = (int)blockIdx.x < p.ntiles ? (p.ntiles - blockIdx.x + gridDim.x - 1) / gridDim.x : 0;
    const int reps = my_tiles * p.T;
    if (warp == kEpiWarps) {
        chain_producer(p.pg, bars, smem, p.wstream, reps);
    } else if (warp == kEpiWarps + 1) {
        chain_mma<TR>(p.pg, bars, smem, tmem_base, reps);
    } else {
        EpiSync es(bars);
        uint8_t* ga = smem + p.ga_off;
        uint8_t* gb = smem + p.gb_off;
        uint8_t* g2 = smem + p.g2_off;
        const EpiGeo<TR> geo;
        const int row = geo.row, grp = epi_group(), lane = threadIdx.x & 31, tid = threadIdx.x;
        const bool lead = geo.half == 0;  // half tiles: lanes 16..31 see the second column half of every accumulator
        const int S = p.S, D = p.D, A = p.A;
        const int Dp = p.cdp * 8;
        const int DpT = kHalf ? Dp / 2 : Dp;
        const int cfeat = p.cs + p.cdp;
        const int colC = max(p.chh, p.cha) * (kHalf ? 4 : 8), colAS = colC + DpT;
        // R_AS: 128-row tiles [stoch Sp | action 16]; half tiles: stoch in lanes 0..15, action in lanes 16..31 of columns 0..Sp
        const int colAct = kHalf ? colAS : colAS + p.Sp;
        auto rng = [&](int nchunks, int& beg, int& end, int& go) {
            const int nct = geo.tmem_chunks(nchunks), per = (nct + kEpiGroups - 1) / kEpiGroups;
            beg = min(nct, grp * per); end = min(nct, beg + per); go = geo.goff(nchunks);
        };
        __nv_bfloat16* stash_in = p.stash + (size_t)blockIdx.x * 2 * p.cdp * kTileRows * 8;
        __nv_bfloat16* stash_hn = stash_in + (size_t)p.cdp * kTileRows * 8;
        const int last = p.an - 1;
        // fp32 staging blocks of the prior head, [128][S] each: GB is dead from the GRU's last GEMM until the next
        // GRU epilogue, GA from the step's last GEMM until the d-prior-hidden epilogue
        float* sEps = reinterpret_cast<float*>(gb);
        float* sStd = sEps + kTileRows * S;
        float* sGst = sStd + kTileRows * S;
        float* sGm = reinterpret_cast<float*>(ga);
        float* sGs = sGm + kTileRows * S;
        const size_t chunk_img = (size_t)kTileRows * 8;  // elements of one chunk of a TCM image
        for (int tile = blockIdx.x; tile < p.ntiles; tile += gridDim.x) {
            const int row0 = tile * kTileRows;
            const int grow = row0 + row;
            const bool live = grow < p.n;
            const int rows_live = min(kTileRows, p.n - row0);
            // asynchronous, coalesced fetch of step t2's prior-head blocks into the staging area
            auto fetch_prior = [&](int t2) {
                if (p.g_wait_flags) {  // the head kernels producing g_* run beside this one (drm_rollout_grad.g_wait_flags)
                    if (lane == 0) {
                        wait_counter_ge(p.g_wait_flags + t2, p.g_wait_target, 100);
                    }
                    __syncwarp();
                }
                const size_t o0 = ((size_t)t2 * p.n + row0) * S;
                for (int i = tid; i < rows_live * S; i += kEpiThreads) {
                    cp_async4(sEps + i, p.eps_prior + o0 + i);
                    cp_async4(sStd + i, p.prior_std + o0 + i);
                    if (p.g_prior_stoch) cp_async4(sGst + i, p.g_prior_stoch + o0 + i);
                    if (p.g_prior_mean) cp_async4(sGm + i, p.g_prior_mean + o0 + i);
                    if (p.g_prior_std) cp_async4(sGs + i, p.g_prior_std + o0 + i);
                }
            };
            // L2 prefetch (one thread) of what the epilogues of step t2 read, issued about one phase ahead
            auto pf_hid = [&](int t2) { if (tid == 0) l2_prefetch(tcm_at(p.sv_hid, p.chh, t2 * p.ntiles + tile, 0, 0), p.chh * chunk_img * 2); };
            auto pf_gru = [&](int t2) {
                if (tid == 0) {
                    const int ts2 = t2 * p.ntiles + tile;
                    l2_prefetch(tcm_at(p.sv_r, p.cdp, ts2, 0, 0), p.cdp * chunk_img * 2);
                    l2_prefetch(tcm_at(p.sv_z, p.cdp, ts2, 0, 0), p.cdp * chunk_img * 2);
                    l2_prefetch(tcm_at(p.sv_n, p.cdp, ts2, 0, 0), p.cdp * chunk_img * 2);
                    l2_prefetch(tcm_at(p.sv_hn, p.cdp, ts2, 0, 0), p.cdp * chunk_img * 2);
                    l2_prefetch(tcm_at(p.sv_lo, p.cdp, ts2, 0, 0), p.cdp * chunk_img * 2);
                    l2_prefetch(tcm_at(p.sv_feat, cfeat, ts2, p.cs, 0), p.cdp * chunk_img * 2);
                    if (p.g_deter) l2_prefetch(p.g_deter + ((size_t)t2 * p.n + row0) * D, (size_t)rows_live * D * 4);
                }
            };
            // ---- zero the carried gradients
            {
                float z[8];
#pragma unroll
                for (int i = 0; i < 8; ++i) z[i] = 0.f;
                for (int c = grp; c < geo.tmem_chunks(p.cdp); c += kEpiGroups) tmem_st8(epi_taddr(tmem_base, colC + c * 8), z);
                for (int c = grp; c < p.cs; c += kEpiGroups) tmem_st8(epi_taddr(tmem_base, colAS + c * 8), z);
                tmem_st_wait();
            }
            fetch_prior(p.T - 1);
            pf_hid(p.T - 1);
            for (int t = p.T - 1; t >= 0; --t) {
                const int ts = t * p.ntiles + tile;
                const size_t trow = (size_t)t * p.n + grow;
                // ------------------------------------------------------------ prior head (pointwise) -> G2
                cp_async_wait_all();
                named_bar<kEpiThreads, 1>();
                // the dY images of step t + 1 are complete: one thread releases them (barrier + cumulative gpu-scope fence)
                if (p.bwd_flags && tid == 0 && t + 1 < p.T) { __threadfence(); atomicAdd(p.bwd_flags + t + 1, 1); }
                pf_gru(t);
                {
                    const int per = (p.cs + kEpiGroups - 1) / kEpiGroups;
                    const int cbeg = grp * per, cend = min(p.cs, cbeg + per);
                    for (int c = cbeg; c < cend; ++c) {
                        float cs_[8], dm[8], ds[8];
                        tmem_ld8(epi_taddr(tmem_base, colAS + c * 8), cs_);
                        tmem_ld_wait();
#pragma unroll
                        for (int i = 0; i < 8; ++i) {
                            const int j = c * 8 + i;
                            dm[i] = 0.f; ds[i] = 0.f;
                            if (live && lead && j < S) {
                                const int o = row * S + j;
                                const float dst = (p.g_prior_stoch ? sGst[o] : 0.f) + cs_[i];
                                dm[i] = (p.g_prior_mean ? sGm[o] : 0.f) + dst;
                                const float dsd = (p.g_prior_std ? sGs[o] : 0.f) + dst * sEps[o];
                                ds[i] = dsd * (1.f - __expf(-(sStd[o] - p.min_std)));  // sigmoid(raw) from softplus(raw)
                            }
                        }
                        if (lead) {
                            st_chunk_bf16(g2 + ((size_t)c * kTileRows + row) * 16, dm);
                            st_chunk_bf16(g2 + ((size_t)(p.cs + c) * kTileRows + row) * 16, ds);
                        }
                    }
                }
                es.release();
                // ------------------------------------------------------------ d prior hidden
                {
                    int cbeg, cend, go;
                    rng(p.chh, cbeg, cend, go);
                    epi_for_chunks_saved(es, tmem_base, 0, cbeg, cend,
                        [&](int c) { return *reinterpret_cast<const uint4*>(tcm_at(p.sv_hid, p.chh, ts, c, row)); },
                        [&](int c, float (&v)[8], const uint4 hq) {
                            elu_bwd8(v, hq, live);
                            st_chunk_bf16(ga + ((size_t)c * kTileRows + row) * 16, v);
                        }, [] {}, go);
                }
                es.release();
                // ------------------------------------------------------------ GRU cell (pointwise)
                es.wait_acc();
                if (tid == 0) l2_prefetch(tcm_at(p.sv_x, p.chh, ts, 0, 0), p.chh * chunk_img * 2);
                if (p.g_deter) {
                    // upstream d deter rows of this tile-step, coalesced (8 rows x 64 bytes per warp instruction), as
                    // fp32 chunk-major halves into GA | GB: exactly the slots this thread overwrites below
                    const int q4 = D >> 2, qblocks = (q4 + 3) >> 2, nblocks = (kTileRows / 8) * qblocks;
                    const float* src_t = p.g_deter + ((size_t)t * p.n + row0) * D;
                    for (int b = tid >> 5; b < nblocks; b += kEpiWarps) {
                        const int rb = b / qblocks, qb = b - rb * qblocks;
                        const int r = rb * 8 + (lane >> 2), q = qb * 4 + (lane & 3);
                        if (q < q4 && r < rows_live)
                            cp_async16(((q & 1) ? gb : ga) + ((size_t)(q >> 1) * kTileRows + r) * 16, src_t + (size_t)r * D + q * 4);
                    }
                    cp_async_wait_all();
                    named_bar<kEpiThreads, 1>();
                }
                {
                    int cbeg, cend, go;
                    rng(p.cdp, cbeg, cend, go);
                    epi_for_chunks(tmem_base, colC, cbeg, cend, [&](int c, float (&dh)[8]) {
                        float rr[8], zz[8], nn[8], hn[8], dpr[8], dpz[8], din[8], dhn[8], gd[8], hpv[8], hlo[8];
                        ld_chunk_bf16(tcm_at(p.sv_r, p.cdp, ts, c, row), rr);
                        ld_chunk_bf16(tcm_at(p.sv_z, p.cdp, ts, c, row), zz);
                        ld_chunk_bf16(tcm_at(p.sv_n, p.cdp, ts, c, row), nn);
                        ld_chunk_bf16(tcm_at(p.sv_hn, p.cdp, ts, c, row), hn);
                        // deter fed into step t = hi (feat image) + lo (side image), both tile-major and coalesced
                        ld_chunk_bf16(tcm_at(p.sv_feat, cfeat, ts, p.cs + c, row), hpv);
                        ld_chunk_bf16(tcm_at(p.sv_lo, p.cdp, ts, c, row), hlo);
#pragma unroll
                        for (int i = 0; i < 8; ++i) { gd[i] = 0.f; hpv[i] += hlo[i]; }
                        if (p.g_deter) {
                            const float4 lo4 = *reinterpret_cast<const float4*>(ga + ((size_t)c * kTileRows + row) * 16);
                            const float4 hi4 = *reinterpret_cast<const float4*>(gb + ((size_t)c * kTileRows + row) * 16);
                            gd[0] = lo4.x; gd[1] = lo4.y; gd[2] = lo4.z; gd[3] = lo4.w;
                            gd[4] = hi4.x; gd[5] = hi4.y; gd[6] = hi4.z; gd[7] = hi4.w;
                        }
#pragma unroll
                        for (int i = 0; i < 8; ++i) {
                            const int col = c * 8 + i;
                            const bool ok = live && col < D;
                            float d = ok ? dh[i] + gd[i] : 0.f;
                            const float hp = hpv[i];
                            const float dn = d * (1.f - zz[i]);
                            const float dz = d * (hp - nn[i]);
                            const float dpn = dn * (1.f - nn[i] * nn[i]);
                            din[i] = dpn;
                            dhn[i] = dpn * rr[i];
                            dpr[i] = dpn * hn[i] * rr[i] * (1.f - rr[i]);
                            dpz[i] = dz * zz[i] * (1.f - zz[i]);
                            dh[i] = d * zz[i];  // carry through h' = (1-z) n + z h
                        }
                        tmem_st8(epi_taddr(tmem_base, colC + (c - go) * 8), dh);
                        st_chunk_bf16(ga + ((size_t)c * kTileRows + row) * 16, dpr);
                        st_chunk_bf16(gb + ((size_t)c * kTileRows + row) * 16, dpz);
                        st_chunk_bf16(stash_in + ((size_t)c * kTileRows + row) * 8, din);
                        st_chunk_bf16(stash_hn + ((size_t)c * kTileRows + row) * 8, dhn);
                    }, go);
                    tmem_st_wait();
                }
                es.release();
                // ------------------------------------------------------------ reload d i_n, d h_n (own writes) -> GA, GB
                {
                    int tb, te, go;
                    rng(p.cdp, tb, te, go);
                    const int cbeg = tb + go, cend = te + go;  // global chunks
                    es.wait_acc();
#pragma unroll 2
                    for (int c = cbeg; c < cend; ++c) {
                        const uint4 a = *reinterpret_cast<const uint4*>(stash_in + ((size_t)c * kTileRows + row) * 8);
                        const uint4 b = *reinterpret_cast<const uint4*>(stash_hn + ((size_t)c * kTileRows + row) * 8);
                        *reinterpret_cast<uint4*>(ga + ((size_t)c * kTileRows + row) * 16) = a;
                        *reinterpret_cast<uint4*>(gb + ((size_t)c * kTileRows + row) * 16) = b;
                    }
                }
                es.release();
                // ------------------------------------------------------------ d rnn-input pre-activation
                {
                    int cbeg, cend, go;
                    rng(p.chh, cbeg, cend, go);
                    epi_for_chunks_saved(es, tmem_base, 0, cbeg, cend,
                        [&](int c) { return *reinterpret_cast<const uint4*>(tcm_at(p.sv_x, p.chh, ts, c, row)); },
                        [&](int c, float (&v)[8], const uint4 hq) {
                            elu_bwd8(v, hq, live);
                            st_chunk_bf16(ga + ((size_t)c * kTileRows + row) * 16, v);
                        }, [] {}, go);
                    if (p.want_actor && tid == 0) l2_prefetch(tcm_at(p.sv_ah[last - 1], p.cha, ts, 0, 0), p.cha * chunk_img * 2);
                }
                es.release();
                // ------------------------------------------------------------ d [stoch | action] is in R_AS now
                es.wait_acc();
                if (!p.want_actor && t > 0) { fetch_prior(t - 1); pf_hid(t - 1); }
                if (p.want_actor) {
                    // actor head (pointwise): d action -> d raw -> G2 (K = 16) [+ TCM copy, fp32 bias gradient]
                    if (grp == 0) {
                        float da[16], draw[16];
                        {
                            float v0[8], v1[8];
                            tmem_ld8(epi_taddr(tmem_base, colAct), v0);
                            tmem_ld8(epi_taddr(tmem_base, colAct + 8), v1);
                            tmem_ld_wait();
                            if (kHalf) {  // the action columns of this row sit in the partner lane (lanes 16..31)
#pragma unroll
                                for (int i = 0; i < 8; ++i) {
                                    v0[i] = __shfl_xor_sync(0xffffffffu, v0[i], 16);
                                    v1[i] = __shfl_xor_sync(0xffffffffu, v1[i], 16);
                                }
                            }
#pragma unroll
                            for (int i = 0; i < 8; ++i) { da[i] = v0[i]; da[8 + i] = v1[i]; }
                        }
#pragma unroll
                        for (int j = 0; j < 16; ++j) draw[j] = 0.f;
                        if (live && lead) {
                            const float* go = p.g_action_out ? p.g_action_out + trow * A : nullptr;
                            const float* raw = p.sv_raw + trow * p.Aout;
                            const float* av = p.action_out + trow * A;
                            if (p.act_dist == DRM_DIST_TANH_NORMAL) {
#pragma unroll
                                for (int j = 0; j < 8; ++j) {
                                    if (j < A) {
                                        const float d = da[j] + (go ? go[j] : 0.f);
                                        const float a = av[j];
                                        const float dpre = d * (1.f - a * a);
                                        const float th = tanh_fast(raw[j] / p.mean_scale);
                                        draw[j] = dpre * (1.f - th * th);
                                        draw[A + j] = dpre * __ldg(p.noise_act + trow * A + j) * sigmoid_fast(raw[A + j] + p.raw_init_std);
                                    }
                                }
                            } else if (p.act_dist == DRM_DIST_ONE_HOT) {
                                float mx = -INFINITY, den = 0.f, dot = 0.f;
#pragma unroll
                                for (int j = 0; j < 16; ++j) if (j < A) mx = fmaxf(mx, raw[j]);
#pragma unroll
                                for (int j = 0; j < 16; ++j) if (j < A) den += __expf(raw[j] - mx);
#pragma unroll
                                for (int j = 0; j < 16; ++j) if (j < A) dot += (da[j] + (go ? go[j] : 0.f)) * (__expf(raw[j] - mx) / den);
#pragma unroll
                                for (int j = 0; j < 16; ++j)
                                    if (j < A) draw[j] = (__expf(raw[j] - mx) / den) * ((da[j] + (go ? go[j] : 0.f)) - dot);
                            } else {
                                float dot = 0.f;
#pragma unroll
                                for (int j = 0; j < 16; ++j) if (j < A) dot += (da[j] + (go ? go[j] : 0.f)) * av[j];
#pragma unroll
                                for (int j = 0; j < 16; ++j)
                                    if (j < A) draw[j] = av[j] * ((da[j] + (go ? go[j] : 0.f)) - dot) / p.temperature;
                            }
                        }
                        for (int c = 0; c < 2; ++c) {
                            float v[8];
#pragma unroll
                            for (int i = 0; i < 8; ++i) v[i] = c == 0 ? draw[i] : draw[8 + i];
                            if (lead) {
                                st_chunk_bf16(g2 + ((size_t)c * kTileRows + row) * 16, v);
                                st_chunk_bf16(tcm_at(p.dys_a[last], 2, ts, c, row), v);
                            }
                        }
#pragma unroll
                        for (int j = 0; j < 16; ++j) {
                            const float sum = warp_sum(draw[j]);
                            if (lane == 0 && j < p.Aout) atomicAdd(p.db_a_last + j, sum);
                        }
                    }
                    es.release();
                    // actor hidden layers, last Linear first; layer 0's input (feat) is detached (rnns.py:238)
                    for (int l = last; l >= 1; --l) {
                        int cbeg, cend, go;
                        rng(p.cha, cbeg, cend, go);
                        epi_for_chunks_saved(es, tmem_base, 0, cbeg, cend,
                            [&](int c) { return *reinterpret_cast<const uint4*>(tcm_at(p.sv_ah[l - 1], p.cha, ts, c, row)); },
                            [&](int c, float (&v)[8], const uint4 hq) {
                                elu_bwd8(v, hq, live);
                                if (l > 1) st_chunk_bf16(ga + ((size_t)c * kTileRows + row) * 16, v);  // layer 0 gets no data gradient
                                st_chunk_bf16(tcm_at(p.dys_a[l - 1], p.cha, ts, c, row), v);
                            },
                            [&] {
                                if (l > 1) {
                                    if (tid == 0) l2_prefetch(tcm_at(p.sv_ah[l - 2], p.cha, ts, 0, 0), p.cha * chunk_img * 2);
                                } else if (t > 0) {
                                    // last epilogue of the step: GA and GB are dead, start on the next step's prior head
                                    fetch_prior(t - 1);
                                    pf_hid(t - 1);
                                }
                            }, go);
                        if (l > 1) es.release();
                    }
                }
            }
            // the next tile's zero-fill of R_C / R_AS and its first release() come after these loads
            if (p.bwd_flags) {
                named_bar<kEpiThreads, 1>();
                if (tid == 0) { __threadfence(); atomicAdd(p.bwd_flags, 1); }
            }
        }
    }
    chain_teardown(tmem_base);
}

// ------------------------------------------------------------------------------------------------
// host side: program, weight images, shared memory and save layout
// ------------------------------------------------------------------------------------------------
struct RollTcPlan {
    RollTcDev dev;
    PackTcJobs jobs;
    size_t wbytes;
    size_t smem_bytes;
    size_t save_bytes;
};

int rollout_tc_supported(const drm_rollout* r) {
    const drm_rssm& w = r->rssm;
    if (r->mode != DRM_MODE_POLICY) return 0;
    // Categorical policies sample through an argmax / a temperature-0.1 softmax of the logits: both amplify the
    // bf16 operand rounding beyond the 1e-2 parity budget (a flipped argmax is an O(1) difference), so those
    // rollouts stay on the fp32 kernels.
    if (r->act_dist != DRM_DIST_TANH_NORMAL) return 0;
    const int Aout = r->actor.out_dim;
    if (Aout > 16 || w.A > 16) return 0;
    if (rup(w.D, 16) > 256 || rup(w.Hd, 16) > 256 || rup(r->actor.hidden, 16) > 256 || w.S > 64) return 0;
    if (r->actor.n_linear < 2) return 0;
    // staged global I/O: fp32 [128][S] x 3 and [128][16] x 3 blocks in the hidden buffer, 8 eps registers per thread,
    // float4 deter rows
    {
        const int hbc = (rup((w.Hd + 7) / 8, 2) > rup((r->actor.hidden + 7) / 8, 2)) ? rup((w.Hd + 7) / 8, 2) : rup((r->actor.hidden + 7) / 8, 2);
        if (w.S > 32 || (w.D & 3) || 3 * kTileRows * w.S * 4 > hbc * kChunkBytes || 3 * kTileRows * 16 * 4 > hbc * kChunkBytes) return 0;
        if (rup((w.D + 7) / 8, 2) > hbc) return 0;  // lo staging of the deter tile
        if (rup((w.D + 7) / 8, 2) > 4 * kMaxPerD) return 0;
    }
    // tensor memory of the backward: R_H | R_C | R_AS
    const int chh = rup((w.Hd + 7) / 8, 2), cha = rup((r->actor.hidden + 7) / 8, 2), cdp = rup((w.D + 7) / 8, 2);
    const int cs = rup((w.S + 7) / 8, 2);
    if ((chh > cha ? chh : cha) * 8 + cdp * 8 + rup(cs * 8 + 16, 16) > kTmemCols) return 0;
    return 1;
}

// Rows per CTA.  64-row half tiles (see chain_mma) put more CTAs on the SMs than 128-row tiles while the batch is small.
// From the point where the 64-row tiles would (nearly) fill the machine in one wave, 128-row tiles win: they leave SMs
// for the head kernels that run BESIDE the rollout (algo.py) and never need a second, partly filled wave.  Measured, ms
// per step 64 / 128 rows (profiles/r2_experiments.md): 4 096 start states 1.11 / 1.45; 8 192 (one GPU's shard of BASELINE
// configs[3] at 8 GPUs) 1.89 / 1.71; 10 240 3.01 / 2.35; 12 288 3.15 / 2.67; 14 336 3.37 / 2.91; 16 384 3.53 / 3.12;
// around the switch: 5 000 1.28 / 1.47; 6 000 1.65 / 1.50; 7 000 1.77 / 1.57.
static int tile_rows_for(int n) {
    static int forced = -1;  // developer / test override DRM_ROLLOUT_TR = 64 | 128, read once
    if (forced < 0) { const char* e = getenv("DRM_ROLLOUT_TR"); forced = e ? atoi(e) : 0; }
    if (forced == 64 || forced == 128) return forced;
    const int sms = num_sms();
    const int t64 = (n + 63) / 64;
    return t64 > sms - 64 ? 128 : 64;  // 64-row tiles as long as they leave 2 x 32 SMs to the head kernels (algo._MIN_FREE)
}
static size_t tcm_bytes_r(int tr, size_t tiles, int chunks) { return tiles * (size_t)chunks * tr * 16; }

static void set_dims(const drm_rollout* r, RollTcDev& d) {
    const drm_rssm& w = r->rssm;
    memset(&d, 0, sizeof(d));
    d.T = r->steps; d.n = r->n; d.tr = tile_rows_for(r->n); d.ntiles = (r->n + d.tr - 1) / d.tr;
    d.act_dist = r->act_dist; d.training = r->training;
    d.A = w.A; d.S = w.S; d.D = w.D; d.Hd = w.Hd; d.ah = r->actor.hidden; d.an = r->actor.n_linear; d.Aout = r->actor.out_dim;
    d.cs = rup((w.S + 7) / 8, 2); d.cd = (w.D + 7) / 8; d.cdp = rup(d.cd, 2); d.ca = rup((w.A + 7) / 8, 2);
    d.chh = rup((w.Hd + 7) / 8, 2); d.cha = rup((d.ah + 7) / 8, 2);
    d.Sp = d.cs * 8;
    d.mean_scale = r->act_mean_scale; d.raw_init_std = r->act_raw_init_std; d.act_min_std = r->act_min_std;
    d.temperature = r->act_temperature; d.min_std = r->min_std;
}

struct ProgBuilder {
    TcProgram* pg;
    PackTcJobs* jobs;
    uint8_t* wbase;
    size_t woff;
    bool ok;
    int tr;  // 64: every op runs as two column halves into the same TMEM columns (chain_mma), so columns halve
    TcOp* add_op(uint32_t a0, int c0, uint32_t a1, int c1, int n, int d_col, int accumulate, int phase_end) {
        if (pg->n_ops >= kMaxOps) { ok = false; return nullptr; }
        TcOp& op = pg->op[pg->n_ops++];
        op.a_off[0] = a0; op.a_off[1] = a1; op.a_chunks[0] = (uint16_t)c0; op.a_chunks[1] = (uint16_t)c1;
        op.n = (uint16_t)n; op.d_col = (uint16_t)(tr == 64 ? d_col / 2 : d_col); op.accumulate = (uint16_t)accumulate; op.phase_end = (uint16_t)phase_end;
        if (phase_end) pg->n_phases++;
        return &op;
    }
    // image of `k_chunks` K chunks for an op with MMA N = n_pad; appended to the weight stream
    PackTcJob* add_job(const float* src, int ld, int transposed, int n0, int n_valid, int n_pad, int k_chunks) {
        if (jobs->n >= kMaxPackTcJobs) { ok = false; return nullptr; }
        PackTcJob& j = jobs->j[jobs->n++];
        memset(&j, 0, sizeof(j));
        j.src = src; j.dst = wbase ? reinterpret_cast<__nv_bfloat16*>(wbase + woff) : nullptr;
        j.ld = ld; j.transposed = transposed; j.n0 = n0; j.n_valid = n_valid; j.n_pad = n_pad; j.k_chunks = k_chunks;
        woff += (size_t)k_chunks * n_pad * 16;
        return &j;
    }
    static void seg(PackTcJob* j, int kstart, int len, int src0) {
        if (!j) return;
        j->seg_kstart[j->nseg] = kstart; j->seg_len[j->nseg] = len; j->seg_src0[j->nseg] = src0; j->nseg++;
    }
};

static int finish_smem(RollTcDev& d, size_t& smem_bytes) {
    TcProgram& pg = d.pg;
    d.bias_off = 256;
    const BiasLayout bl = bias_layout(d.an, d.cha * 8, d.chh * 8, d.cdp * 8, d.Sp);
    d.st_off = (uint32_t)rup_sz(d.bias_off + (size_t)bl.total * sizeof(float), 1024);
    d.hb_off = d.st_off + (uint32_t)(d.cs + d.cdp + d.ca) * (d.tr * 16);
    const int hb_chunks = d.chh > d.cha ? d.chh : d.cha;
    pg.stage_off = (uint32_t)rup_sz(d.hb_off + (size_t)hb_chunks * (d.tr * 16), 1024);
    pg.stage_bytes = 32 * 1024;
    int stages = (int)(((size_t)kMaxSmem - pg.stage_off) / pg.stage_bytes);
    if (stages > 4) stages = 4;
    DRM_REQUIRE(stages >= 2, DRM_E_SHAPE, "model too wide for the bf16 rollout's shared memory");
    pg.n_stages = stages;
    smem_bytes = pg.stage_off + (size_t)stages * pg.stage_bytes;
    return DRM_OK;
}

static int plan_fwd(const drm_rollout* r, uint8_t* wbase, RollTcPlan& pl) {
    const drm_rssm& w = r->rssm;
    const drm_mlp& a = r->actor;
    RollTcDev& d = pl.dev;
    set_dims(r, d);
    int rc = finish_smem(d, pl.smem_bytes); if (rc) return rc;
    const int S = w.S, D = w.D, Hd = w.Hd, A = w.A;
    const int Dp = d.cdp * 8, Hp = d.chh * 8, Ap = d.cha * 8;
    const uint32_t cb = (uint32_t)d.tr * 16;  // bytes of one operand chunk
    const uint32_t st_stoch = d.st_off, st_deter = d.st_off + (uint32_t)d.cs * cb;
    const uint32_t st_act = d.st_off + (uint32_t)(d.cs + d.cdp) * cb, hb = d.hb_off;
    pl.jobs.n = 0; d.pg.n_ops = 0; d.pg.n_phases = 0;
    ProgBuilder b{&d.pg, &pl.jobs, wbase, 0, true, d.tr};
    PackTcJob* j;
    // actor layer 0 on [stoch | deter]
    b.add_op(st_stoch, d.cs + d.cdp, 0, 0, Ap, 0, 0, 1);
    j = b.add_job(a.W[0], S + D, 0, 0, a.hidden, Ap, d.cs + d.cdp);
    ProgBuilder::seg(j, 0, S, 0); ProgBuilder::seg(j, d.cs * 8, D, S);
    for (int l = 1; l < a.n_linear; ++l) {
        const bool lastl = l == a.n_linear - 1;
        const int npad = lastl ? 16 : Ap;
        b.add_op(hb, d.cha, 0, 0, npad, 0, 0, 1);
        j = b.add_job(a.W[l], a.hidden, 0, 0, lastl ? a.out_dim : a.hidden, npad, d.cha);
        ProgBuilder::seg(j, 0, a.hidden, 0);
    }
    // x = W_in [action | stoch]: A operand order is [stoch | action]
    b.add_op(st_stoch, d.cs, st_act, d.ca, Hp, 0, 0, 1);
    j = b.add_job(w.W_in, A + S, 0, 0, Hd, Hp, d.cs + d.ca);
    ProgBuilder::seg(j, 0, S, A); ProgBuilder::seg(j, d.cs * 8, A, 0);
    // GRU part 1: r = [W_ir x + W_hr h], h_n = W_hn h
    b.add_op(hb, d.chh, st_deter, d.cdp, Dp, 0, 0, 0);
    j = b.add_job(w.W_ih, Hd, 0, 0, D, Dp, d.chh); ProgBuilder::seg(j, 0, Hd, 0);
    j = b.add_job(w.W_hh, D, 0, 0, D, Dp, d.cdp); ProgBuilder::seg(j, 0, D, 0);
    b.add_op(st_deter, d.cdp, 0, 0, Dp, Dp, 0, 1);
    j = b.add_job(w.W_hh, D, 0, 2 * D, D, Dp, d.cdp); ProgBuilder::seg(j, 0, D, 0);
    // GRU part 2: i_n accumulated onto r * h_n, z
    b.add_op(hb, d.chh, 0, 0, Dp, 0, 1, 0);
    j = b.add_job(w.W_ih, Hd, 0, 2 * D, D, Dp, d.chh); ProgBuilder::seg(j, 0, Hd, 0);
    b.add_op(hb, d.chh, st_deter, d.cdp, Dp, Dp, 0, 1);
    j = b.add_job(w.W_ih, Hd, 0, D, D, Dp, d.chh); ProgBuilder::seg(j, 0, Hd, 0);
    j = b.add_job(w.W_hh, D, 0, D, D, Dp, d.cdp); ProgBuilder::seg(j, 0, D, 0);
    // prior
    b.add_op(st_deter, d.cdp, 0, 0, Hp, 0, 0, 1);
    j = b.add_job(w.W_p1, D, 0, 0, Hd, Hp, d.cdp); ProgBuilder::seg(j, 0, D, 0);
    const int n2 = rup(d.Sp + d.cs * 8, 16);
    b.add_op(hb, d.chh, 0, 0, n2, 0, 0, 1);
    j = b.add_job(w.W_p2, Hd, 0, 0, S, n2, d.chh); ProgBuilder::seg(j, 0, Hd, 0);
    if (j) { j->n2_start = d.Sp; j->n2_valid = S; j->n2_src0 = S; }
    DRM_REQUIRE(b.ok, DRM_E_SHAPE, "actor too deep for the bf16 rollout");
    pl.wbytes = b.woff;
    return DRM_OK;
}

// carve the save buffer; returns bytes
static size_t carve_save_tc(const RollTcDev& dd, uint8_t* base, RollTcDev* out) {
    const size_t TS = (size_t)dd.T * dd.ntiles;
    size_t off = 0;
    auto take = [&](size_t bytes) -> uint8_t* {
        uint8_t* ptr = base ? base + off : nullptr;
        off += rup_sz(bytes, 256);
        return ptr;
    };
    RollTcDev tmp = dd;
    tmp.sv_feat = (__nv_bfloat16*)take(tcm_bytes_r(dd.tr, TS, dd.cs + dd.cdp));
    for (int l = 0; l < dd.an - 1; ++l) tmp.sv_ah[l] = (__nv_bfloat16*)take(tcm_bytes_r(dd.tr, TS, dd.cha));
    tmp.sv_x = (__nv_bfloat16*)take(tcm_bytes_r(dd.tr, TS, dd.chh));
    tmp.sv_r = (__nv_bfloat16*)take(tcm_bytes_r(dd.tr, TS, dd.cdp));
    tmp.sv_z = (__nv_bfloat16*)take(tcm_bytes_r(dd.tr, TS, dd.cdp));
    tmp.sv_n = (__nv_bfloat16*)take(tcm_bytes_r(dd.tr, TS, dd.cdp));
    tmp.sv_hn = (__nv_bfloat16*)take(tcm_bytes_r(dd.tr, TS, dd.cdp));
    tmp.sv_hid = (__nv_bfloat16*)take(tcm_bytes_r(dd.tr, TS, dd.chh));
    tmp.sv_lo = (__nv_bfloat16*)take(tcm_bytes_r(dd.tr, TS, dd.cdp));
    tmp.sv_raw = (float*)take((size_t)dd.T * dd.n * dd.Aout * sizeof(float));
    if (out) *out = tmp;
    return off;
}

size_t rollout_tc_save_bytes(const drm_rollout* r) {
    RollTcDev d;
    set_dims(r, d);
    return carve_save_tc(d, nullptr, nullptr) + 512;
}

static void fill_io_tc(const drm_rollout* r, RollTcDev& d) {
    const drm_rssm& w = r->rssm;
    d.b_in = w.b_in; d.b_ih = w.b_ih; d.b_hh = w.b_hh; d.b_p1 = w.b_p1; d.b_p2 = w.b_p2;
    for (int l = 0; l < r->actor.n_linear; ++l) d.ba[l] = r->actor.b[l];
    d.start_stoch = r->start_stoch; d.start_deter = r->start_deter; d.eps_prior = r->eps_prior; d.noise_act = r->noise_act;
    d.prior_mean = r->prior_mean; d.prior_std = r->prior_std; d.prior_stoch = r->prior_stoch; d.deter = r->deter;
    d.action_out = r->action_out;
    d.step_flags = r->step_flags;
}

int rollout_tc_fwd(const drm_rollout* r, cudaStream_t stream) {
    DRM_REQUIRE(rollout_tc_supported(r), DRM_E_SHAPE, "shape not supported by the bf16 rollout");
    RollTcPlan pl;
    uint8_t* ws = reinterpret_cast<uint8_t*>(rup_sz((size_t)r->workspace, 256));
    int rc = plan_fwd(r, ws, pl); if (rc) return rc;
    DRM_REQUIRE(r->workspace && r->workspace_bytes >= pl.wbytes + 512, DRM_E_WORKSPACE, "workspace too small (bf16 rollout)");
    RollTcDev& d = pl.dev;
    fill_io_tc(r, d);
    d.wstream = ws;
    const int grid = d.ntiles < num_sms() ? d.ntiles : num_sms();
    if (r->training) {
        DRM_REQUIRE(r->save && r->save_bytes >= rollout_tc_save_bytes(r), DRM_E_WORKSPACE, "save buffer too small (bf16 rollout)");
        RollTcDev carved;
        carve_save_tc(d, reinterpret_cast<uint8_t*>(rup_sz((size_t)r->save, 256)), &carved);
        d = carved;
        d.lo_per_cta = 0;
    } else {
        // no save buffer: the lo image of the carried deter lives in one workspace slot per CTA
        DRM_REQUIRE(r->workspace_bytes >= rollout_tc_fwd_workspace_bytes(r), DRM_E_WORKSPACE, "workspace too small (bf16 rollout)");
        d.sv_lo = reinterpret_cast<__nv_bfloat16*>(ws + rup_sz(pl.wbytes, 256));
        d.lo_per_cta = 1;
    }
    rc = launch_pack_tc(pl.jobs, stream); if (rc) return rc;
    ProfScope prof(PROF_ROLLOUT_FWD, stream);
    if (d.tr == 64) {
        DRM_CUDA_OK(cudaFuncSetAttribute(rollout_fwd_tc_kernel<64>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)pl.smem_bytes));
        rollout_fwd_tc_kernel<64><<<grid, kChainThreads, pl.smem_bytes, stream>>>(d);
    } else {
        DRM_CUDA_OK(cudaFuncSetAttribute(rollout_fwd_tc_kernel<128>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)pl.smem_bytes));
        rollout_fwd_tc_kernel<128><<<grid, kChainThreads, pl.smem_bytes, stream>>>(d);
    }
    DRM_CUDA_OK(cudaGetLastError());
    return DRM_OK;
}

static int finish_smem_bwd(RollTcDev& d, size_t& smem_bytes) {
    TcProgram& pg = d.pg;
    const int hb_chunks = max(max(d.chh, d.cha), d.cdp);
    d.ga_off = 1024;
    d.gb_off = d.ga_off + (uint32_t)hb_chunks * (d.tr * 16);
    d.g2_off = d.gb_off + (uint32_t)d.cdp * (d.tr * 16);
    const int g2_chunks = 2 * d.cs > 2 ? 2 * d.cs : 2;
    pg.stage_off = (uint32_t)rup_sz(d.g2_off + (size_t)g2_chunks * (d.tr * 16), 1024);
    pg.stage_bytes = 32 * 1024;
    int stages = (int)(((size_t)kMaxSmem - pg.stage_off) / pg.stage_bytes);
    if (stages > 4) stages = 4;
    DRM_REQUIRE(stages >= 2, DRM_E_SHAPE, "model too wide for the bf16 rollout's shared memory");
    pg.n_stages = stages;
    smem_bytes = pg.stage_off + (size_t)stages * pg.stage_bytes;
    return DRM_OK;
}

static int plan_bwd(const drm_rollout* r, bool want_actor, uint8_t* wbase, RollTcPlan& pl) {
    const drm_rssm& w = r->rssm;
    const drm_mlp& a = r->actor;
    RollTcDev& d = pl.dev;
    set_dims(r, d);
    d.want_actor = want_actor ? 1 : 0;
    int rc = finish_smem_bwd(d, pl.smem_bytes); if (rc) return rc;
    const int S = w.S, D = w.D, Hd = w.Hd, A = w.A;
    const int Dp = d.cdp * 8, Hp = d.chh * 8, Ap = d.cha * 8;
    const int colC = (d.chh > d.cha ? d.chh : d.cha) * 8, colAS = colC + Dp;
    // half tiles: the stoch columns must fill the first column half exactly (lanes 0..15), the action the second
    const int nAS = d.tr == 64 ? 2 * d.Sp : rup(d.Sp + 16, 16);
    DRM_REQUIRE(d.tr != 64 || d.Sp >= 16, DRM_E_SHAPE, "stoch too narrow for half tiles");
    DRM_REQUIRE(colAS + nAS <= kTmemCols, DRM_E_SHAPE, "model too wide for the bf16 rollout's tensor memory");
    const uint32_t ga = d.ga_off, gb = d.gb_off, g2 = d.g2_off;
    pl.jobs.n = 0; d.pg.n_ops = 0; d.pg.n_phases = 0;
    ProgBuilder b{&d.pg, &pl.jobs, wbase, 0, true, d.tr};
    PackTcJob* j;
    // 1: d prior hidden = d[mean|std] . W_p2
    b.add_op(g2, 2 * d.cs, 0, 0, Hp, 0, 0, 1);
    j = b.add_job(w.W_p2, Hd, 1, 0, Hd, Hp, 2 * d.cs);
    ProgBuilder::seg(j, 0, S, 0); ProgBuilder::seg(j, d.Sp, S, S);
    // 2: d deter += dhid . W_p1
    b.add_op(ga, d.chh, 0, 0, Dp, colC, 1, 1);
    j = b.add_job(w.W_p1, D, 1, 0, D, Dp, d.chh); ProgBuilder::seg(j, 0, Hd, 0);
    // 3: dx = dpr.W_ir + dpz.W_iz ; d deter_prev += dpr.W_hr + dpz.W_hz
    b.add_op(ga, d.cdp, gb, d.cdp, Hp, 0, 0, 0);
    j = b.add_job(w.W_ih, Hd, 1, 0, Hd, Hp, d.cdp); ProgBuilder::seg(j, 0, D, 0);
    j = b.add_job(w.W_ih, Hd, 1, 0, Hd, Hp, d.cdp); ProgBuilder::seg(j, 0, D, D);
    b.add_op(ga, d.cdp, gb, d.cdp, Dp, colC, 1, 1);
    j = b.add_job(w.W_hh, D, 1, 0, D, Dp, d.cdp); ProgBuilder::seg(j, 0, D, 0);
    j = b.add_job(w.W_hh, D, 1, 0, D, Dp, d.cdp); ProgBuilder::seg(j, 0, D, D);
    // 4: dx += din.W_in ; d deter_prev += dhn.W_hn
    b.add_op(ga, d.cdp, 0, 0, Hp, 0, 1, 0);
    j = b.add_job(w.W_ih, Hd, 1, 0, Hd, Hp, d.cdp); ProgBuilder::seg(j, 0, D, 2 * D);
    b.add_op(gb, d.cdp, 0, 0, Dp, colC, 1, 1);
    j = b.add_job(w.W_hh, D, 1, 0, D, Dp, d.cdp); ProgBuilder::seg(j, 0, D, 2 * D);
    // 5: d[stoch | action] = dx_pre . W_in   (columns: stoch at 0.., action at Sp..)
    b.add_op(ga, d.chh, 0, 0, nAS, colAS, 0, 1);
    j = b.add_job(w.W_in, A + S, 1, A, S, nAS, d.chh); ProgBuilder::seg(j, 0, Hd, 0);
    if (j) { j->n2_start = d.Sp; j->n2_valid = A; j->n2_src0 = 0; }
    if (want_actor) {
        const int last = a.n_linear - 1;
        b.add_op(g2, 2, 0, 0, Ap, 0, 0, 1);
        j = b.add_job(a.W[last], a.hidden, 1, 0, a.hidden, Ap, 2); ProgBuilder::seg(j, 0, a.out_dim, 0);
        for (int l = last - 1; l >= 1; --l) {
            b.add_op(ga, d.cha, 0, 0, Ap, 0, 0, 1);
            j = b.add_job(a.W[l], a.hidden, 1, 0, a.hidden, Ap, d.cha); ProgBuilder::seg(j, 0, a.hidden, 0);
        }
    }
    DRM_REQUIRE(b.ok, DRM_E_SHAPE, "actor too deep for the bf16 rollout");
    pl.wbytes = b.woff;
    return DRM_OK;
}

// workspace of the backward: [weight images][dY images of the actor][per-CTA stash]
static size_t carve_bwd_ws(const RollTcDev& dd, size_t wbytes, uint8_t* base, RollTcDev* out) {
    const size_t TS = (size_t)dd.T * dd.ntiles;
    size_t off = rup_sz(wbytes, 256);
    auto take = [&](size_t bytes) -> uint8_t* {
        uint8_t* ptr = base ? base + off : nullptr;
        off += rup_sz(bytes, 256);
        return ptr;
    };
    RollTcDev tmp = dd;
    for (int l = 0; l < dd.an; ++l) tmp.dys_a[l] = (__nv_bfloat16*)take(tcm_bytes_r(dd.tr, TS, l == dd.an - 1 ? 2 : dd.cha));
    const int grid = dd.ntiles < num_sms() ? dd.ntiles : num_sms();
    tmp.stash = (__nv_bfloat16*)take((size_t)grid * 2 * dd.cdp * (dd.tr * 16));
    tmp.bwd_flags = (int*)take(sizeof(int) * dd.T);
    if (out) *out = tmp;
    return off;
}

size_t rollout_tc_bwd_workspace_bytes(const drm_rollout* r) {
    RollTcPlan pl;
    if (!rollout_tc_supported(r) || plan_bwd(r, true, nullptr, pl) != DRM_OK) return 0;
    return carve_bwd_ws(pl.dev, pl.wbytes, nullptr, nullptr) + 1024;
}

int rollout_tc_bwd(const drm_rollout* r, const drm_rollout_grad* g, cudaStream_t stream) {
    DRM_REQUIRE(rollout_tc_supported(r), DRM_E_SHAPE, "shape not supported by the bf16 rollout");
    DRM_REQUIRE(!g->want_rssm_grad, DRM_E_SHAPE,
                "the bf16 imagination rollout computes actor gradients only (world model frozen, dreamer_algo.py:293); "
                "run the rollout in fp32 for transition weight gradients");
    const drm_mlp& a = r->actor;
    const bool want_actor = g->want_actor_grad != 0;
    RollTcPlan pl;
    uint8_t* ws = reinterpret_cast<uint8_t*>(rup_sz((size_t)r->workspace, 256));
    int rc = plan_bwd(r, want_actor, ws, pl); if (rc) return rc;
    DRM_REQUIRE(r->workspace && r->workspace_bytes >= rollout_tc_bwd_workspace_bytes(r), DRM_E_WORKSPACE,
                "workspace too small (bf16 rollout backward)");
    DRM_REQUIRE(r->save && r->save_bytes >= rollout_tc_save_bytes(r), DRM_E_WORKSPACE,
                "save buffer missing: run drm_rollout_fwd with training=1 first");
    RollTcDev d = pl.dev;
    {
        RollTcDev carved;
        carve_save_tc(d, reinterpret_cast<uint8_t*>(rup_sz((size_t)r->save, 256)), &carved);
        d = carved;
        carve_bwd_ws(d, pl.wbytes, ws, &carved);
        d = carved;
    }
    fill_io_tc(r, d);
    d.wstream = ws;
    d.g_prior_mean = g->g_prior_mean; d.g_prior_std = g->g_prior_std; d.g_prior_stoch = g->g_prior_stoch;
    d.g_deter = g->g_deter; d.g_action_out = g->g_action_out;
    d.g_wait_flags = g->g_wait_flags; d.g_wait_target = g->g_wait_target;
    if (want_actor) {
        for (int l = 0; l < a.n_linear; ++l) {
            const int nout = l == a.n_linear - 1 ? a.out_dim : a.hidden;
            const int nin = l == 0 ? a.in_dim : a.hidden;
            DRM_REQUIRE(g->actor.dW[l] && g->actor.db[l], DRM_E_ARG, "null actor gradient sink");
            pl.jobs.zero(g->actor.dW[l], (size_t)nout * nin);
            pl.jobs.zero(g->actor.db[l], (size_t)nout);
        }
        d.db_a_last = g->actor.db[a.n_linear - 1];
    }
    // The weight-gradient GEMMs can follow the rollout step by step on another stream (per-step counters) instead of
    // starting when it has finished.  Only after a PREPARE call: the counters are zeroed there.
    cudaStream_t wg_stream = stream;
    int* flags = d.bwd_flags;
    d.bwd_flags = nullptr;
    if (want_actor) pl.jobs.zero(reinterpret_cast<float*>(flags), (size_t)d.T);
    // Both grids must be resident, and a waiting weight-gradient CTA (64-row images: 49 KB, 256 TMEM columns) must never share
    // an SM with a rollout CTA, whose TMEM allocation it could block: the rollout's shared memory has to exclude that.
    const bool excl = pl.smem_bytes + 50 * 1024 > (size_t)kMaxSmem;
    if (want_actor && g->wgrad_stream && g->phase == DRM_PHASE_RUN && d.ntiles + 16 <= num_sms() && excl) {
        wg_stream = (cudaStream_t)g->wgrad_stream;
        d.bwd_flags = flags;
    }
    if (g->phase != DRM_PHASE_RUN) {
        rc = launch_pack_tc(pl.jobs, stream); if (rc) return rc;
        if (g->phase == DRM_PHASE_PREPARE) return DRM_OK;
    }
    {
        const int grid = d.ntiles < num_sms() ? d.ntiles : num_sms();
        ProfScope prof(PROF_ROLLOUT_BWD, stream);
        if (d.tr == 64) {
            DRM_CUDA_OK(cudaFuncSetAttribute(rollout_bwd_tc_kernel<64>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)pl.smem_bytes));
            rollout_bwd_tc_kernel<64><<<grid, kChainThreads, pl.smem_bytes, stream>>>(d);
        } else {
            DRM_CUDA_OK(cudaFuncSetAttribute(rollout_bwd_tc_kernel<128>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)pl.smem_bytes));
            rollout_bwd_tc_kernel<128><<<grid, kChainThreads, pl.smem_bytes, stream>>>(d);
        }
        DRM_CUDA_OK(cudaGetLastError());
    }
    if (want_actor) {
        WgParams wg;
        memset(&wg, 0, sizeof(wg));
        wg.n_layers = a.n_linear; wg.row_tiles = d.T * d.ntiles; wg.tile_rows = d.tr;
        const int last = a.n_linear - 1;
        for (int l = 0; l <= last; ++l) {
            WgLayer& L = wg.L[l];
            L.dy = d.dys_a[l]; L.dy_cpt = l == last ? 2 : d.cha;
            L.nout = l == last ? a.out_dim : a.hidden;
            if (l == 0) {
                L.nseg = 2;
                L.seg[0] = WgSeg{d.sv_feat, d.cs + d.cdp, 0, (d.S + 7) / 8, 0, d.S};
                L.seg[1] = WgSeg{d.sv_feat, d.cs + d.cdp, d.cs, d.cd, d.S, d.D};
                L.ld_dw = a.in_dim;
            } else {
                L.nseg = 1;
                L.seg[0] = WgSeg{d.sv_ah[l - 1], d.cha, 0, (a.hidden + 7) / 8, 0, a.hidden};
                L.ld_dw = a.hidden;
            }
            L.dW = g->actor.dW[l];
            L.db = l == last ? nullptr : g->actor.db[l];
        }
        int slots = 0;
        if (d.bwd_flags) {
            // row tiles in reverse time order, each gated by its step's counter; the grid must be resident BESIDE the rollout
            wg.wait_flags = d.bwd_flags; wg.wait_target = d.ntiles;
            wg.tiles_per_step = d.ntiles; wg.reverse = 1;
            slots = (num_sms() - d.ntiles) * (d.tr == 64 ? 2 : 1);
        }
        wgrad_tc_plan(wg, slots);
        rc = launch_wgrad_tc(wg, wg_stream); if (rc) return rc;
    }
    return DRM_OK;
}

int rollout_tc_tiles(const drm_rollout* r) {
    const int tr = tile_rows_for(r->n);
    return (r->n + tr - 1) / tr;
}

size_t rollout_tc_fwd_workspace_bytes(const drm_rollout* r) {
    RollTcPlan pl;
    if (!rollout_tc_supported(r) || plan_fwd(r, nullptr, pl) != DRM_OK) return 0;
    const int grid = pl.dev.ntiles < num_sms() ? pl.dev.ntiles : num_sms();
    return rup_sz(pl.wbytes, 256) + tcm_bytes_r(pl.dev.tr, (size_t)num_sms(), pl.dev.cdp) + 1024;
}

}  // namespace tc
void warm_rollout_tc() {
    warm_kernel(tc::rollout_fwd_tc_kernel<64>); warm_kernel(tc::rollout_fwd_tc_kernel<128>);
    warm_kernel(tc::rollout_bwd_tc_kernel<64>); warm_kernel(tc::rollout_bwd_tc_kernel<128>);
}

}  // namespace drm

#ifdef DRM_TIMELINE
// Debug only: copy out (and reset) the handshake stamps CTA 0 of the rollout kernels recorded.
extern "C" int drm_debug_timeline(long long* stamps, int* counts) {
    cudaDeviceSynchronize();
    cudaMemcpyFromSymbol(stamps, drm::tc::g_tl, sizeof(long long) * 4 * 4096);
    cudaMemcpyFromSymbol(counts, drm::tc::g_tl_n, sizeof(int) * 4);
    int zero[4] = {0, 0, 0, 0};
    cudaMemcpyToSymbol(drm::tc::g_tl_n, zero, sizeof(zero));
    return 0;
}
#endif
