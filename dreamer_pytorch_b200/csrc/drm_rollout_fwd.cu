// Persistent forward rollout kernel: one launch runs all `T` sequential steps of
//   POLICY          rollout_policy        rnns.py:224-244 + agent.py:101-117 + action.py:51-66
//   TRANSITION      rollout_transition    rnns.py:207-222 (step: rnns.py:98-107)
//   REPRESENTATION  rollout_representation rnns.py:179-205 (step: rnns.py:149-159)
// for a tile of R rows per CTA.  The recurrent state (stoch, deter) never leaves shared memory
// between steps; weights stream from L2 (packed, coalesced).  fp32 FFMA path.
#include "drm_rollout.cuh"
#include "drm_philox.cuh"

namespace drm {

template <int R>
__global__ void __launch_bounds__(rollout_threads(R)) rollout_fwd_kernel(const RollDev p) {
    extern __shared__ __align__(16) float smem[];
    const int S = p.S, D = p.D, Hd = p.Hd, A = p.A;
    const int ldXH = Hd + D;
    float* sS = smem;                    // [R][S4]    stoch carried between steps
    float* sXH = sS + R * p.S4;          // [R][Hd+D]  x | deter (h)
    float* sAct = sXH + R * ldXH;        // [R][A4]
    float* sU = sAct + R * p.A4;         // [R][U]     actor hidden ping-pong / GRU gates / MLP hidden
    float* sMS = sU + R * p.U;           // [R][MS4]   (mean | raw std)
    float* sRaw = sMS + R * p.MS4;       // [R][Aout4] actor output
    const int U = p.U;
    const size_t n = (size_t)p.n;
    const bool policy = p.mode == DRM_MODE_POLICY;
    const bool repr = p.mode == DRM_MODE_REPRESENTATION;
    const bool save = p.training != 0;
    const Seg none{nullptr, nullptr, 0, 0};
    const NoiseKey nk = p.nz ? load_noise_key(p.nz) : NoiseKey{0u, 0u, 0u, 0u};

    // wide CTAs (few rows): split-K scratch behind the row buffers, see cta_linear_splitk
    // wide CTAs: whatever the launch left of shared memory behind the row buffers is split-K scratch (cta_linear)
    unsigned dyn_bytes;
    asm("mov.u32 %0, %%dynamic_smem_size;" : "=r"(dyn_bytes));
    const int row_floats = R * fwd_smem_floats_per_row(p);
    const KScratch ksc = (R <= 20 && blockDim.x > kThreads && (int)(dyn_bytes / 4) > row_floats)
                             ? KScratch{smem + row_floats, (int)(dyn_bytes / 4) - row_floats} : KScratch{nullptr, 0};
    for (int tile = blockIdx.x; tile * R < p.n; tile += gridDim.x) {
        const int row0 = tile * R;
        const int total = R * fwd_smem_floats_per_row(p);
        for (int i = threadIdx.x; i < total; i += blockDim.x) smem[i] = 0.f;
        __syncthreads();
        for (int i = threadIdx.x; i < R * S; i += blockDim.x) {
            const int r = i / S, c = i - r * S;
            if (row0 + r < p.n) sS[r * p.S4 + c] = p.start_stoch[(size_t)(row0 + r) * S + c];
        }
        for (int i = threadIdx.x; i < R * D; i += blockDim.x) {
            const int r = i / D, c = i - r * D;
            if (row0 + r < p.n) sXH[r * ldXH + Hd + c] = p.start_deter[(size_t)(row0 + r) * D + c];
        }
        __syncthreads();

        for (int t = 0; t < p.T; ++t) {
            const size_t trow0 = (size_t)t * n + row0;  // first row of this tile inside [T*n]
            // ------------------------------------------------------------------ action
            if (policy) {
                const int ah = p.ah;
                // layer 0 on feat = [stoch | deter]
                {
                    const float* bias = p.ba[0];
                    float* svh = save ? p.sv.h[0] : nullptr;
                    auto epi = [&](int r, int col, float v) {
                        v = elu(v + bias[col]);
                        sU[r * U + col] = v;
                        if (svh && row0 + r < p.n) svh[(trow0 + r) * ah + col] = v;
                    };
                    Seg s0{p.Wa[0], sS, p.S4, p.S4};
                    Seg s1{p.Wa[0] + (size_t)p.S4 * p.ld_a[0], sXH + Hd, D, ldXH};
                    cta_linear<R>(s0, s1, p.ld_a[0], ah, epi, ksc);
                }
                __syncthreads();
                int cur = 0;  // offset of the current hidden buffer inside sU (0 or ah)
                for (int l = 1; l < p.an - 1; ++l) {
                    const float* bias = p.ba[l];
                    float* svh = save ? p.sv.h[l] : nullptr;
                    const int nxt = cur ? 0 : ah;
                    auto epi = [&](int r, int col, float v) {
                        v = elu(v + bias[col]);
                        sU[r * U + nxt + col] = v;
                        if (svh && row0 + r < p.n) svh[(trow0 + r) * ah + col] = v;
                    };
                    Seg s0{p.Wa[l], sU + cur, ah, U};
                    cta_linear<R>(s0, none, p.ld_a[l], ah, epi, ksc);
                    __syncthreads();
                    cur = nxt;
                }
                {
                    const int l = p.an - 1;
                    const float* bias = p.ba[l];
                    auto epi = [&](int r, int col, float v) { sRaw[r * p.Aout4 + col] = v + bias[col]; };
                    if (p.Aout <= 16) {
                        cta_linear_narrow(p.Wa_last, ah, ah, sU + cur, U, R, p.Aout, epi);
                    } else {
                        Seg s0{p.Wa[l], sU + cur, ah, U};
                        cta_linear<R>(s0, none, p.ld_a[l], p.Aout, epi, ksc);
                    }
                }
                __syncthreads();
                // sample the action (agent.py:101-117)
                if (p.act_dist == DRM_DIST_TANH_NORMAL) {
                    for (int i = threadIdx.x; i < R * A; i += blockDim.x) {
                        const int r = i / A, j = i - r * A;
                        const bool ok = row0 + r < p.n;
                        const float rm = sRaw[r * p.Aout4 + j], rs = sRaw[r * p.Aout4 + A + j];
                        const float mean = p.mean_scale * tanhf(rm / p.mean_scale);
                        const float std = softplusf_(rs + p.raw_init_std) + p.act_min_std;
                        const float e = ok ? noise_at(p.noise_act, nk, NZ_NOISE_ACT, NZ_NORMAL, (trow0 + r) * A + j) : 0.f;
                        const float a = tanhf(mean + std * e);
                        sAct[r * p.A4 + j] = a;
                        if (ok) {
                            p.action_out[(trow0 + r) * A + j] = a;
                            if (save) {
                                p.sv.raw[(trow0 + r) * p.Aout + j] = rm;
                                p.sv.raw[(trow0 + r) * p.Aout + A + j] = rs;
                            }
                        }
                    }
                } else {
                    // categorical variants: one thread per row (A is small)
                    for (int r = threadIdx.x; r < R; r += blockDim.x) {
                        const bool ok = row0 + r < p.n;
                        const float* lg = sRaw + r * p.Aout4;
                        auto gz = [&](int j) { return noise_at(p.noise_act, nk, NZ_NOISE_ACT, NZ_GUMBEL, (trow0 + r) * A + j); };
                        float mx = -INFINITY;
                        for (int j = 0; j < A; ++j) mx = fmaxf(mx, lg[j]);
                        if (p.act_dist == DRM_DIST_ONE_HOT) {
                            float den = 0.f;
                            for (int j = 0; j < A; ++j) den += expf(lg[j] - mx);
                            int best = 0; float bv = -INFINITY;
                            for (int j = 0; j < A; ++j) {
                                const float s = lg[j] + (ok ? gz(j) : 0.f);
                                if (s > bv) { bv = s; best = j; }
                            }
                            for (int j = 0; j < A; ++j) {
                                const float pj = expf(lg[j] - mx) / den;
                                // straight-through value: (sample + probs) - probs, agent.py:110-112
                                const float a = ((j == best ? 1.f : 0.f) + pj) - pj;
                                sAct[r * p.A4 + j] = a;
                                if (ok) {
                                    p.action_out[(trow0 + r) * A + j] = a;
                                    if (save) p.sv.raw[(trow0 + r) * p.Aout + j] = lg[j];
                                }
                            }
                        } else {  // relaxed one-hot: softmax((logits + g) / temperature)
                            float smx = -INFINITY;
                            for (int j = 0; j < A; ++j) smx = fmaxf(smx, (lg[j] + (ok ? gz(j) : 0.f)) / p.temperature);
                            float den = 0.f;
                            for (int j = 0; j < A; ++j) den += expf((lg[j] + (ok ? gz(j) : 0.f)) / p.temperature - smx);
                            for (int j = 0; j < A; ++j) {
                                const float a = expf((lg[j] + (ok ? gz(j) : 0.f)) / p.temperature - smx) / den;
                                sAct[r * p.A4 + j] = a;
                                if (ok) {
                                    p.action_out[(trow0 + r) * A + j] = a;
                                    if (save) p.sv.raw[(trow0 + r) * p.Aout + j] = lg[j];
                                }
                            }
                        }
                    }
                }
            } else {
                for (int i = threadIdx.x; i < R * A; i += blockDim.x) {
                    const int r = i / A, j = i - r * A;
                    sAct[r * p.A4 + j] = (row0 + r < p.n) ? p.action_in[(trow0 + r) * A + j] : 0.f;
                }
            }
            __syncthreads();
            // ------------------------------------------------------------------ x = ELU(W_in [a|stoch] + b)
            {
                float* svx = save ? p.sv.x : nullptr;
                auto epi = [&](int r, int col, float v) {
                    v = elu(v + p.b_in[col]);
                    sXH[r * ldXH + col] = v;
                    if (svx && row0 + r < p.n) svx[(trow0 + r) * Hd + col] = v;
                };
                Seg s0{p.Win, sAct, p.A4, p.A4};
                Seg s1{p.Win + (size_t)p.A4 * p.ld_in, sS, p.S4, p.S4};
                cta_linear<R>(s0, s1, p.ld_in, Hd, epi, ksc);
            }
            __syncthreads();
            // ------------------------------------------------------------------ GRU gates (torch GRUCell)
            {
                auto epi_rz = [&](int r, int col, float v) { sU[r * U + col] = sigmoidf_(v + p.b_ih[col] + p.b_hh[col]); };
                Seg s0{p.Wrz, sXH, Hd, ldXH};
                Seg s1{p.Wrz + (size_t)Hd * p.ld_rz, sXH + Hd, D, ldXH};
                cta_linear<R>(s0, s1, p.ld_rz, 2 * D, epi_rz, ksc);
                auto epi_in = [&](int r, int col, float v) { sU[r * U + 2 * D + col] = v + p.b_ih[2 * D + col]; };
                Seg s2{p.Wni, sXH, Hd, ldXH};
                cta_linear<R>(s2, none, p.ld_n, D, epi_in, ksc);
                auto epi_hn = [&](int r, int col, float v) { sU[r * U + 3 * D + col] = v + p.b_hh[2 * D + col]; };
                Seg s3{p.Wnh, sXH + Hd, D, ldXH};
                cta_linear<R>(s3, none, p.ld_n, D, epi_hn, ksc);
            }
            __syncthreads();
            for (int i = threadIdx.x; i < R * D; i += blockDim.x) {
                const int r = i / D, c = i - r * D;
                const float rr = sU[r * U + c], zz = sU[r * U + D + c];
                const float in_ = sU[r * U + 2 * D + c], hn = sU[r * U + 3 * D + c];
                const float nn = tanhf(in_ + rr * hn);
                const float hprev = sXH[r * ldXH + Hd + c];
                const float hnew = (1.f - zz) * nn + zz * hprev;
                sXH[r * ldXH + Hd + c] = hnew;
                if (row0 + r < p.n) {
                    const size_t o = (trow0 + r) * D + c;
                    p.deter[o] = hnew;
                    if (save) { p.sv.r[o] = rr; p.sv.z[o] = zz; p.sv.nn[o] = nn; p.sv.hn[o] = hn; }
                }
            }
            __syncthreads();
            // ------------------------------------------------------------------ prior MLP
            {
                float* svh = save ? p.sv.hid : nullptr;
                auto epi = [&](int r, int col, float v) {
                    v = elu(v + p.b_p1[col]);
                    sU[r * U + col] = v;
                    if (svh && row0 + r < p.n) svh[(trow0 + r) * Hd + col] = v;
                };
                Seg s0{p.Wp1, sXH + Hd, D, ldXH};
                cta_linear<R>(s0, none, p.ld_p1, Hd, epi, ksc);
            }
            __syncthreads();
            {
                auto epi = [&](int r, int col, float v) { sMS[r * p.MS4 + col] = v + p.b_p2[col]; };
                Seg s0{p.Wp2, sU, Hd, U};
                cta_linear<R>(s0, none, p.ld_p2, 2 * S, epi, ksc);
            }
            __syncthreads();
            for (int i = threadIdx.x; i < R * S; i += blockDim.x) {
                const int r = i / S, c = i - r * S;
                const bool ok = row0 + r < p.n;
                const float mean = sMS[r * p.MS4 + c];
                const float std = softplusf_(sMS[r * p.MS4 + S + c]) + p.min_std;
                const size_t o = (trow0 + r) * S + c;
                const float e = ok ? noise_at(p.eps_prior, nk, NZ_EPS_PRIOR, NZ_NORMAL, o) : 0.f;
                const float st = mean + std * e;
                if (!repr) sS[r * p.S4 + c] = st;
                if (ok) { p.prior_mean[o] = mean; p.prior_std[o] = std; p.prior_stoch[o] = st; }
            }
            __syncthreads();
            // ------------------------------------------------------------------ posterior MLP
            if (repr) {
                {
                    float* svh = save ? p.sv.hidq : nullptr;
                    auto epi = [&](int r, int col, float v) {
                        const bool ok = row0 + r < p.n;
                        v += p.b_q1[col] + (ok ? p.embproj[(trow0 + r) * Hd + col] : 0.f);
                        v = elu(v);
                        sU[r * U + col] = v;
                        if (svh && ok) svh[(trow0 + r) * Hd + col] = v;
                    };
                    Seg s0{p.Wq1, sXH + Hd, D, ldXH};
                    cta_linear<R>(s0, none, p.ld_q1, Hd, epi, ksc);
                }
                __syncthreads();
                {
                    auto epi = [&](int r, int col, float v) { sMS[r * p.MS4 + col] = v + p.b_q2[col]; };
                    Seg s0{p.Wq2, sU, Hd, U};
                    cta_linear<R>(s0, none, p.ld_q2, 2 * S, epi, ksc);
                }
                __syncthreads();
                for (int i = threadIdx.x; i < R * S; i += blockDim.x) {
                    const int r = i / S, c = i - r * S;
                    const bool ok = row0 + r < p.n;
                    const float mean = sMS[r * p.MS4 + c];
                    const float std = softplusf_(sMS[r * p.MS4 + S + c]) + p.min_std;
                    const size_t o = (trow0 + r) * S + c;
                    const float e = ok ? noise_at(p.eps_post, nk, NZ_EPS_POST, NZ_NORMAL, o) : 0.f;
                    const float st = mean + std * e;
                    sS[r * p.S4 + c] = st;
                    if (ok) { p.post_mean[o] = mean; p.post_std[o] = std; p.post_stoch[o] = st; }
                }
                __syncthreads();
            }
        }
        __syncthreads();
    }
}

int launch_rollout_fwd(const RollDev& p, cudaStream_t stream) {
    const size_t per_row = (size_t)fwd_smem_floats_per_row(p) * sizeof(float);
    const int Rv = pick_rows(p.n, per_row, 0);
    DRM_REQUIRE(Rv > 0, DRM_E_SHAPE, "model too wide for shared memory");
    // wide CTAs (<= 20 rows) take split-K scratch behind the row buffers: as much as they can use, or what is left
    size_t scratch = (size_t)splitk_scratch_floats(Rv) * sizeof(float);
    if (per_row * Rv + scratch > (size_t)kMaxSmem) scratch = ((size_t)kMaxSmem - per_row * Rv) & ~(size_t)15;
    if (Rv == 8 && scratch < (size_t)splitk_scratch_floats(Rv) * sizeof(float)) scratch = 0;  // the 32-column form wants all of it
    const bool wide = Rv <= 20 && (Rv != 8 || scratch > 0);
    const size_t smem = per_row * Rv + (wide ? scratch : 0);
    const int nthreads = wide ? rollout_threads(Rv) : kThreads;
    const int tiles = (p.n + Rv - 1) / Rv;
    const int grid = tiles < num_sms() ? tiles : num_sms();
    ProfScope prof(PROF_ROLLOUT_FWD, stream);
    DRM_DISPATCH_R(Rv, {
        DRM_CUDA_OK(cudaFuncSetAttribute(rollout_fwd_kernel<R>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        rollout_fwd_kernel<R><<<grid, nthreads, smem, stream>>>(p);
    });
    DRM_CUDA_OK(cudaGetLastError());
    return DRM_OK;
}

void warm_rollout_fwd() { warm_kernel(rollout_fwd_kernel<1>); warm_kernel(rollout_fwd_kernel<2>); warm_kernel(rollout_fwd_kernel<4>); warm_kernel(rollout_fwd_kernel<8>); warm_kernel(rollout_fwd_kernel<12>); warm_kernel(rollout_fwd_kernel<16>); warm_kernel(rollout_fwd_kernel<20>); warm_kernel(rollout_fwd_kernel<24>); warm_kernel(rollout_fwd_kernel<32>); }

}  // namespace drm
