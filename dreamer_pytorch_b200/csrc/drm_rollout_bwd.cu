// Persistent reverse-time rollout kernel: data-gradients of all T steps in one launch, for a
// tile of R rows per CTA.  The carried state gradient (d stoch, d deter) stays in shared memory;
// pre-activation gradients of every Linear are streamed out as [T,n,width] "dY" arrays that the
// weight-gradient GEMMs (drm_mlp.cu: wgrad_kernel) consume afterwards.
// Mirrors autograd through rnns.py:98-107,149-159,224-244, action.py:51-66, agent.py:101-117.
#include "drm_rollout.cuh"
#include "drm_philox.cuh"

namespace drm {

template <int R>
__global__ void __launch_bounds__(rollout_threads(R)) rollout_bwd_kernel(const RollDev p) {
    extern __shared__ __align__(16) float smem[];
    const int S = p.S, D = p.D, Hd = p.Hd, A = p.A, U = p.U;
    float* cS = smem;                 // [R][S4]   carried d stoch
    float* cD = cS + R * p.S4;        // [R][D]    carried d deter
    float* sMS = cD + R * D;          // [R][MS4]
    float* sV = sMS + R * p.MS4;      // [R][Hd]
    float* sDH = sV + R * Hd;         // [R][D]
    float* sU = sDH + R * D;          // [R][U]    gate gradients / actor ping-pong
    float* sAS = sU + R * U;          // [R][AS4]  d[action | stoch_prev]
    float* sRaw = sAS + R * p.AS4;    // [R][Aout4]
    const size_t n = (size_t)p.n;
    const bool policy = p.mode == DRM_MODE_POLICY;
    const bool repr = p.mode == DRM_MODE_REPRESENTATION;
    const Seg none{nullptr, nullptr, 0, 0};
    const NoiseKey nk = p.nz ? load_noise_key(p.nz) : NoiseKey{0u, 0u, 0u, 0u};

    // wide CTAs (few rows): split-K scratch behind the row buffers, see cta_linear_splitk
    // wide CTAs: whatever the launch left of shared memory behind the row buffers is split-K scratch (cta_linear)
    unsigned dyn_bytes;
    asm("mov.u32 %0, %%dynamic_smem_size;" : "=r"(dyn_bytes));
    const int row_floats = R * bwd_smem_floats_per_row(p);
    const KScratch ksc = (R <= 20 && blockDim.x > kThreads && (int)(dyn_bytes / 4) > row_floats)
                             ? KScratch{smem + row_floats, (int)(dyn_bytes / 4) - row_floats} : KScratch{nullptr, 0};
    for (int tile = blockIdx.x; tile * R < p.n; tile += gridDim.x) {
        const int row0 = tile * R;
        const int total = R * bwd_smem_floats_per_row(p);
        for (int i = threadIdx.x; i < total; i += blockDim.x) smem[i] = 0.f;
        __syncthreads();

        for (int t = p.T - 1; t >= 0; --t) {
            const size_t trow0 = (size_t)t * n + row0;
            // ------------------------------------------------------------ posterior head (REPRESENTATION)
            if (repr) {
                for (int i = threadIdx.x; i < R * S; i += blockDim.x) {
                    const int r = i / S, c = i - r * S;
                    float dm = 0.f, ds = 0.f;
                    if (row0 + r < p.n) {
                        const size_t o = (trow0 + r) * S + c;
                        const float dst = (p.g_post_stoch ? p.g_post_stoch[o] : 0.f) + cS[r * p.S4 + c];
                        dm = (p.g_post_mean ? p.g_post_mean[o] : 0.f) + dst;
                        const float dsd = (p.g_post_std ? p.g_post_std[o] : 0.f) + dst * noise_at(p.eps_post, nk, NZ_EPS_POST, NZ_NORMAL, o);
                        ds = dsd * (1.f - expf(-(p.post_std[o] - p.min_std)));  // sigmoid(raw) from softplus(raw)
                        if (p.want_rssm_dy) {
                            p.dy.q2[(trow0 + r) * 2 * S + c] = dm;
                            p.dy.q2[(trow0 + r) * 2 * S + S + c] = ds;
                        }
                    }
                    sMS[r * p.MS4 + c] = dm;
                    sMS[r * p.MS4 + S + c] = ds;
                }
                __syncthreads();
                {
                    auto epi = [&](int r, int col, float v) {
                        float g = 0.f;
                        if (row0 + r < p.n) {
                            g = v * elu_grad_from_out(p.sv.hidq[(trow0 + r) * Hd + col]);
                            if (p.want_q1_dy) p.dy.q1[(trow0 + r) * Hd + col] = g;
                        }
                        sV[r * Hd + col] = g;
                    };
                    Seg s0{p.Wq2, sMS, p.MS4, p.MS4};
                    cta_linear<R>(s0, none, p.ld_q2, Hd, epi, ksc);
                }
                __syncthreads();
                {
                    auto epi = [&](int r, int col, float v) { sDH[r * D + col] = v; };
                    Seg s0{p.Wq1, sV, Hd, Hd};
                    cta_linear<R>(s0, none, p.ld_q1, D, epi, ksc);
                }
                __syncthreads();
            }
            // ------------------------------------------------------------ prior head
            for (int i = threadIdx.x; i < R * S; i += blockDim.x) {
                const int r = i / S, c = i - r * S;
                float dm = 0.f, ds = 0.f;
                if (row0 + r < p.n) {
                    const size_t o = (trow0 + r) * S + c;
                    const float dst = (p.g_prior_stoch ? p.g_prior_stoch[o] : 0.f) + (repr ? 0.f : cS[r * p.S4 + c]);
                    dm = (p.g_prior_mean ? p.g_prior_mean[o] : 0.f) + dst;
                    const float dsd = (p.g_prior_std ? p.g_prior_std[o] : 0.f) + dst * noise_at(p.eps_prior, nk, NZ_EPS_PRIOR, NZ_NORMAL, o);
                    ds = dsd * (1.f - expf(-(p.prior_std[o] - p.min_std)));
                    if (p.want_rssm_dy) {
                        p.dy.p2[(trow0 + r) * 2 * S + c] = dm;
                        p.dy.p2[(trow0 + r) * 2 * S + S + c] = ds;
                    }
                }
                sMS[r * p.MS4 + c] = dm;
                sMS[r * p.MS4 + S + c] = ds;
            }
            __syncthreads();
            {
                auto epi = [&](int r, int col, float v) {
                    float g = 0.f;
                    if (row0 + r < p.n) {
                        g = v * elu_grad_from_out(p.sv.hid[(trow0 + r) * Hd + col]);
                        if (p.want_rssm_dy) p.dy.p1[(trow0 + r) * Hd + col] = g;
                    }
                    sV[r * Hd + col] = g;
                };
                Seg s0{p.Wp2, sMS, p.MS4, p.MS4};
                cta_linear<R>(s0, none, p.ld_p2, Hd, epi, ksc);
            }
            __syncthreads();
            {
                auto epi = [&](int r, int col, float v) {
                    float tot = v + cD[r * D + col] + (repr ? sDH[r * D + col] : 0.f);
                    if (p.g_deter && row0 + r < p.n) tot += p.g_deter[(trow0 + r) * D + col];
                    sDH[r * D + col] = tot;
                };
                Seg s0{p.Wp1, sV, Hd, Hd};
                cta_linear<R>(s0, none, p.ld_p1, D, epi, ksc);
            }
            __syncthreads();
            // ------------------------------------------------------------ GRU cell
            for (int i = threadIdx.x; i < R * D; i += blockDim.x) {
                const int r = i / D, c = i - r * D;
                float dpr = 0.f, dpz = 0.f, din = 0.f, dhn = 0.f, carry = 0.f;
                if (row0 + r < p.n) {
                    const size_t o = (trow0 + r) * D + c;
                    const float dh = sDH[r * D + c];
                    const float rr = p.sv.r[o], zz = p.sv.z[o], nn = p.sv.nn[o], hn = p.sv.hn[o];
                    const float hprev = t > 0 ? p.deter[o - n * D] : p.start_deter[(size_t)(row0 + r) * D + c];
                    const float dn = dh * (1.f - zz);
                    const float dz = dh * (hprev - nn);
                    const float dpn = dn * (1.f - nn * nn);
                    din = dpn;
                    dhn = dpn * rr;
                    dpr = dpn * hn * rr * (1.f - rr);
                    dpz = dz * zz * (1.f - zz);
                    carry = dh * zz;
                    if (p.want_rssm_dy) {
                        const size_t o3 = (trow0 + r) * 3 * D + c;
                        p.dy.ih[o3] = dpr; p.dy.ih[o3 + D] = dpz; p.dy.ih[o3 + 2 * D] = din;
                        p.dy.hh[o3] = dpr; p.dy.hh[o3 + D] = dpz; p.dy.hh[o3 + 2 * D] = dhn;
                    }
                }
                sU[r * U + c] = dpr; sU[r * U + D + c] = dpz; sU[r * U + 2 * D + c] = din; sU[r * U + 3 * D + c] = dhn;
                cD[r * D + c] = carry;
            }
            __syncthreads();
            {
                auto epi_x = [&](int r, int col, float v) {
                    float g = 0.f;
                    if (row0 + r < p.n) {
                        g = v * elu_grad_from_out(p.sv.x[(trow0 + r) * Hd + col]);
                        if (p.want_rssm_dy) p.dy.in[(trow0 + r) * Hd + col] = g;
                    }
                    sV[r * Hd + col] = g;
                };
                Seg s0{p.Wih, sU, 3 * D, U};
                cta_linear<R>(s0, none, p.ld_ih, Hd, epi_x, ksc);
                auto epi_h = [&](int r, int col, float v) { cD[r * D + col] += v; };
                Seg s1{p.Whh, sU, 2 * D, U};
                Seg s2{p.Whh + (size_t)2 * D * p.ld_hh, sU + 3 * D, D, U};
                cta_linear<R>(s1, s2, p.ld_hh, D, epi_h, ksc);
            }
            __syncthreads();
            {
                auto epi = [&](int r, int col, float v) { sAS[r * p.AS4 + col] = v; };
                Seg s0{p.Win, sV, Hd, Hd};
                cta_linear<R>(s0, none, p.ld_in, A + S, epi, ksc);
            }
            __syncthreads();
            for (int i = threadIdx.x; i < R * S; i += blockDim.x) {
                const int r = i / S, c = i - r * S;
                cS[r * p.S4 + c] = sAS[r * p.AS4 + A + c];
            }
            if (p.d_action_in) {
                for (int i = threadIdx.x; i < R * A; i += blockDim.x) {
                    const int r = i / A, j = i - r * A;
                    if (row0 + r < p.n) p.d_action_in[(trow0 + r) * A + j] = sAS[r * p.AS4 + j];
                }
            }
            // ------------------------------------------------------------ actor (POLICY)
            if (policy && p.want_actor_dy) {
                const int ah = p.ah, last = p.an - 1;
                if (p.act_dist == DRM_DIST_TANH_NORMAL) {
                    for (int i = threadIdx.x; i < R * A; i += blockDim.x) {
                        const int r = i / A, j = i - r * A;
                        float dm = 0.f, ds = 0.f;
                        if (row0 + r < p.n) {
                            const size_t o = (trow0 + r) * A + j;
                            const float da = sAS[r * p.AS4 + j] + (p.g_action_out ? p.g_action_out[o] : 0.f);
                            const float a = p.action_out[o];
                            const float dpre = da * (1.f - a * a);
                            const float rm = p.sv.raw[(trow0 + r) * p.Aout + j];
                            const float rs = p.sv.raw[(trow0 + r) * p.Aout + A + j];
                            const float th = tanhf(rm / p.mean_scale);
                            dm = dpre * (1.f - th * th);
                            ds = dpre * noise_at(p.noise_act, nk, NZ_NOISE_ACT, NZ_NORMAL, o) * sigmoidf_(rs + p.raw_init_std);
                            p.dy.a[last][(trow0 + r) * p.Aout + j] = dm;
                            p.dy.a[last][(trow0 + r) * p.Aout + A + j] = ds;
                        }
                        sRaw[r * p.Aout4 + j] = dm;
                        sRaw[r * p.Aout4 + A + j] = ds;
                    }
                } else {
                    for (int r = threadIdx.x; r < R; r += blockDim.x) {
                        const bool ok = row0 + r < p.n;
                        if (!ok) { for (int j = 0; j < A; ++j) sRaw[r * p.Aout4 + j] = 0.f; continue; }
                        const float* lg = p.sv.raw + (trow0 + r) * p.Aout;
                        const float* go = p.g_action_out ? p.g_action_out + (trow0 + r) * A : nullptr;
                        if (p.act_dist == DRM_DIST_ONE_HOT) {
                            float mx = -INFINITY;
                            for (int j = 0; j < A; ++j) mx = fmaxf(mx, lg[j]);
                            float den = 0.f;
                            for (int j = 0; j < A; ++j) den += expf(lg[j] - mx);
                            float dot = 0.f;
                            for (int j = 0; j < A; ++j) dot += (sAS[r * p.AS4 + j] + (go ? go[j] : 0.f)) * (expf(lg[j] - mx) / den);
                            for (int j = 0; j < A; ++j) {
                                const float pj = expf(lg[j] - mx) / den;
                                const float g = pj * ((sAS[r * p.AS4 + j] + (go ? go[j] : 0.f)) - dot);
                                sRaw[r * p.Aout4 + j] = g;
                                p.dy.a[last][(trow0 + r) * p.Aout + j] = g;
                            }
                        } else {
                            const float* av = p.action_out + (trow0 + r) * A;
                            float dot = 0.f;
                            for (int j = 0; j < A; ++j) dot += (sAS[r * p.AS4 + j] + (go ? go[j] : 0.f)) * av[j];
                            for (int j = 0; j < A; ++j) {
                                const float g = av[j] * ((sAS[r * p.AS4 + j] + (go ? go[j] : 0.f)) - dot) / p.temperature;
                                sRaw[r * p.Aout4 + j] = g;
                                p.dy.a[last][(trow0 + r) * p.Aout + j] = g;
                            }
                        }
                    }
                }
                __syncthreads();
                // hidden layers, last Linear first; layer 0's input (feat) is detached (rnns.py:238)
                const float* cur = sRaw;
                int curK4 = p.Aout4, curLd = p.Aout4;
                int off = 0;
                for (int l = last; l >= 1; --l) {
                    const float* hs = p.sv.h[l - 1];
                    float* dsv = p.dy.a[l - 1];
                    float* dst = sU + off;
                    auto epi = [&](int r, int col, float v) {
                        float g = 0.f;
                        if (row0 + r < p.n) {
                            g = v * elu_grad_from_out(hs[(trow0 + r) * ah + col]);
                            dsv[(trow0 + r) * ah + col] = g;
                        }
                        dst[r * U + col] = g;
                    };
                    Seg s0{p.Wa[l], cur, curK4, curLd};
                    cta_linear<R>(s0, none, p.ld_a[l], ah, epi, ksc);
                    __syncthreads();
                    cur = dst; curK4 = ah; curLd = U;
                    off = off ? 0 : ah;
                }
            }
            __syncthreads();
        }
        // gradient w.r.t. the start state
        if (p.d_start_stoch) {
            for (int i = threadIdx.x; i < R * S; i += blockDim.x) {
                const int r = i / S, c = i - r * S;
                if (row0 + r < p.n) p.d_start_stoch[(size_t)(row0 + r) * S + c] = cS[r * p.S4 + c];
            }
        }
        if (p.d_start_deter) {
            for (int i = threadIdx.x; i < R * D; i += blockDim.x) {
                const int r = i / D, c = i - r * D;
                if (row0 + r < p.n) p.d_start_deter[(size_t)(row0 + r) * D + c] = cD[r * D + c];
            }
        }
        __syncthreads();
    }
}

int launch_rollout_bwd(const RollDev& p, cudaStream_t stream) {
    const size_t per_row = (size_t)bwd_smem_floats_per_row(p) * sizeof(float);
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
    ProfScope prof(PROF_ROLLOUT_BWD, stream);
    DRM_DISPATCH_R(Rv, {
        DRM_CUDA_OK(cudaFuncSetAttribute(rollout_bwd_kernel<R>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        rollout_bwd_kernel<R><<<grid, nthreads, smem, stream>>>(p);
    });
    DRM_CUDA_OK(cudaGetLastError());
    return DRM_OK;
}

void warm_rollout_bwd() { warm_kernel(rollout_bwd_kernel<1>); warm_kernel(rollout_bwd_kernel<2>); warm_kernel(rollout_bwd_kernel<4>); warm_kernel(rollout_bwd_kernel<8>); warm_kernel(rollout_bwd_kernel<12>); warm_kernel(rollout_bwd_kernel<16>); warm_kernel(rollout_bwd_kernel<20>); warm_kernel(rollout_bwd_kernel<24>); warm_kernel(rollout_bwd_kernel<32>); }

}  // namespace drm
