// Loss-section kernels: lambda-returns + actor/value losses with their analytic gradients
// (dreamer_algo.py:314-337,398-422) and KL / free-nats / entropies (dreamer_algo.py:270-273,342-343).
#include <cuda_bf16.h>
#include "drm_common.cuh"

namespace drm {

__device__ __forceinline__ float block_sum(float v, float* red) {
    v = warp_sum(v);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    __syncthreads();
    if (lane == 0) red[warp] = v;
    __syncthreads();
    float s = 0.f;
    if (warp == 0) {
        s = lane < (blockDim.x >> 5) ? red[lane] : 0.f;
        s = warp_sum(s);
    }
    return s;  // valid in warp 0
}

// One thread per imagined trajectory (column n); arrays are [H,N] so accesses are coalesced.
__global__ void __launch_bounds__(256) returns_loss_kernel(const drm_returns_loss_call c) {
    __shared__ float red[8];
    const int N = c.N, H = c.H, L = c.H - 1;
    const int n = blockIdx.x * blockDim.x + threadIdx.x;
    const float lam = c.lambda_;
    const float cmean = -1.f / ((float)L * (float)N);
    float sum_actor = 0.f, sum_value = 0.f;
    if (n < N) {
        auto dsc = [&](int h) -> float { return c.pcont ? sigmoidf_(c.pcont[(size_t)h * N + n]) : c.discount; };
        // pass 1 (reverse): lambda-returns, compute_return (dreamer_algo.py:398-422)
        float acc = c.value[(size_t)(H - 1) * N + n];  // bootstrap
        for (int h = L - 1; h >= 0; --h) {
            const float d = dsc(h);
            const float vnext = c.value[(size_t)(h + 1) * N + n];
            const float target = c.reward[(size_t)h * N + n] + d * vnext * (1.f - lam);
            acc = target + d * lam * acc;
            c.returns[(size_t)h * N + n] = acc;
        }
        // pass 2 (forward): cumulative discount, losses, adjoints of the return recursion
        float disc = 1.f, gR = 0.f, dprev = 0.f;
        c.d_value[n] = 0.f;
        for (int h = 0; h < L; ++h) {
            const float d = dsc(h);
            if (h > 0) disc *= d;  // cumprod([1, d_1, ..., d_{H-2}])  (dreamer_algo.py:322-323)
            const float Rh = c.returns[(size_t)h * N + n];
            const float v = c.value[(size_t)h * N + n];
            c.disc[(size_t)h * N + n] = disc;
            sum_actor += disc * Rh;
            const float diff = Rh - v;
            sum_value += disc * (-0.5f * diff * diff - 0.9189385332046727f);  // Normal(v,1).log_prob(R)
            c.d_value_pred[(size_t)h * N + n] = cmean * disc * diff;
            gR = cmean * disc + gR * dprev * lam;  // adjoint of returns[h]
            dprev = d;
            const float vnext = c.value[(size_t)(h + 1) * N + n];
            const float Rnext = (h + 1 < L) ? c.returns[(size_t)(h + 1) * N + n] : vnext;
            c.d_reward[(size_t)h * N + n] = gR;
            float dv = gR * d * (1.f - lam);
            if (h == L - 1) dv += gR * d * lam;  // bootstrap = value[H-1]
            c.d_value[(size_t)(h + 1) * N + n] = dv;
            if (c.d_pcont) c.d_pcont[(size_t)h * N + n] = gR * (lam * Rnext + (1.f - lam) * vnext);  // d loss / d d_h (returns part)
        }
        c.d_reward[(size_t)L * N + n] = 0.f;
        c.d_value_pred[(size_t)L * N + n] = 0.f;
        if (c.d_pcont) {
            c.d_pcont[(size_t)L * N + n] = 0.f;
            // pass 3 (reverse): adjoint of the cumprod, then through the sigmoid
            float carry = 0.f;
            for (int h = L - 1; h >= 0; --h) {
                const float d = dsc(h);
                float gd = c.d_pcont[(size_t)h * N + n];
                if (h >= 1) {
                    const float gdisc = cmean * c.returns[(size_t)h * N + n] + carry;
                    gd += gdisc * c.disc[(size_t)(h - 1) * N + n];
                    carry = gdisc * d;
                }
                c.d_pcont[(size_t)h * N + n] = gd * d * (1.f - d);
            }
        }
    }
    const float sa = block_sum(sum_actor, red);
    const float sv = block_sum(sum_value, red);
    if (threadIdx.x == 0) {
        atomicAdd(c.losses + 0, cmean * sa);
        atomicAdd(c.losses + 1, cmean * sv);
    }
}

// Same computation with the trajectory held in registers (H <= MAXH): every global load of a column is issued up front
// instead of one dependent load per step of the recursions (the generic kernel above spends ~H load latencies per pass).
template <int MAXH>
__global__ void __launch_bounds__(128) returns_loss_reg_kernel(const drm_returns_loss_call c) {
    __shared__ float red[8];
    const int N = c.N, H = c.H, L = c.H - 1;
    const int n = blockIdx.x * blockDim.x + threadIdx.x;
    const float lam = c.lambda_;
    const float cmean = -1.f / ((float)L * (float)N);
    float sum_actor = 0.f, sum_value = 0.f;
    if (n < N) {
        float V[MAXH], Rw[MAXH], D[MAXH], Ret[MAXH], Disc[MAXH], gP[MAXH];
#pragma unroll
        for (int h = 0; h < MAXH; ++h) {
            V[h] = h < H ? c.value[(size_t)h * N + n] : 0.f;
            Rw[h] = h < L ? c.reward[(size_t)h * N + n] : 0.f;
            D[h] = c.pcont ? (h < L ? sigmoidf_(c.pcont[(size_t)h * N + n]) : 0.f) : c.discount;
            Ret[h] = 0.f; Disc[h] = 0.f; gP[h] = 0.f;
        }
        // pass 1 (reverse): lambda-returns
        float acc = 0.f;
#pragma unroll
        for (int h = MAXH - 1; h >= 0; --h) {
            if (h == H - 1) acc = V[h];  // bootstrap
            if (h < L) {
                const float vnext = h + 1 < MAXH ? V[h + 1 < MAXH ? h + 1 : h] : 0.f;
                const float target = Rw[h] + D[h] * vnext * (1.f - lam);
                acc = target + D[h] * lam * acc;
                Ret[h] = acc;
                c.returns[(size_t)h * N + n] = acc;
            }
        }
        // pass 2 (forward): cumulative discount, losses, adjoints of the return recursion
        float disc = 1.f, gR = 0.f, dprev = 0.f;
        c.d_value[n] = 0.f;
#pragma unroll
        for (int h = 0; h < MAXH; ++h) {
            if (h < L) {
                const float d = D[h];
                if (h > 0) disc *= d;
                const float Rh = Ret[h], v = V[h];
                Disc[h] = disc;
                c.disc[(size_t)h * N + n] = disc;
                sum_actor += disc * Rh;
                const float diff = Rh - v;
                sum_value += disc * (-0.5f * diff * diff - 0.9189385332046727f);
                c.d_value_pred[(size_t)h * N + n] = cmean * disc * diff;
                gR = cmean * disc + gR * dprev * lam;
                dprev = d;
                const float vnext = V[h + 1 < MAXH ? h + 1 : h];
                const float Rnext = (h + 1 < L) ? Ret[h + 1 < MAXH ? h + 1 : h] : vnext;
                c.d_reward[(size_t)h * N + n] = gR;
                float dv = gR * d * (1.f - lam);
                if (h == L - 1) dv += gR * d * lam;
                c.d_value[(size_t)(h + 1) * N + n] = dv;
                gP[h] = gR * (lam * Rnext + (1.f - lam) * vnext);
            }
        }
        c.d_reward[(size_t)L * N + n] = 0.f;
        c.d_value_pred[(size_t)L * N + n] = 0.f;
        if (c.d_pcont) {
            c.d_pcont[(size_t)L * N + n] = 0.f;
            float carry = 0.f;
#pragma unroll
            for (int h = MAXH - 1; h >= 0; --h) {
                if (h < L) {
                    const float d = D[h];
                    float gd = gP[h];
                    if (h >= 1) {
                        const float gdisc = cmean * Ret[h] + carry;
                        gd += gdisc * Disc[h >= 1 ? h - 1 : 0];
                        carry = gdisc * d;
                    }
                    c.d_pcont[(size_t)h * N + n] = gd * d * (1.f - d);
                }
            }
        }
    }
    const float sa = block_sum(sum_actor, red);
    const float sv = block_sum(sum_value, red);
    if (threadIdx.x == 0) {
        atomicAdd(c.losses + 0, cmean * sa);
        atomicAdd(c.losses + 1, cmean * sv);
    }
}

__global__ void __launch_bounds__(256) kl_fwd_kernel(const drm_kl_call c) {
    __shared__ float red[8];
    const size_t total = (size_t)c.M * c.S;
    float kl = 0.f, ep = 0.f, eq = 0.f;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
        const float mq = c.post_mean[i], sq = c.post_std[i], mp = c.prior_mean[i], sp = c.prior_std[i];
        const float ratio = sq / sp;
        const float vr = ratio * ratio;
        const float t1 = (mq - mp) / sp;
        kl += 0.5f * (vr + t1 * t1 - 1.f - logf(vr));
        ep += 1.4189385332046727f + logf(sp);  // 0.5 + 0.5 ln(2 pi) + ln(sigma)
        eq += 1.4189385332046727f + logf(sq);
    }
    const float inv = 1.f / (float)c.M;
    const float a = block_sum(kl, red), b = block_sum(ep, red), d = block_sum(eq, red);
    if (threadIdx.x == 0) {
        atomicAdd(c.out + 1, a * inv);
        atomicAdd(c.out + 2, b * inv);
        atomicAdd(c.out + 3, d * inv);
    }
}
__global__ void kl_finalize_kernel(float* out, float free_nats) { out[0] = fmaxf(out[1], free_nats); }

__global__ void __launch_bounds__(256) kl_bwd_kernel(const drm_kl_call c) {
    const size_t total = (size_t)c.M * c.S;
    const bool active = c.out[1] > c.free_nats;  // torch.max(div, free_nats): gradient only above the clamp
    const float inv = active ? 1.f / (float)c.M : 0.f;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
        const float mq = c.post_mean[i], sq = c.post_std[i], mp = c.prior_mean[i], sp = c.prior_std[i];
        const float isp2 = 1.f / (sp * sp);
        const float dm = (mq - mp) * isp2;
        if (c.d_post_mean) c.d_post_mean[i] = inv * dm;
        if (c.d_prior_mean) c.d_prior_mean[i] = -inv * dm;
        if (c.d_post_std) c.d_post_std[i] = inv * (sq * isp2 - 1.f / sq);
        if (c.d_prior_std) c.d_prior_std[i] = inv * (1.f / sp - (sq * sq + (mq - mp) * (mq - mp)) * isp2 / sp);
    }
}


// ------------------------------------------------------------------------------------------------
// Image log-likelihood of the observation decoder: Independent(Normal(mean, 1), 3).log_prob(obs)
// (dreamer/models/observation.py:117, dreamer_algo.py:264) as ONE reduction per image instead of a chain of ATen
// pointwise kernels over [T*B, 3*64*64]:  logp[row] = -0.5 * sum_i (obs - mean)^2 - D * 0.5 ln(2 pi).
// HBM-bound: 2 reads per element forward, 2 reads + 1 write backward.  `mean` is fp32 or bf16 (autocast decoder).
// ------------------------------------------------------------------------------------------------
template <bool kBf16>
__device__ __forceinline__ float4 ld_mean4(const void* mean, size_t i4) {
    if (kBf16) {
        const uint2 q = __ldg(reinterpret_cast<const uint2*>(mean) + i4);
        return make_float4(__uint_as_float(q.x << 16), __uint_as_float(q.x & 0xFFFF0000u), __uint_as_float(q.y << 16),
                           __uint_as_float(q.y & 0xFFFF0000u));
    }
    return __ldg(reinterpret_cast<const float4*>(mean) + i4);
}

template <bool kBf16>
__global__ void __launch_bounds__(256) image_logprob_fwd_kernel(const drm_image_logprob_call c) {
    __shared__ float red[8];
    const size_t row = blockIdx.x;
    const int D4 = c.D >> 2;
    const size_t base4 = row * (size_t)D4;
    float acc = 0.f;
    for (int i = threadIdx.x; i < D4; i += 256) {
        const float4 m = ld_mean4<kBf16>(c.mean, base4 + i);
        const float4 x = __ldg(reinterpret_cast<const float4*>(c.obs) + base4 + i);
        const float a = x.x - m.x, b = x.y - m.y, d = x.z - m.z, e = x.w - m.w;
        acc += (a * a + b * b) + (d * d + e * e);
    }
    const float s = block_sum(acc, red);
    if (threadIdx.x == 0) c.logp[row] = -0.5f * s - (float)c.D * 0.9189385332046727f;
}

template <bool kBf16>
__global__ void __launch_bounds__(256) image_logprob_bwd_kernel(const drm_image_logprob_call c) {
    const size_t row = blockIdx.x;
    const int D4 = c.D >> 2;
    const size_t base4 = row * (size_t)D4;
    const float g = __ldg(c.g + row);
    for (int i = threadIdx.x; i < D4; i += 256) {
        const float4 m = ld_mean4<kBf16>(c.mean, base4 + i);
        const float4 x = __ldg(reinterpret_cast<const float4*>(c.obs) + base4 + i);
        const float4 d = make_float4(g * (x.x - m.x), g * (x.y - m.y), g * (x.z - m.z), g * (x.w - m.w));
        if (kBf16) {
            __nv_bfloat162 lo = __floats2bfloat162_rn(d.x, d.y), hi = __floats2bfloat162_rn(d.z, d.w);
            uint2 q;
            q.x = *reinterpret_cast<uint32_t*>(&lo); q.y = *reinterpret_cast<uint32_t*>(&hi);
            reinterpret_cast<uint2*>(c.dmean)[base4 + i] = q;
        } else {
            reinterpret_cast<float4*>(c.dmean)[base4 + i] = d;
        }
    }
}

void warm_loss() { warm_kernel(returns_loss_kernel); warm_kernel(returns_loss_reg_kernel<16>); warm_kernel(returns_loss_reg_kernel<32>); warm_kernel(kl_fwd_kernel); warm_kernel(kl_finalize_kernel); warm_kernel(kl_bwd_kernel);
    warm_kernel(image_logprob_fwd_kernel<false>); warm_kernel(image_logprob_fwd_kernel<true>);
    warm_kernel(image_logprob_bwd_kernel<false>); warm_kernel(image_logprob_bwd_kernel<true>); }

}  // namespace drm

using namespace drm;

extern "C" {

int drm_returns_loss(const drm_returns_loss_call* c, void* stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    DRM_REQUIRE(c != nullptr, DRM_E_ARG, "null call");
    DRM_REQUIRE(c->H >= 2 && c->N >= 1, DRM_E_SHAPE, "need H >= 2 and N >= 1");
    DRM_REQUIRE(c->reward && c->value && c->returns && c->disc && c->losses && c->d_reward && c->d_value &&
                    c->d_value_pred, DRM_E_ARG, "null pointer");
    DRM_REQUIRE((c->pcont == nullptr) == (c->d_pcont == nullptr), DRM_E_ARG, "pcont and d_pcont must come together");
    int rc = check_device(); if (rc) return rc;
    ProfScope prof(PROF_LOSS, stream);
    if (c->H <= 16) returns_loss_reg_kernel<16><<<(c->N + 127) / 128, 128, 0, stream>>>(*c);
    else if (c->H <= 32) returns_loss_reg_kernel<32><<<(c->N + 127) / 128, 128, 0, stream>>>(*c);
    else returns_loss_kernel<<<(c->N + 255) / 256, 256, 0, stream>>>(*c);
    DRM_CUDA_OK(cudaGetLastError());
    return DRM_OK;
}

int drm_kl_fwd(const drm_kl_call* c, void* stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    DRM_REQUIRE(c != nullptr, DRM_E_ARG, "null call");
    DRM_REQUIRE(c->M >= 1 && c->S >= 1, DRM_E_SHAPE, "bad dims");
    DRM_REQUIRE(c->post_mean && c->post_std && c->prior_mean && c->prior_std && c->out, DRM_E_ARG, "null pointer");
    int rc = check_device(); if (rc) return rc;
    const size_t total = (size_t)c->M * c->S;
    int grid = (int)((total + 255) / 256);
    if (grid > 2 * num_sms()) grid = 2 * num_sms();
    ProfScope prof(PROF_LOSS, stream);
    kl_fwd_kernel<<<grid, 256, 0, stream>>>(*c);
    kl_finalize_kernel<<<1, 1, 0, stream>>>(c->out, c->free_nats);
    DRM_CUDA_OK(cudaGetLastError());
    return DRM_OK;
}

int drm_kl_bwd(const drm_kl_call* c, void* stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    DRM_REQUIRE(c != nullptr, DRM_E_ARG, "null call");
    DRM_REQUIRE(c->M >= 1 && c->S >= 1, DRM_E_SHAPE, "bad dims");
    DRM_REQUIRE(c->post_mean && c->post_std && c->prior_mean && c->prior_std && c->out, DRM_E_ARG, "null pointer");
    int rc = check_device(); if (rc) return rc;
    const size_t total = (size_t)c->M * c->S;
    int grid = (int)((total + 255) / 256);
    if (grid > 4 * num_sms()) grid = 4 * num_sms();
    ProfScope prof(PROF_LOSS, stream);
    kl_bwd_kernel<<<grid, 256, 0, stream>>>(*c);
    DRM_CUDA_OK(cudaGetLastError());
    return DRM_OK;
}

static int image_logprob_check(const drm_image_logprob_call* c, bool bwd) {
    DRM_REQUIRE(c != nullptr, DRM_E_ARG, "null call");
    DRM_REQUIRE(c->rows >= 1 && c->rows < (1ll << 31) && c->D >= 4 && (c->D & 3) == 0, DRM_E_SHAPE, "need rows >= 1 and D % 4 == 0");
    DRM_REQUIRE(c->mean && c->obs, DRM_E_ARG, "null pointer");
    DRM_REQUIRE(((uintptr_t)c->mean & 15) == 0 && ((uintptr_t)c->obs & 15) == 0, DRM_E_ALIGN, "mean / obs must be 16-byte aligned");
    if (bwd) DRM_REQUIRE(c->g && c->dmean && ((uintptr_t)c->dmean & 15) == 0, DRM_E_ARG, "backward needs g and an aligned dmean");
    else DRM_REQUIRE(c->logp != nullptr, DRM_E_ARG, "null logp");
    return check_device();
}

int drm_image_logprob_fwd(const drm_image_logprob_call* c, void* stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    int rc = image_logprob_check(c, false); if (rc) return rc;
    ProfScope prof(PROF_LOSS, stream);
    if (c->mean_bf16) image_logprob_fwd_kernel<true><<<(unsigned)c->rows, 256, 0, stream>>>(*c);
    else image_logprob_fwd_kernel<false><<<(unsigned)c->rows, 256, 0, stream>>>(*c);
    DRM_CUDA_OK(cudaGetLastError());
    return DRM_OK;
}

int drm_image_logprob_bwd(const drm_image_logprob_call* c, void* stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    int rc = image_logprob_check(c, true); if (rc) return rc;
    ProfScope prof(PROF_LOSS, stream);
    if (c->mean_bf16) image_logprob_bwd_kernel<true><<<(unsigned)c->rows, 256, 0, stream>>>(*c);
    else image_logprob_bwd_kernel<false><<<(unsigned)c->rows, 256, 0, stream>>>(*c);
    DRM_CUDA_OK(cudaGetLastError());
    return DRM_OK;
}

}  // extern "C"
