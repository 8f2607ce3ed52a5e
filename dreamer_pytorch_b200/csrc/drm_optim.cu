// Optimizer tail of one parameter group: torch.nn.utils.clip_grad_norm_ followed by torch.optim.Adam.step
// (dreamer_algo.py:182-194) as two launches.  HBM-bound, tiny (an actor or value group is ~0.13 M parameters): what
// matters is that it is 2 launches instead of ~16 and needs no host synchronisation.
#include "drm_common.cuh"

namespace drm {

constexpr int kOptThreads = 256;
// Work decomposition: a CTA owns one chunk of kOptChunk consecutive elements of one tensor; the grid is the total number of
// chunks of the group (a fixed 16 CTAs per tensor left the 0.2-0.5 M element tensors of the world-model group -- 5.3 M
// parameters in 48 tensors -- latency-bound on 16 CTAs each: 0.8 ms per step at 2 % of the HBM rate).
constexpr int kOptChunk = 4096;

struct OptChunk {
    int t;               // tensor
    long long lo, hi;    // element range
};
__device__ __forceinline__ OptChunk opt_chunk(const drm_clip_adam_call& c) {
    long long b = blockIdx.x;
    for (int t = 0; t < c.n_tensors; ++t) {
        const long long n = (c.numel[t] + kOptChunk - 1) / kOptChunk;
        if (b < n) {
            const long long lo = b * kOptChunk;
            return OptChunk{t, lo, lo + kOptChunk < c.numel[t] ? lo + kOptChunk : c.numel[t]};
        }
        b -= n;
    }
    return OptChunk{0, 0, 0};
}

// sum of squares of every gradient element of the group -> acc[0] (double)
__global__ void __launch_bounds__(kOptThreads) grad_sumsq_kernel(const drm_clip_adam_call c, double* acc) {
    const OptChunk ch = opt_chunk(c);
    const int t = ch.t;
    if (c.step_dev && blockIdx.x == 0 && threadIdx.x == 0) atomicAdd(c.step_dev, 1);  // read by the update kernel
    const float* g = c.grad[t];
    float s = 0.f;
    for (long long i = ch.lo + threadIdx.x; i < ch.hi; i += kOptThreads) {
        const float v = g[i];
        s = fmaf(v, v, s);
    }
    if (c.mult[t] > 1) s *= (float)c.mult[t];  // a gradient listed k times counts k times in clip_grad_norm_
    s = warp_sum(s);
    __shared__ float part[kOptThreads / 32];
    if ((threadIdx.x & 31) == 0) part[threadIdx.x >> 5] = s;
    __syncthreads();
    if (threadIdx.x < 32) {
        float v = threadIdx.x < kOptThreads / 32 ? part[threadIdx.x] : 0.f;
        v = warp_sum(v);
        if (threadIdx.x == 0 && v != 0.f) atomicAdd(acc, (double)v);
    }
}

__global__ void __launch_bounds__(kOptThreads) clip_adam_kernel(const drm_clip_adam_call c, const double* acc) {
    const float total = (float)sqrt(*acc);
    if (blockIdx.x == 0 && threadIdx.x == 0 && c.norm_out) *c.norm_out = total;
    float coef = 1.f;
    if (c.max_norm > 0.f) coef = fminf(c.max_norm / (total + 1e-6f), 1.f);  // clip_grad.py: clamp(max_norm / (norm + 1e-6), max=1)
    const OptChunk ch = opt_chunk(c);
    const int t = ch.t;
    float* p = c.param[t];
    float* g = c.grad[t];
    float* m = c.exp_avg[t];
    float* v = c.exp_avg_sq[t];
    const float w1 = c.one_minus_beta1, w2 = c.one_minus_beta2;
    float step_size = c.step_size, bias2_sqrt = c.bias2_sqrt;
    if (c.step_dev) {  // bias corrections from the device-side step counter (adam.py: 1 - beta ** step), in double
        const double st = (double)*c.step_dev;
        step_size = (float)((double)c.lr / (1.0 - pow(c.beta1_d, st)));
        bias2_sqrt = (float)sqrt(1.0 - pow(c.beta2_d, st));
    }
    const int k = c.mult[t];
    if (k > 1) {
        // parameter listed k times (see drm_clip_adam_call.mult): gradient scaled k times, k sequential Adam updates
        const double upd = (double)(c.step_dev ? *c.step_dev : c.update);
        // the bias corrections of the k sub-steps depend on the step only: once per thread, not per element
        constexpr int kMaxHoist = 4;
        float ssr[kMaxHoist], b2r[kMaxHoist];
#pragma unroll
        for (int r = 1; r <= kMaxHoist; ++r) {
            const double st = (upd - 1.0) * k + (r <= k ? r : k);
            ssr[r - 1] = (float)((double)c.lr / (1.0 - pow(c.beta1_d, st)));
            b2r[r - 1] = (float)sqrt(1.0 - pow(c.beta2_d, st));
        }
        for (long long i = ch.lo + threadIdx.x; i < ch.hi; i += kOptThreads) {
            float gi = g[i];
            for (int r = 0; r < k; ++r) gi *= coef;
            float mi = m[i], vi = v[i], pi = p[i];
#pragma unroll
            for (int r = 1; r <= kMaxHoist; ++r) {  // unrolled: ssr / b2r stay in registers
                if (r <= k) {
                    mi = mi + w1 * (gi - mi);
                    vi = vi * c.beta2 + w2 * gi * gi;
                    pi = pi - ssr[r - 1] * (mi / (sqrtf(vi) / b2r[r - 1] + c.eps));
                }
            }
            for (int r = kMaxHoist + 1; r <= k; ++r) {  // a parameter listed more than four times: per element
                const double st = (upd - 1.0) * k + r;
                const float ss = (float)((double)c.lr / (1.0 - pow(c.beta1_d, st)));
                const float b2 = (float)sqrt(1.0 - pow(c.beta2_d, st));
                mi = mi + w1 * (gi - mi);
                vi = vi * c.beta2 + w2 * gi * gi;
                pi = pi - ss * (mi / (sqrtf(vi) / b2 + c.eps));
            }
            g[i] = gi; m[i] = mi; v[i] = vi; p[i] = pi;
        }
        return;
    }
    for (long long i = ch.lo + threadIdx.x; i < ch.hi; i += kOptThreads) {
        const float gi = g[i] * coef;
        const float mi = m[i] + w1 * (gi - m[i]);         // exp_avg.lerp_(grad, 1 - beta1)
        const float vi = v[i] * c.beta2 + w2 * gi * gi;   // exp_avg_sq.mul_(beta2).addcmul_(grad, grad, value=1 - beta2)
        const float denom = sqrtf(vi) / bias2_sqrt + c.eps;
        g[i] = gi;
        m[i] = mi;
        v[i] = vi;
        p[i] = p[i] - step_size * (mi / denom);           // param.addcdiv_(exp_avg, denom, value=-step_size)
    }
}

void warm_optim() { warm_kernel(grad_sumsq_kernel); warm_kernel(clip_adam_kernel); }

}  // namespace drm

extern "C" int drm_clip_adam(const drm_clip_adam_call* c, void* stream_) {
    using namespace drm;
    cudaStream_t stream = (cudaStream_t)stream_;
    DRM_REQUIRE(c != nullptr, DRM_E_ARG, "null call");
    DRM_REQUIRE(c->n_tensors >= 1 && c->n_tensors <= DRM_MAX_OPT_TENSORS, DRM_E_SHAPE, "1..64 tensors per call");
    DRM_REQUIRE(c->workspace && c->workspace_bytes >= 16, DRM_E_WORKSPACE, "workspace too small");
    for (int t = 0; t < c->n_tensors; ++t)
        DRM_REQUIRE(c->param[t] && c->grad[t] && c->exp_avg[t] && c->exp_avg_sq[t] && c->numel[t] >= 0, DRM_E_ARG,
                    "null tensor pointer");
    int rc = check_device(); if (rc) return rc;
    double* acc = reinterpret_cast<double*>(rup_sz((size_t)c->workspace, 8));
    DRM_CUDA_OK(cudaMemsetAsync(acc, 0, sizeof(double), stream));
    long long chunks = 0;
    for (int t = 0; t < c->n_tensors; ++t) chunks += (c->numel[t] + kOptChunk - 1) / kOptChunk;
    if (chunks == 0) chunks = 1;
    DRM_REQUIRE(chunks < (1ll << 31), DRM_E_SHAPE, "parameter group too large for one launch");
    const unsigned grid = (unsigned)chunks;
    {
        ProfScope prof(PROF_LOSS, stream);
        grad_sumsq_kernel<<<grid, kOptThreads, 0, stream>>>(*c, acc);
        DRM_CUDA_OK(cudaGetLastError());
    }
    {
        ProfScope prof(PROF_LOSS, stream);
        clip_adam_kernel<<<grid, kOptThreads, 0, stream>>>(*c, acc);
        DRM_CUDA_OK(cudaGetLastError());
    }
    return DRM_OK;
}
