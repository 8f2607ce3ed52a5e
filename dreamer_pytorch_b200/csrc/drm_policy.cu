// Environment-time policy head: SampleDist.mode() of the tanh-Normal action distribution
// (dreamer/models/distribution.py:24-35 on action.py:54-61) as ONE launch instead of ~15 torch.distributions kernels:
// `samples` reparameterised draws per row, their log-probabilities under the transformed distribution, arg-max, gather.
//   mean = mean_scale * tanh(raw_m / mean_scale);  std = softplus(raw_s + raw_init_std) + min_std
//   z = mean + std * eps;  y = tanh(z)                                  (TransformedDistribution.rsample)
//   x = atanh(y)                                                        (TanhTransform._inverse, torch/distributions/transforms.py)
//   logp = sum_j [ Normal(mean, std).log_prob(x) - 2 (ln 2 - x - softplus(-2 x)) ]
// One warp per row; lanes stride over the samples; the winner (first arg-max) is found with warp shuffles.
#include "drm_common.cuh"

namespace drm {

constexpr int kPolWarps = 4;

__global__ void __launch_bounds__(kPolWarps * 32) tanh_normal_mode_kernel(const drm_policy_mode_call c) {
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int row = blockIdx.x * kPolWarps + warp;
    if (row >= c.n) return;
    const int A = c.A;
    const float* raw = c.raw + (size_t)row * 2 * A;
    // torch.argmax semantics: NaN counts as the maximum, the first occurrence wins.  (A draw whose tanh saturates to +-1 in
    // fp32 has atanh = +-inf and a NaN log-probability in the reference -- TanhTransform._inverse does not clamp -- so with
    // the wide initial std the reference's mode() returns its first saturated draw; reproduced here, not "fixed".)
    float best = -INFINITY;
    int best_s = 0x7fffffff;
    bool best_nan = false;
    auto better = [](float v, int s, bool vn, float b, int bs, bool bn) {
        if (vn != bn) return vn;
        if (vn) return s < bs;
        return v > b || (v == b && s < bs);
    };
    for (int s = lane; s < c.samples; s += 32) {
        const float* e = c.noise + ((size_t)s * c.n + row) * A;
        float lp = 0.f;
        for (int j = 0; j < A; ++j) {
            const float mean = c.mean_scale * tanhf(raw[j] / c.mean_scale);
            const float std = softplusf_(raw[A + j] + c.raw_init_std) + c.min_std;
            const float y = tanhf(mean + std * e[j]);
            const float x = atanhf(y);                                     // TanhTransform._inverse
            const float d = x - mean;
            const float base = -(d * d) / (2.f * std * std) - logf(std) - 0.9189385332046727f;  // Normal.log_prob
            const float ladj = 2.f * (0.6931471805599453f - x - softplusf_(-2.f * x));          // log_abs_det_jacobian
            lp += base - ladj;
        }
        if (c.logp) c.logp[(size_t)s * c.n + row] = lp;
        const bool ln_ = lp != lp;
        if (best_s == 0x7fffffff || better(lp, s, ln_, best, best_s, best_nan)) { best = lp; best_s = s; best_nan = ln_; }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        const float ob = __shfl_xor_sync(0xffffffffu, best, o);
        const int os = __shfl_xor_sync(0xffffffffu, best_s, o);
        const bool on = __shfl_xor_sync(0xffffffffu, (int)best_nan, o) != 0;
        if (os != 0x7fffffff && (best_s == 0x7fffffff || better(ob, os, on, best, best_s, best_nan))) {
            best = ob; best_s = os; best_nan = on;
        }
    }
    if (best_s == 0x7fffffff) best_s = 0;
    for (int j = lane; j < A; j += 32) {
        const float mean = c.mean_scale * tanhf(raw[j] / c.mean_scale);
        const float std = softplusf_(raw[A + j] + c.raw_init_std) + c.min_std;
        c.action[(size_t)row * A + j] = tanhf(mean + std * c.noise[((size_t)best_s * c.n + row) * A + j]);
    }
    if (lane == 0 && c.index) c.index[row] = best_s;
}

void warm_policy() { warm_kernel(tanh_normal_mode_kernel); }

}  // namespace drm

extern "C" int drm_policy_mode(const drm_policy_mode_call* c, void* stream_) {
    using namespace drm;
    cudaStream_t stream = (cudaStream_t)stream_;
    DRM_REQUIRE(c != nullptr, DRM_E_ARG, "null call");
    DRM_REQUIRE(c->n >= 1 && c->A >= 1 && c->samples >= 1, DRM_E_SHAPE, "need n, A, samples >= 1");
    DRM_REQUIRE(c->raw && c->noise && c->action, DRM_E_ARG, "null pointer");
    int rc = check_device(); if (rc) return rc;
    ProfScope prof(PROF_LOSS, stream);
    tanh_normal_mode_kernel<<<(c->n + kPolWarps - 1) / kPolWarps, kPolWarps * 32, 0, stream>>>(*c);
    DRM_CUDA_OK(cudaGetLastError());
    return DRM_OK;
}
