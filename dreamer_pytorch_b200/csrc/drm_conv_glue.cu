// Pointwise glue of the conv encoder / decoder (dreamer/models/observation.py:18-25,83-93: every
// Conv2d / ConvTranspose2d is followed by ReLU).  The convolutions themselves are library calls (cuDNN); around them
// torch runs, per layer, a broadcast bias add (strided, not vectorised on channels-last tensors), a ReLU, and in the
// backward pass a threshold kernel plus a bias-gradient reduction: four passes over activations of up to 307 MB each.
// Here they are two:   forward   y = act(x + bias[c])             in place on the conv output (called without bias)
//                      backward  g_pre = g * (y > 0),  dbias[c] = sum g_pre      one pass, the reduction rides along
// on channels-last fp32 tensors viewed as [rows = N*H*W, C], C a multiple of 4.  HBM-bound.
#include "drm_common.cuh"

namespace drm {

constexpr int kGlueThreads = 256;

__global__ void __launch_bounds__(kGlueThreads) bias_act_fwd_kernel(const drm_bias_act_call c) {
    const int C4 = c.C >> 2;
    const size_t total4 = (size_t)c.rows * C4;
    float4* x = reinterpret_cast<float4*>(c.x);
    const float4* b = reinterpret_cast<const float4*>(c.bias);
    const size_t step = (size_t)gridDim.x * kGlueThreads;
    const int stepq = (int)(step % C4);
    size_t i = (size_t)blockIdx.x * kGlueThreads + threadIdx.x;
    int q = (int)(i % C4);  // channel quad of element i, carried along instead of a 64-bit modulo per element
    for (; i < total4; i += step, q = (q + stepq >= C4) ? q + stepq - C4 : q + stepq) {
        float4 v = x[i];
        const float4 bb = __ldg(b + q);
        v.x += bb.x; v.y += bb.y; v.z += bb.z; v.w += bb.w;
        if (c.relu) { v.x = fmaxf(v.x, 0.f); v.y = fmaxf(v.y, 0.f); v.z = fmaxf(v.z, 0.f); v.w = fmaxf(v.w, 0.f); }
        x[i] = v;
    }
}

// thread -> channel quad q = tid % C4 (fixed for the whole kernel), row lane = tid / C4; partial bias sums stay in
// registers, meet in shared memory, one atomicAdd per channel and CTA.  Requires C4 to divide the CTA size.
__global__ void __launch_bounds__(kGlueThreads) bias_act_bwd_kernel(const drm_bias_act_call c) {
    __shared__ float4 part[kGlueThreads];
    const int C4 = c.C >> 2;
    const int q = threadIdx.x % C4, lane = threadIdx.x / C4, lanes = kGlueThreads / C4;
    const float4* g = reinterpret_cast<const float4*>(c.g);
    const float4* y = reinterpret_cast<const float4*>(c.y);
    float4* out = reinterpret_cast<float4*>(c.g_pre);
    float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
    for (long long r = (long long)blockIdx.x * lanes + lane; r < c.rows; r += (long long)gridDim.x * lanes) {
        const size_t i = (size_t)r * C4 + q;
        float4 v = __ldg(g + i);
        if (c.relu) {
            const float4 yy = __ldg(y + i);
            v.x = yy.x > 0.f ? v.x : 0.f; v.y = yy.y > 0.f ? v.y : 0.f;
            v.z = yy.z > 0.f ? v.z : 0.f; v.w = yy.w > 0.f ? v.w : 0.f;
            out[i] = v;
        }
        acc.x += v.x; acc.y += v.y; acc.z += v.z; acc.w += v.w;
    }
    part[threadIdx.x] = acc;
    __syncthreads();
    if (lane == 0 && c.dbias) {
        float4 s = acc;
        for (int l = 1; l < lanes; ++l) {
            const float4 p = part[l * C4 + q];
            s.x += p.x; s.y += p.y; s.z += p.z; s.w += p.w;
        }
        atomicAdd(c.dbias + 4 * q + 0, s.x);
        atomicAdd(c.dbias + 4 * q + 1, s.y);
        atomicAdd(c.dbias + 4 * q + 2, s.z);
        atomicAdd(c.dbias + 4 * q + 3, s.w);
    }
}

void warm_conv_glue() { warm_kernel(bias_act_fwd_kernel); warm_kernel(bias_act_bwd_kernel); }

}  // namespace drm

using namespace drm;

static int check_call(const drm_bias_act_call* c) {
    DRM_REQUIRE(c != nullptr, DRM_E_ARG, "null call");
    DRM_REQUIRE(c->rows >= 0 && c->C >= 4 && c->C % 4 == 0 && c->C <= 4 * kGlueThreads && kGlueThreads % (c->C / 4) == 0,
                DRM_E_SHAPE, "channels must be a multiple of 4 whose quarter divides 256");
    return DRM_OK;
}

static int glue_grid(long long work_items) {
    long long blocks = (work_items + kGlueThreads - 1) / kGlueThreads;
    const long long cap = (long long)num_sms() * 8;
    if (blocks > cap) blocks = cap;
    return blocks < 1 ? 1 : (int)blocks;
}

extern "C" {

int drm_bias_act_fwd(const drm_bias_act_call* c, void* stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    int rc = check_call(c); if (rc) return rc;
    DRM_REQUIRE(c->x && c->bias, DRM_E_ARG, "null tensor");
    DRM_REQUIRE(((size_t)c->x & 15) == 0 && ((size_t)c->bias & 15) == 0, DRM_E_ALIGN, "16-byte alignment");
    rc = check_device(); if (rc) return rc;
    if (c->rows == 0) return DRM_OK;
    ProfScope prof(PROF_LOSS, stream);
    bias_act_fwd_kernel<<<glue_grid(c->rows * (c->C / 4)), kGlueThreads, 0, stream>>>(*c);
    DRM_CUDA_OK(cudaGetLastError());
    return DRM_OK;
}

int drm_bias_act_bwd(const drm_bias_act_call* c, void* stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    int rc = check_call(c); if (rc) return rc;
    DRM_REQUIRE(c->g && (!c->relu || (c->y && c->g_pre)), DRM_E_ARG, "null tensor");
    DRM_REQUIRE(((size_t)c->g & 15) == 0 && ((size_t)c->y & 15) == 0 && ((size_t)c->g_pre & 15) == 0, DRM_E_ALIGN,
                "16-byte alignment");
    rc = check_device(); if (rc) return rc;
    if (c->dbias) DRM_CUDA_OK(cudaMemsetAsync(c->dbias, 0, sizeof(float) * c->C, stream));
    if (c->rows == 0) return DRM_OK;
    ProfScope prof(PROF_LOSS, stream);
    const int lanes = kGlueThreads / (c->C / 4);
    bias_act_bwd_kernel<<<glue_grid((c->rows + lanes - 1) / lanes * kGlueThreads / 4), kGlueThreads, 0, stream>>>(*c);
    DRM_CUDA_OK(cudaGetLastError());
    return DRM_OK;
}

}  // extern "C"
