// Error plumbing, device check, weight packing kernel and ABI self-description.
#include <stdarg.h>
#include <string.h>
#include "drm_common.cuh"

namespace drm {

static thread_local char g_err[512] = "";

void set_error(const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}

int check_device() {
    int dev = 0;
    DRM_CUDA_OK(cudaGetDevice(&dev));
    static int cached_dev = -1, cached_major = 0;
    if (cached_dev != dev) {
        int major = 0;
        DRM_CUDA_OK(cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, dev));
        cached_major = major;
        cached_dev = dev;
    }
    if (cached_major != 10) {
        set_error("libdreamer_b200 is built for sm_100a only; device %d has compute capability major %d", dev,
                  cached_major);
        return DRM_E_ARCH;
    }
    return DRM_OK;
}

__global__ void pack_kernel(PackJobs jobs) {
    const PackJob j = jobs.j[blockIdx.y];
    const int total = j.dst_rows * j.dst_ld;
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
        const int dr = i / j.dst_ld, dc = i - dr * j.dst_ld;
        // transpose: dst row = in-feature (col of src), dst col = out-feature (row of src)
        const int sr = j.transpose ? dc : dr;
        const int sc = j.transpose ? dr : dc;
        float v = 0.f;
        if (sr < j.nr && sc < j.nc) v = j.src[(size_t)(j.r0 + sr) * j.src_ld + j.c0 + sc];
        j.dst[i] = v;
    }
}

int launch_pack(const PackJobs& jobs, cudaStream_t stream) {
    if (jobs.n == 0) return DRM_OK;
    DRM_REQUIRE(jobs.n <= kMaxPackJobs, DRM_E_SHAPE, "too many pack jobs");
    dim3 grid(16, jobs.n);
    pack_kernel<<<grid, 256, 0, stream>>>(jobs);
    DRM_CUDA_OK(cudaGetLastError());
    return DRM_OK;
}

}  // namespace drm

extern "C" {

int drm_version(void) { return DRM_ABI_VERSION; }
const char* drm_last_error(void) { return drm::g_err; }

size_t drm_sizeof(const char* name) {
    if (!name) return 0;
#define DRM_SZ(T) if (strcmp(name, #T) == 0) return sizeof(T);
    DRM_SZ(drm_mlp)
    DRM_SZ(drm_mlp_grad)
    DRM_SZ(drm_rssm)
    DRM_SZ(drm_rssm_grad)
    DRM_SZ(drm_rollout)
    DRM_SZ(drm_rollout_grad)
    DRM_SZ(drm_mlp_call)
    DRM_SZ(drm_mlp_bwd_call)
    DRM_SZ(drm_returns_loss_call)
    DRM_SZ(drm_kl_call)
#undef DRM_SZ
    return 0;
}

}  // extern "C"
