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

int num_sms() {
    static int cached_dev = -1, cached = 148;
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess) return cached;
    if (dev != cached_dev) {
        int n = 0;
        if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) == cudaSuccess && n > 0) cached = n;
        cached_dev = dev;
    }
    return cached;
}

// ---------------------------------------------------------------------------------------------
// launch counter + optional per-kernel CUDA-event timing
// ---------------------------------------------------------------------------------------------
static unsigned long long g_launches = 0;
static int g_prof_on = 0;
constexpr int kProfCap = 8192;
struct ProfRec { cudaEvent_t a, b; int slot; };
static ProfRec g_prof[kProfCap];
static int g_prof_n = 0;       // records in use
static int g_prof_made = 0;    // events created so far (reused across enable/reset)

ProfScope::ProfScope(int slot_, cudaStream_t stream_) : slot(slot_), stream(stream_), idx(-1) {
    ++g_launches;
    if (!g_prof_on || g_prof_n >= kProfCap) return;
    idx = g_prof_n++;
    if (idx >= g_prof_made) {
        cudaEventCreate(&g_prof[idx].a);
        cudaEventCreate(&g_prof[idx].b);
        g_prof_made = idx + 1;
    }
    g_prof[idx].slot = slot;
    cudaEventRecord(g_prof[idx].a, stream);
}
ProfScope::~ProfScope() {
    if (idx >= 0) cudaEventRecord(g_prof[idx].b, stream);
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
    ProfScope prof(PROF_PACK, stream);
    pack_kernel<<<grid, 256, 0, stream>>>(jobs);
    DRM_CUDA_OK(cudaGetLastError());
    return DRM_OK;
}

void warm_util() { warm_kernel(pack_kernel); }

}  // namespace drm

extern "C" {

int drm_version(void) { return DRM_ABI_VERSION; }

int drm_init(void) {
    int rc = drm::check_device(); if (rc) return rc;
    static int done_dev = -1;
    int dev = 0;
    DRM_CUDA_OK(cudaGetDevice(&dev));
    if (dev == done_dev) return DRM_OK;
    drm::warm_util(); drm::warm_mlp(); drm::warm_mlp_tc(); drm::warm_wgrad_tc(); drm::warm_rollout_fwd();
    drm::warm_rollout_bwd(); drm::warm_rollout_tc(); drm::warm_loss(); drm::warm_optim(); drm::warm_policy(); drm::warm_replay(); drm::warm_noise(); drm::warm_allreduce(); drm::warm_conv_glue();
    DRM_CUDA_OK(cudaGetLastError());
    done_dev = dev;
    return DRM_OK;
}
int drm_num_sms(void) { return drm::num_sms(); }
const char* drm_last_error(void) { return drm::g_err; }

unsigned long long drm_launch_count(void) { return drm::g_launches; }

void drm_prof_enable(int on) {
    drm::g_prof_on = on ? 1 : 0;
    drm::g_prof_n = 0;
}

int drm_prof_read(int slot, double* total_ms, int* count) {
    if (slot < 0 || slot >= drm::PROF_N || !total_ms || !count) return DRM_E_ARG;
    double tot = 0.0;
    int n = 0;
    for (int i = 0; i < drm::g_prof_n; ++i) {
        if (drm::g_prof[i].slot != slot) continue;
        if (cudaEventSynchronize(drm::g_prof[i].b) != cudaSuccess) return DRM_E_CUDA;
        float ms = 0.f;
        if (cudaEventElapsedTime(&ms, drm::g_prof[i].a, drm::g_prof[i].b) != cudaSuccess) return DRM_E_CUDA;
        tot += ms;
        ++n;
    }
    *total_ms = tot;
    *count = n;
    return DRM_OK;
}

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
    DRM_SZ(drm_clip_adam_call)
    DRM_SZ(drm_image_logprob_call)
    DRM_SZ(drm_policy_mode_call)
    DRM_SZ(drm_replay_gather_call)
    DRM_SZ(drm_noise_fill_call)
    DRM_SZ(drm_allreduce_call)
    DRM_SZ(drm_bias_act_call)
#undef DRM_SZ
    return 0;
}

}  // extern "C"
