// Device-side parameter block and buffer layouts of the persistent rollout kernels.
#pragma once
#include "drm_common.cuh"

namespace drm {

// Saved-for-backward activations, all [steps, n, width] fp32, carved out of drm_rollout.save.
struct RollSave {
    float* h[DRM_MAX_LINEAR];  // actor hidden activations (post-ELU), width ah        (POLICY)
    float* raw;                // actor output before the distribution, width Aout      (POLICY)
    float* x;                  // ELU(rnn input), width Hd
    float* r; float* z; float* nn; float* hn;  // GRU gates and W_hn h + b_hn, width D
    float* hid;                // prior hidden (post-ELU), width Hd
    float* hidq;               // posterior hidden (post-ELU), width Hd                 (REPRESENTATION)
};

// dY streams written by the backward kernel for the weight-gradient GEMMs, [steps, n, width].
struct RollDy {
    float* p2;    // 2S   d(prior mean | raw std)
    float* p1;    // Hd
    float* ih;    // 3D   d(gi) = (dpre_r, dpre_z, d i_n)
    float* hh;    // 3D   d(gh) = (dpre_r, dpre_z, d h_n)
    float* in;    // Hd
    float* q2;    // 2S
    float* q1;    // Hd   (also d embproj, source of d_obs_embed)
    float* a[DRM_MAX_LINEAR];  // actor layers: width ah, last Aout
};

struct RollDev {
    int mode, T, n, act_dist, training;
    int A, S, D, Hd, E, A4, S4, MS4, AS4;
    int an, ah, Aout, Aout4;      // actor: number of Linear layers, hidden width, output width
    int U;                         // width of the union scratch region (floats per row)
    float mean_scale, raw_init_std, act_min_std, temperature, min_std;
    // packed weights (see Packer): forward kernel gets the transposed set, backward the plain set
    const float* Win;   int ld_in;    // fwd: rows A4+S4, ld rup(Hd);    bwd: rows Hd,  ld rup(A+S)
    const float* Wrz;   int ld_rz;    // fwd: rows Hd+D, ld rup(2D)
    const float* Wni;   int ld_n;     // fwd: rows Hd,  ld rup(D)
    const float* Wnh;                 // fwd: rows D,   ld rup(D)
    const float* Wih;   int ld_ih;    // bwd: rows 3D,  ld rup(Hd)
    const float* Whh;   int ld_hh;    // bwd: rows 3D,  ld rup(D)
    const float* Wp1;   int ld_p1;    // fwd: rows D,  ld rup(Hd);       bwd: rows Hd, ld rup(D)
    const float* Wp2;   int ld_p2;    // fwd: rows Hd, ld rup(2S);       bwd: rows MS4, ld rup(Hd)
    const float* Wq1;   int ld_q1;    // deter block of W_q1, same shapes as Wp1
    const float* Wq2;   int ld_q2;
    const float* Wa[DRM_MAX_LINEAR]; int ld_a[DRM_MAX_LINEAR];
    // original tensors (biases; last actor layer for the narrow path)
    const float *b_in, *b_ih, *b_hh, *b_p1, *b_p2, *b_q1, *b_q2;
    const float* ba[DRM_MAX_LINEAR];
    const float* Wa_last;
    // io
    const float *start_stoch, *start_deter, *action_in, *embproj, *eps_prior, *eps_post, *noise_act;
    const unsigned long long* nz;  // {seed, offset} of the in-kernel noise streams (drm_philox.cuh) for NULL eps / noise pointers
    float *prior_mean, *prior_std, *prior_stoch, *deter, *post_mean, *post_std, *post_stoch, *action_out;
    RollSave sv;
    // backward
    const float *g_prior_mean, *g_prior_std, *g_prior_stoch, *g_deter, *g_post_mean, *g_post_std, *g_post_stoch,
        *g_action_out;
    float *d_start_stoch, *d_start_deter, *d_action_in;
    RollDy dy;
    int want_rssm_dy, want_actor_dy, want_q1_dy;
};

// floats of shared memory per tile row
__host__ __device__ inline int fwd_smem_floats_per_row(const RollDev& p) { return p.S4 + (p.Hd + p.D) + p.A4 + p.U + p.MS4 + p.Aout4; }
__host__ __device__ inline int bwd_smem_floats_per_row(const RollDev& p) {
    return p.S4 + p.D + p.MS4 + p.Hd + p.D + p.U + p.AS4 + p.Aout4;
}

// bf16 tensor-core path (drm_rollout_tc.cu): POLICY mode only
namespace tc {
int rollout_tc_supported(const drm_rollout* r);
size_t rollout_tc_save_bytes(const drm_rollout* r);
size_t rollout_tc_fwd_workspace_bytes(const drm_rollout* r);
int rollout_tc_tiles(const drm_rollout* r);
size_t rollout_tc_bwd_workspace_bytes(const drm_rollout* r);
int rollout_tc_fwd(const drm_rollout* r, cudaStream_t stream);
int rollout_tc_bwd(const drm_rollout* r, const drm_rollout_grad* g, cudaStream_t stream);
}  // namespace tc

int launch_rollout_fwd(const RollDev& p, cudaStream_t stream);
int launch_rollout_bwd(const RollDev& p, cudaStream_t stream);

}  // namespace drm
