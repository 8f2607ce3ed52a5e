// Internal interface of the batched-MLP and weight-gradient kernels (shared by the C-ABI entry
// points and by the rollout backward, which reuses the weight-gradient GEMM).
#pragma once
#include "drm_common.cuh"

namespace drm {

// dW[n][koff+k] (+)= sum_m dY[m][n] * X[m][k];  db[n] (+)= sum_m dY[m][n]   (fp32 atomics)
// dW / db must be zero-filled (or hold the value to accumulate onto) before the launch.
int launch_wgrad(const float* dY, int64_t ldy, const float* X, int64_t ldx, int M, int Nout, int K, float* dW,
                 int ld_dw, int koff, float* db, cudaStream_t stream);

// y[M,nout] = x[M,K] . Wblk^T  where Wblk = W[0:nout, c0:c0+K] of a row-major matrix with row stride w_ld
// (no bias, no activation).  pack_ws: >= pack_size(K, nout) floats of scratch.
int launch_linear_nobias(const float* x, int64_t ldx, int M, int K, const float* W, int w_ld, int w_c0, int nout,
                         float* y, float* pack_ws, cudaStream_t stream);
// dx[M,K] = dy[M,nout] . Wblk   (same block addressing).  pack_ws: >= pack_size(nout, K) floats.
int launch_linear_dx(const float* dy, int M, int nout, const float* W, int w_ld, int w_c0, int K, float* dx,
                     int64_t lddx, float* pack_ws, cudaStream_t stream);

// bf16 tensor-core path (drm_mlp_tc.cu)
namespace tc {
int mlp_tc_supported(const drm_mlp_call* c);  // every layer and the input at most 512 wide
size_t mlp_tc_workspace_bytes(const drm_mlp_call* c);
size_t mlp_tc_save_bytes(const drm_mlp_call* c);
int mlp_tc_fwd(const drm_mlp_call* c, cudaStream_t stream);
int mlp_tc_bwd(const drm_mlp_bwd_call* bc, cudaStream_t stream);
}  // namespace tc

}  // namespace drm
