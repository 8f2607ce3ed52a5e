// Batched MLP (DenseModel / ActionDecoder feed-forward stack) forward, data-gradient backward and
// the weight-gradient GEMM.  fp32 FFMA path.  Reference: dreamer/models/dense.py:28-43.
#include "drm_mlp.cuh"
#include <mutex>

namespace drm {

struct MlpDev {
    int M, k0, k1, k0p, k1p, n_linear, hidden, out_dim;
    const float* x0; int64_t ld0;
    const float* x1; int64_t ld1;
    float* y;
    const float* Wp[DRM_MAX_LINEAR];  // packed (fwd: reduce over in-features; bwd: over out-features)
    int ldw[DRM_MAX_LINEAR];
    const float* W[DRM_MAX_LINEAR];   // original
    const float* b[DRM_MAX_LINEAR];
    float* save_h;                    // [n_linear-1][M][hidden] or nullptr
    // backward only
    const float* dy; const float* dy_scale;
    float* dx0; float* dx1; int accumulate_dx;
    float* dysave[DRM_MAX_LINEAR];    // [M][nout_l] or nullptr
};

template <int R>
__global__ void __launch_bounds__(kThreads) mlp_fwd_kernel(const MlpDev p) {
    extern __shared__ __align__(16) float smem[];
    const int inp = p.k0p + p.k1p;
    float* sX = smem;                 // [R][inp]
    float* sA = sX + R * inp;         // [R][hidden]
    float* sB = sA + R * p.hidden;    // [R][hidden]
    const int last = p.n_linear - 1;
    for (int tile = blockIdx.x; tile * R < p.M; tile += gridDim.x) {
        const int row0 = tile * R;
        for (int i = threadIdx.x; i < R * inp; i += kThreads) {
            const int r = i / inp, k = i - r * inp;
            const int row = row0 + r;
            float v = 0.f;
            if (row < p.M) {
                if (k < p.k0p) { if (k < p.k0) v = p.x0[row * p.ld0 + k]; }
                else { const int kk = k - p.k0p; if (kk < p.k1) v = p.x1[row * p.ld1 + kk]; }
            }
            sX[i] = v;
        }
        __syncthreads();
        const float* cur = sX;
        int curK4 = inp, curK = inp;
        float* nxt = sA;
        for (int l = 0; l <= last; ++l) {
            const int nout = (l == last) ? p.out_dim : p.hidden;
            const float* bias = p.b[l];
            if (l == last) {
                auto epi = [&](int r, int col, float v) {
                    const int row = row0 + r;
                    if (row < p.M) p.y[(size_t)row * p.out_dim + col] = v + (bias ? bias[col] : 0.f);
                };
                if (l > 0 && nout <= 16) {
                    const int nrows = min(R, p.M - row0);
                    cta_linear_narrow(p.W[l], curK, curK, cur, curK4, nrows, nout, epi);
                } else {
                    Seg s0{p.Wp[l], cur, curK4, curK4}, s1{nullptr, nullptr, 0, 0};
                    cta_linear<R>(s0, s1, p.ldw[l], nout, epi);
                }
            } else {
                float* sv = p.save_h ? p.save_h + (size_t)l * p.M * p.hidden : nullptr;
                float* dst = nxt;
                const int hid = p.hidden;
                auto epi = [&](int r, int col, float v) {
                    v = elu(v + bias[col]);
                    dst[r * hid + col] = v;
                    const int row = row0 + r;
                    if (sv && row < p.M) sv[(size_t)row * hid + col] = v;
                };
                Seg s0{p.Wp[l], cur, curK4, curK4}, s1{nullptr, nullptr, 0, 0};
                cta_linear<R>(s0, s1, p.ldw[l], nout, epi);
            }
            __syncthreads();
            cur = nxt;
            curK4 = p.hidden; curK = p.hidden;
            nxt = (nxt == sA) ? sB : sA;
        }
    }
}

template <int R>
__global__ void __launch_bounds__(kThreads) mlp_bwd_kernel(const MlpDev p) {
    extern __shared__ __align__(16) float smem[];
    const int out4 = rup(p.out_dim, 4);
    const int wd = max(p.hidden, out4);
    float* sA = smem;            // [R][wd]
    float* sB = sA + R * wd;     // [R][wd]
    const int last = p.n_linear - 1;
    const float scale = p.dy_scale ? __ldg(p.dy_scale) : 1.f;
    const bool want_dx = (p.dx0 != nullptr) || (p.dx1 != nullptr);
    for (int tile = blockIdx.x; tile * R < p.M; tile += gridDim.x) {
        const int row0 = tile * R;
        // dY of the last Linear (no activation after it)
        for (int i = threadIdx.x; i < R * out4; i += kThreads) {
            const int r = i / out4, c = i - r * out4;
            const int row = row0 + r;
            float v = 0.f;
            if (row < p.M && c < p.out_dim) {
                v = p.dy[(size_t)row * p.out_dim + c] * scale;
                if (p.dysave[last]) p.dysave[last][(size_t)row * p.out_dim + c] = v;
            }
            sA[r * out4 + c] = v;
        }
        __syncthreads();
        float* cur = sA;
        float* nxt = sB;
        int curK4 = out4;
        for (int l = last; l >= 0; --l) {
            if (l == 0) {
                if (!want_dx) break;
                const int nin = p.k0 + p.k1;
                auto epi = [&](int r, int col, float v) {
                    const int row = row0 + r;
                    if (row >= p.M) return;
                    float* dst;
                    if (col < p.k0) { if (!p.dx0) return; dst = p.dx0 + row * p.ld0 + col; }
                    else { if (!p.dx1) return; dst = p.dx1 + row * p.ld1 + (col - p.k0); }
                    *dst = p.accumulate_dx ? *dst + v : v;
                };
                Seg s0{p.Wp[0], cur, curK4, curK4}, s1{nullptr, nullptr, 0, 0};
                cta_linear<R>(s0, s1, p.ldw[0], nin, epi);
            } else {
                const int hid = p.hidden;
                const float* hsave = p.save_h + (size_t)(l - 1) * p.M * hid;
                float* dsv = p.dysave[l - 1];
                float* dst = nxt;
                auto epi = [&](int r, int col, float v) {
                    const int row = row0 + r;
                    float g = 0.f;
                    if (row < p.M) {
                        g = v * elu_grad_from_out(hsave[(size_t)row * hid + col]);
                        if (dsv) dsv[(size_t)row * hid + col] = g;
                    }
                    dst[r * hid + col] = g;
                };
                Seg s0{p.Wp[l], cur, curK4, curK4}, s1{nullptr, nullptr, 0, 0};
                cta_linear<R>(s0, s1, p.ldw[l], hid, epi);
            }
            __syncthreads();
            float* t = cur; cur = nxt; nxt = t;
            curK4 = p.hidden;
        }
        __syncthreads();
    }
}

// ------------------------------------------------------------------------------------------
// weight gradient: dW[n][koff+k] += sum_m dY[m][n] X[m][k]   (64x64 output tile, split over m)
// Stages of 32 rows, double-buffered in shared memory; the global loads of stage s+1 are issued (into registers) before
// the FFMAs of stage s, so their latency hides under the math (the first version loaded, synchronised and computed in
// turn and ran at 12 % of the FFMA peak: 4.6 of the 10.1 ms of an fp32 walker step).
// ------------------------------------------------------------------------------------------
constexpr int WG_T = 64, WG_M = 32;

__global__ void __launch_bounds__(256) wgrad_kernel(const float* __restrict__ dY, int64_t ldy,
                                                    const float* __restrict__ X, int64_t ldx, int M, int Nout,
                                                    int K, float* dW, int ld_dw, int koff, float* db,
                                                    int m_per_split) {
    __shared__ __align__(16) float sY[2][WG_M][WG_T];
    __shared__ __align__(16) float sX[2][WG_M][WG_T];
    const int n0 = blockIdx.x * WG_T, k0 = blockIdx.y * WG_T;
    const int m_begin = blockIdx.z * m_per_split;
    const int m_end = min(M, m_begin + m_per_split);
    const int tx = threadIdx.x & 15, ty = threadIdx.x >> 4;
    // loader geometry: thread -> rows lr, lr + 16 of the stage, 4 adjacent columns lc .. lc + 3
    const int lr = threadIdx.x >> 4, lc = (threadIdx.x & 15) * 4;
    // 16-byte loads when the 4 columns are inside the matrix and aligned for every row
    const bool vy = (n0 + lc + 3 < Nout) && ((ldy & 3) == 0) && (((size_t)(dY + n0 + lc) & 15) == 0);
    const bool vx = (k0 + lc + 3 < K) && ((ldx & 3) == 0) && (((size_t)(X + k0 + lc) & 15) == 0);
    float4 ry[2], rx[2];
    auto load_stage = [&](int m0) {
#pragma unroll
        for (int i = 0; i < 2; ++i) {
            const int m = m0 + lr + 16 * i;
            float4 a = make_float4(0.f, 0.f, 0.f, 0.f), b = a;
            if (m < m_end) {
                const float* py = dY + (size_t)m * ldy + n0 + lc;
                const float* px = X + (size_t)m * ldx + k0 + lc;
                if (vy) a = __ldg(reinterpret_cast<const float4*>(py));
                else {
                    if (n0 + lc + 0 < Nout) a.x = __ldg(py + 0);
                    if (n0 + lc + 1 < Nout) a.y = __ldg(py + 1);
                    if (n0 + lc + 2 < Nout) a.z = __ldg(py + 2);
                    if (n0 + lc + 3 < Nout) a.w = __ldg(py + 3);
                }
                if (vx) b = __ldg(reinterpret_cast<const float4*>(px));
                else {
                    if (k0 + lc + 0 < K) b.x = __ldg(px + 0);
                    if (k0 + lc + 1 < K) b.y = __ldg(px + 1);
                    if (k0 + lc + 2 < K) b.z = __ldg(px + 2);
                    if (k0 + lc + 3 < K) b.w = __ldg(px + 3);
                }
            }
            ry[i] = a; rx[i] = b;
        }
    };
    auto store_stage = [&](int buf) {
#pragma unroll
        for (int i = 0; i < 2; ++i) {
            *reinterpret_cast<float4*>(&sY[buf][lr + 16 * i][lc]) = ry[i];
            *reinterpret_cast<float4*>(&sX[buf][lr + 16 * i][lc]) = rx[i];
        }
    };
    float acc[4][4];
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;
    float bsum[4] = {0.f, 0.f, 0.f, 0.f};
    const bool do_bias = (db != nullptr) && (blockIdx.y == 0) && (tx == 0);
    if (m_begin < m_end) {
        load_stage(m_begin);
        store_stage(0);
    }
    __syncthreads();
    int buf = 0;
    for (int m0 = m_begin; m0 < m_end; m0 += WG_M, buf ^= 1) {
        const bool more = m0 + WG_M < m_end;
        if (more) load_stage(m0 + WG_M);  // in flight during the FFMAs below
#pragma unroll
        for (int mm = 0; mm < WG_M; ++mm) {
            const float4 a = *reinterpret_cast<const float4*>(&sY[buf][mm][ty * 4]);
            const float4 b = *reinterpret_cast<const float4*>(&sX[buf][mm][tx * 4]);
            const float av[4] = {a.x, a.y, a.z, a.w};
            const float bv[4] = {b.x, b.y, b.z, b.w};
#pragma unroll
            for (int i = 0; i < 4; ++i) {
#pragma unroll
                for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(av[i], bv[j], acc[i][j]);
                bsum[i] += av[i];
            }
        }
        if (more) store_stage(buf ^ 1);
        __syncthreads();
    }
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const int n = n0 + ty * 4 + i;
        if (n >= Nout) continue;
        // vector reductions where the four columns are inside and aligned: one L2 operation instead of four (the ~20
        // row splits of a launch add up on the same addresses, and the L2 atomic rate is what bounds this epilogue)
        float* dst = dW + (size_t)n * ld_dw + koff + k0 + tx * 4;
        if (k0 + tx * 4 + 3 < K && ((size_t)dst & 15) == 0) {
            asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(dst), "f"(acc[i][0]), "f"(acc[i][1]),
                         "f"(acc[i][2]), "f"(acc[i][3]) : "memory");
        } else if (k0 + tx * 4 + 3 < K && ((size_t)dst & 7) == 0) {
            asm volatile("red.global.add.v2.f32 [%0], {%1, %2};" ::"l"(dst), "f"(acc[i][0]), "f"(acc[i][1]) : "memory");
            asm volatile("red.global.add.v2.f32 [%0], {%1, %2};" ::"l"(dst + 2), "f"(acc[i][2]), "f"(acc[i][3]) : "memory");
        } else {
#pragma unroll
            for (int j = 0; j < 4; ++j)
                if (k0 + tx * 4 + j < K) atomicAdd(dst + j, acc[i][j]);
        }
        if (do_bias) atomicAdd(db + n, bsum[i]);
    }
}

int launch_wgrad(const float* dY, int64_t ldy, const float* X, int64_t ldx, int M, int Nout, int K, float* dW,
                 int ld_dw, int koff, float* db, cudaStream_t stream) {
    if (M <= 0 || Nout <= 0 || K <= 0) return DRM_OK;
    const int tn = (Nout + WG_T - 1) / WG_T, tk = (K + WG_T - 1) / WG_T;
    int splits = (4 * num_sms() + tn * tk - 1) / (tn * tk);
    const int max_splits = (M + 4 * WG_M - 1) / (4 * WG_M);
    if (splits > max_splits) splits = max_splits;
    if (splits < 1) splits = 1;
    if (splits > 65535) splits = 65535;
    int mps = rup((M + splits - 1) / splits, WG_M);
    splits = (M + mps - 1) / mps;
    dim3 grid(tn, tk, splits);
    ProfScope prof(PROF_WGRAD, stream);
    wgrad_kernel<<<grid, 256, 0, stream>>>(dY, ldy, X, ldx, M, Nout, K, dW, ld_dw, koff, db, mps);
    DRM_CUDA_OK(cudaGetLastError());
    return DRM_OK;
}

// ------------------------------------------------------------------------------------------
// host side
// ------------------------------------------------------------------------------------------
static int validate(const drm_mlp_call* c) {
    DRM_REQUIRE(c != nullptr, DRM_E_ARG, "null call");
    const drm_mlp& m = c->mlp;
    DRM_REQUIRE(m.n_linear >= 1 && m.n_linear <= DRM_MAX_LINEAR, DRM_E_SHAPE, "n_linear out of range");
    DRM_REQUIRE(m.in_dim > 0 && m.out_dim > 0 && c->M >= 0, DRM_E_SHAPE, "bad dims");
    DRM_REQUIRE(m.n_linear == 1 || (m.hidden > 0 && m.hidden % 4 == 0), DRM_E_SHAPE,
                "hidden size must be a positive multiple of 4");
    DRM_REQUIRE(c->k0 > 0 && c->k1 >= 0 && c->k0 + c->k1 == m.in_dim, DRM_E_SHAPE, "k0 + k1 != in_dim");
    DRM_REQUIRE(c->precision == DRM_FP32 || c->precision == DRM_BF16, DRM_E_SHAPE, "unknown precision");
    return DRM_OK;
}

static size_t fwd_pack_floats(const drm_mlp_call* c) {
    const drm_mlp& m = c->mlp;
    size_t f = 0;
    for (int l = 0; l < m.n_linear; ++l) {
        const int nout = l == m.n_linear - 1 ? m.out_dim : m.hidden;
        const int kin4 = l == 0 ? rup(c->k0, 4) + rup(c->k1, 4) : m.hidden;
        f += (size_t)kin4 * rup(nout, 32);
    }
    return f;
}
static size_t bwd_pack_floats(const drm_mlp_call* c) {
    const drm_mlp& m = c->mlp;
    size_t f = 0;
    for (int l = 0; l < m.n_linear; ++l) {
        const int nout = l == m.n_linear - 1 ? m.out_dim : m.hidden;
        const int nin = l == 0 ? m.in_dim : m.hidden;
        f += pack_size(nout, nin);
    }
    return f;
}
static size_t dysave_floats(const drm_mlp_call* c) {
    const drm_mlp& m = c->mlp;
    size_t f = 0;
    for (int l = 0; l < m.n_linear; ++l) f += rup_sz((size_t)c->M * (l == m.n_linear - 1 ? m.out_dim : m.hidden), 4);
    return f;
}

static void fill_common(MlpDev& p, const drm_mlp_call* c) {
    const drm_mlp& m = c->mlp;
    p.M = c->M; p.k0 = c->k0; p.k1 = c->k1; p.k0p = rup(c->k0, 4); p.k1p = rup(c->k1, 4);
    p.n_linear = m.n_linear; p.hidden = m.n_linear > 1 ? m.hidden : 4; p.out_dim = m.out_dim;
    p.x0 = c->x0; p.ld0 = c->ld0; p.x1 = c->x1; p.ld1 = c->ld1; p.y = c->y;
    for (int l = 0; l < DRM_MAX_LINEAR; ++l) {
        p.W[l] = l < m.n_linear ? m.W[l] : nullptr;
        p.b[l] = l < m.n_linear ? m.b[l] : nullptr;
        p.Wp[l] = nullptr; p.ldw[l] = 0; p.dysave[l] = nullptr;
    }
    p.save_h = nullptr; p.dy = nullptr; p.dy_scale = nullptr; p.dx0 = p.dx1 = nullptr; p.accumulate_dx = 0;
}


static int launch_mlp_kernel(const MlpDev& p, bool backward, cudaStream_t stream) {
    size_t per_row;
    if (!backward) per_row = (size_t)(p.k0p + p.k1p + 2 * p.hidden) * sizeof(float);
    else {
        const int wd = p.hidden > rup(p.out_dim, 4) ? p.hidden : rup(p.out_dim, 4);
        per_row = (size_t)2 * wd * sizeof(float);
    }
    const int Rv = pick_rows(p.M, per_row, 0);
    DRM_REQUIRE(Rv > 0, DRM_E_SHAPE, "layer too wide for shared memory");
    const size_t smem = per_row * Rv;
    const int tiles = (p.M + Rv - 1) / Rv;
    const int grid = tiles < 2 * num_sms() ? tiles : 2 * num_sms();
    ProfScope prof(backward ? PROF_MLP_BWD : PROF_MLP_FWD, stream);
    if (!backward) {
        DRM_DISPATCH_R(Rv, {
            DRM_CUDA_OK(cudaFuncSetAttribute(mlp_fwd_kernel<R>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
            mlp_fwd_kernel<R><<<grid, kThreads, smem, stream>>>(p);
        });
    } else {
        DRM_DISPATCH_R(Rv, {
            DRM_CUDA_OK(cudaFuncSetAttribute(mlp_bwd_kernel<R>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
            mlp_bwd_kernel<R><<<grid, kThreads, smem, stream>>>(p);
        });
    }
    DRM_CUDA_OK(cudaGetLastError());
    return DRM_OK;
}

static void blank(MlpDev& p) {
    for (int l = 0; l < DRM_MAX_LINEAR; ++l) { p.W[l] = nullptr; p.b[l] = nullptr; p.Wp[l] = nullptr; p.ldw[l] = 0; p.dysave[l] = nullptr; }
    p.x0 = p.x1 = nullptr; p.ld0 = p.ld1 = 0; p.y = nullptr; p.save_h = nullptr; p.dy = nullptr; p.dy_scale = nullptr;
    p.dx0 = p.dx1 = nullptr; p.accumulate_dx = 0; p.k1 = 0; p.k1p = 0; p.hidden = 4; p.n_linear = 1;
}

int launch_linear_nobias(const float* x, int64_t ldx, int M, int K, const float* W, int w_ld, int w_c0, int nout,
                         float* y, float* pack_ws, cudaStream_t stream) {
    if (M <= 0) return DRM_OK;
    MlpDev p; blank(p);
    p.M = M; p.k0 = K; p.k0p = rup(K, 4); p.out_dim = nout; p.x0 = x; p.ld0 = ldx; p.y = y;
    Packer pk(pack_ws);
    p.Wp[0] = pk.fwd(W, w_ld, 0, nout, w_c0, K);
    p.ldw[0] = rup(nout, 32);
    int rc = launch_pack(pk.jobs, stream); if (rc) return rc;
    return launch_mlp_kernel(p, false, stream);
}

int launch_linear_dx(const float* dy, int M, int nout, const float* W, int w_ld, int w_c0, int K, float* dx,
                     int64_t lddx, float* pack_ws, cudaStream_t stream) {
    if (M <= 0) return DRM_OK;
    MlpDev p; blank(p);
    p.M = M; p.k0 = K; p.k0p = rup(K, 4); p.out_dim = nout; p.dy = dy; p.dx0 = dx; p.ld0 = lddx;
    Packer pk(pack_ws);
    p.Wp[0] = pk.bwd(W, w_ld, 0, nout, w_c0, K);
    p.ldw[0] = rup(K, 32);
    int rc = launch_pack(pk.jobs, stream); if (rc) return rc;
    return launch_mlp_kernel(p, true, stream);
}

void warm_mlp() { warm_kernel(mlp_fwd_kernel<1>); warm_kernel(mlp_fwd_kernel<2>); warm_kernel(mlp_bwd_kernel<1>); warm_kernel(mlp_bwd_kernel<2>); warm_kernel(mlp_fwd_kernel<4>); warm_kernel(mlp_fwd_kernel<8>); warm_kernel(mlp_fwd_kernel<12>); warm_kernel(mlp_fwd_kernel<16>); warm_kernel(mlp_fwd_kernel<20>); warm_kernel(mlp_fwd_kernel<24>); warm_kernel(mlp_fwd_kernel<32>); warm_kernel(mlp_bwd_kernel<4>); warm_kernel(mlp_bwd_kernel<8>); warm_kernel(mlp_bwd_kernel<12>); warm_kernel(mlp_bwd_kernel<16>); warm_kernel(mlp_bwd_kernel<20>); warm_kernel(mlp_bwd_kernel<24>); warm_kernel(mlp_bwd_kernel<32>); warm_kernel(wgrad_kernel); }

}  // namespace drm

using namespace drm;

// bf16 requests run on the tcgen05 kernels when they cover the shape (every layer and the input at most 512 wide: the
// reference's heads are 200-300 wide on 230 features); wider MLPs (BASELINE configs[4]: 660 input features) run on the
// fp32 kernels of this library -- still CUDA, more accurate, slower.
static bool use_tc(const drm_mlp_call* c) { return c->precision == DRM_BF16 && tc::mlp_tc_supported(c); }

extern "C" {

size_t drm_mlp_save_bytes(const drm_mlp_call* c) {
    if (c && use_tc(c)) return tc::mlp_tc_save_bytes(c) + 256;
    if (!c || c->mlp.n_linear < 2) return 16;
    return rup_sz((size_t)(c->mlp.n_linear - 1) * c->M * c->mlp.hidden * sizeof(float), 16) + 16;
}

size_t drm_mlp_workspace_bytes(const drm_mlp_call* c) {
    if (!c) return 0;
    size_t f = fwd_pack_floats(c);
    size_t b = bwd_pack_floats(c) + dysave_floats(c);
    size_t bytes = (f > b ? f : b) * sizeof(float) + 256;
    if (use_tc(c)) {
        const size_t t = tc::mlp_tc_workspace_bytes(c) + 256;
        if (t > bytes) bytes = t;
    }
    return bytes;
}

// fp32 path + producer gating: the FFMA kernels do not poll, so one thread waits for the LAST step's flag first
__global__ void wait_flag_kernel(const int* flag, int target) {
    unsigned long long t0 = 0;
    for (;;) {
        int v;
        asm volatile("ld.acquire.gpu.global.s32 %0, [%1];" : "=r"(v) : "l"(flag) : "memory");
        if (v >= target) return;
        __nanosleep(500);
        unsigned long long now;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(now));
        if (t0 == 0) t0 = now;
        else if (now - t0 > 30000000000ull) __trap();  // 30 s: the producer is not running
    }
}

int drm_mlp_fwd(const drm_mlp_call* c, void* stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    int rc = validate(c); if (rc) return rc;
    rc = check_device(); if (rc) return rc;
    if (c->M == 0) return DRM_OK;
    const drm_mlp& m = c->mlp;
    DRM_REQUIRE(c->x0 && c->y && (c->k1 == 0 || c->x1), DRM_E_ARG, "null io pointer");
    DRM_REQUIRE(c->workspace && c->workspace_bytes >= drm_mlp_workspace_bytes(c), DRM_E_WORKSPACE, "workspace too small");
    if (use_tc(c)) return tc::mlp_tc_fwd(c, stream);
    if (c->wait_flags) {
        DRM_REQUIRE(c->wait_steps > 0, DRM_E_ARG, "wait_flags need wait_steps");
        wait_flag_kernel<<<1, 1, 0, stream>>>(c->wait_flags + c->wait_steps - 1, c->wait_target);
        DRM_CUDA_OK(cudaGetLastError());
    }
    const bool save = c->training && m.n_linear > 1;
    if (save) DRM_REQUIRE(c->save && c->save_bytes >= drm_mlp_save_bytes(c), DRM_E_WORKSPACE, "save buffer too small");
    MlpDev p; fill_common(p, c);
    Packer pk((float*)c->workspace);
    for (int l = 0; l < m.n_linear; ++l) {
        const int nout = l == m.n_linear - 1 ? m.out_dim : m.hidden;
        if (l == 0) {
            p.Wp[0] = pk.fwd(m.W[0], m.in_dim, 0, nout, 0, c->k0);
            if (c->k1 > 0) pk.fwd(m.W[0], m.in_dim, 0, nout, c->k0, c->k1);
        } else {
            p.Wp[l] = pk.fwd(m.W[l], m.hidden, 0, nout, 0, m.hidden);
        }
        p.ldw[l] = rup(nout, 32);
    }
    rc = launch_pack(pk.jobs, stream); if (rc) return rc;
    p.save_h = save ? (float*)c->save : nullptr;
    rc = launch_mlp_kernel(p, false, stream); if (rc) return rc;
    return DRM_OK;
}

int drm_mlp_bwd(const drm_mlp_bwd_call* bc, void* stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    DRM_REQUIRE(bc != nullptr, DRM_E_ARG, "null call");
    const drm_mlp_call* c = &bc->fwd;
    int rc = validate(c); if (rc) return rc;
    rc = check_device(); if (rc) return rc;
    const drm_mlp& m = c->mlp;
    DRM_REQUIRE(bc->phase >= DRM_PHASE_ALL && bc->phase <= DRM_PHASE_RUN, DRM_E_ARG, "bad phase");
    const bool tc_path = use_tc(c);
    DRM_REQUIRE(tc_path || !(bc->tile_done || bc->tile_wait || bc->step_rows_done), DRM_E_ARG,
                "tile_done / tile_wait / step_rows_done are honoured by the bf16 kernels only");
    if (bc->phase == DRM_PHASE_PREPARE && !tc_path) return DRM_OK;  // fp32: everything happens in RUN
    DRM_REQUIRE(bc->dy || bc->phase == DRM_PHASE_PREPARE, DRM_E_ARG, "null dy");
    DRM_REQUIRE(c->workspace && c->workspace_bytes >= drm_mlp_workspace_bytes(c), DRM_E_WORKSPACE, "workspace too small");
    if (m.n_linear > 1) DRM_REQUIRE(c->save && c->save_bytes >= drm_mlp_save_bytes(c), DRM_E_WORKSPACE, "save buffer missing");
    const bool want_dx = bc->dx0 || bc->dx1;
    // zero-fill gradient sinks first (they are accumulated with atomics)
    if (bc->want_wgrad) {
        for (int l = 0; l < m.n_linear; ++l) {
            const int nout = l == m.n_linear - 1 ? m.out_dim : m.hidden;
            const int nin = l == 0 ? m.in_dim : m.hidden;
            DRM_REQUIRE(bc->grad.dW[l] && bc->grad.db[l], DRM_E_ARG, "null weight-gradient sink");
            if (tc_path && c->M > 0) continue;  // zero-filled by the weight packing launch
            DRM_CUDA_OK(cudaMemsetAsync(bc->grad.dW[l], 0, (size_t)nout * nin * sizeof(float), stream));
            DRM_CUDA_OK(cudaMemsetAsync(bc->grad.db[l], 0, (size_t)nout * sizeof(float), stream));
        }
    }
    if (c->M == 0) return DRM_OK;
    if (tc_path) return tc::mlp_tc_bwd(bc, stream);
    MlpDev p; fill_common(p, c);
    Packer pk((float*)c->workspace);
    for (int l = 0; l < m.n_linear; ++l) {
        const int nout = l == m.n_linear - 1 ? m.out_dim : m.hidden;
        const int nin = l == 0 ? m.in_dim : m.hidden;
        if (l == 0 && !want_dx) continue;
        p.Wp[l] = pk.bwd(m.W[l], nin, 0, nout, 0, nin);
        p.ldw[l] = rup(nin, 32);
    }
    float* wsf = (float*)c->workspace + bwd_pack_floats(c);
    if (bc->want_wgrad) {
        for (int l = 0; l < m.n_linear; ++l) {
            p.dysave[l] = wsf;
            wsf += rup_sz((size_t)c->M * (l == m.n_linear - 1 ? m.out_dim : m.hidden), 4);
        }
    }
    rc = launch_pack(pk.jobs, stream); if (rc) return rc;
    p.save_h = (float*)c->save;
    p.dy = bc->dy; p.dy_scale = bc->dy_scale; p.dx0 = bc->dx0; p.dx1 = bc->dx1; p.accumulate_dx = bc->accumulate_dx;
    rc = launch_mlp_kernel(p, true, stream); if (rc) return rc;
    if (bc->want_wgrad) {
        for (int l = 0; l < m.n_linear; ++l) {
            const int nout = l == m.n_linear - 1 ? m.out_dim : m.hidden;
            if (l == 0) {
                rc = launch_wgrad(p.dysave[0], nout, c->x0, c->ld0, c->M, nout, c->k0, bc->grad.dW[0], m.in_dim, 0,
                                  bc->grad.db[0], stream);
                if (rc) return rc;
                if (c->k1 > 0) {
                    rc = launch_wgrad(p.dysave[0], nout, c->x1, c->ld1, c->M, nout, c->k1, bc->grad.dW[0], m.in_dim,
                                      c->k0, nullptr, stream);
                    if (rc) return rc;
                }
            } else {
                const float* h = (const float*)c->save + (size_t)(l - 1) * c->M * m.hidden;
                rc = launch_wgrad(p.dysave[l], nout, h, m.hidden, c->M, nout, m.hidden, bc->grad.dW[l], m.hidden, 0,
                                  bc->grad.db[l], stream);
                if (rc) return rc;
            }
        }
    }
    return DRM_OK;
}

}  // extern "C"

// ---------------------------------------------------------------------------------------------
// drm_mlp_fwd_noalloc: the fp32 inference forward of drm_mlp_fwd for callers that cannot allocate device memory (code
// under CUDA-graph capture, tight serving loops, harnesses that forbid cudaMalloc).  No counterpart in the reference.
// The only scratch the fp32 forward needs -- the re-laid (packed) weight
// matrices -- lives in a file-scope __device__ array instead of a caller-provided workspace: DRM_NOALLOC_SCRATCH_BYTES per
// GPU that loads the library, held until the process exits.  Every call on a device shares that array, so calls that
// overlap on the GPU (two streams, two threads) race; issue them in one stream order.  The array keeps the previous
// call's contents; nothing needs to start at zero, because the packing launch rewrites every float (padding included)
// of the region the MLP kernel then reads.
// ---------------------------------------------------------------------------------------------
namespace drm {
constexpr size_t kNoAllocScratchFloats = DRM_NOALLOC_SCRATCH_BYTES / sizeof(float);
static __device__ __align__(256) float g_noalloc_scratch[kNoAllocScratchFloats];

// In host code `g_noalloc_scratch` is only a stand-in for the symbol, not a device address: the address is looked up
// once per device, with that device current, and cached.
constexpr int kNoAllocMaxDevices = 64;
static float* g_noalloc_ptr[kNoAllocMaxDevices] = {};
static std::mutex g_noalloc_mu;

static int noalloc_scratch(float** out) {
    int dev = 0;
    DRM_CUDA_OK(cudaGetDevice(&dev));
    DRM_REQUIRE(dev >= 0 && dev < kNoAllocMaxDevices, DRM_E_ARG, "device ordinal beyond 63");
    std::lock_guard<std::mutex> lock(g_noalloc_mu);
    if (g_noalloc_ptr[dev] == nullptr) {
        float* ptr = nullptr;
        DRM_CUDA_OK(cudaGetSymbolAddress((void**)&ptr, g_noalloc_scratch));
        DRM_REQUIRE(ptr != nullptr && ((size_t)ptr & 255) == 0, DRM_E_CUDA, "bad address for the static scratch array");
        g_noalloc_ptr[dev] = ptr;
    }
    *out = g_noalloc_ptr[dev];
    return DRM_OK;
}
}  // namespace drm

extern "C" {

size_t drm_mlp_noalloc_scratch_bytes(const drm_mlp_call* c) { return c ? fwd_pack_floats(c) * sizeof(float) : 0; }

int drm_mlp_fwd_noalloc(const drm_mlp_call* c, void* stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    int rc = validate(c); if (rc) return rc;
    rc = check_device(); if (rc) return rc;
    const drm_mlp& m = c->mlp;
    DRM_REQUIRE(c->precision == DRM_FP32, DRM_E_ARG, "drm_mlp_fwd_noalloc runs the fp32 kernels only");
    DRM_REQUIRE(!c->training, DRM_E_ARG, "drm_mlp_fwd_noalloc is inference only (training = 0, no save buffer)");
    DRM_REQUIRE(c->wait_flags == nullptr, DRM_E_ARG, "drm_mlp_fwd_noalloc does not take producer gating");
    const size_t need = fwd_pack_floats(c);
    if (need > kNoAllocScratchFloats) {
        set_error("drm_mlp_fwd_noalloc: packed weights need %zu bytes, the static scratch holds %zu "
                  "(DRM_NOALLOC_SCRATCH_BYTES); use drm_mlp_fwd with a caller workspace", need * sizeof(float),
                  (size_t)DRM_NOALLOC_SCRATCH_BYTES);
        return DRM_E_WORKSPACE;
    }
    if (c->M == 0) return DRM_OK;
    DRM_REQUIRE(c->x0 && c->y && (c->k1 == 0 || c->x1), DRM_E_ARG, "null io pointer");
    float* scratch = nullptr;
    rc = noalloc_scratch(&scratch); if (rc) return rc;
    MlpDev p; fill_common(p, c);
    Packer pk(scratch);
    for (int l = 0; l < m.n_linear; ++l) {
        const int nout = l == m.n_linear - 1 ? m.out_dim : m.hidden;
        if (l == 0) {
            p.Wp[0] = pk.fwd(m.W[0], m.in_dim, 0, nout, 0, c->k0);
            if (c->k1 > 0) pk.fwd(m.W[0], m.in_dim, 0, nout, c->k0, c->k1);
        } else {
            p.Wp[l] = pk.fwd(m.W[l], m.hidden, 0, nout, 0, m.hidden);
        }
        p.ldw[l] = rup(nout, 32);
    }
    DRM_REQUIRE(pk.off <= kNoAllocScratchFloats, DRM_E_WORKSPACE, "packed weights overflow the static scratch");
    rc = launch_pack(pk.jobs, stream); if (rc) return rc;
    return launch_mlp_kernel(p, false, stream);
}

}  // extern "C"
