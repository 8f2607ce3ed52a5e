// C-ABI entry points of the rollouts: argument validation, buffer carving, weight packing,
// kernel launches and the weight-gradient GEMMs that follow the reverse-time kernel.
#include "drm_mlp.cuh"
#include "drm_rollout.cuh"

namespace drm {

static int validate(const drm_rollout* r) {
    DRM_REQUIRE(r != nullptr, DRM_E_ARG, "null rollout");
    const drm_rssm& w = r->rssm;
    DRM_REQUIRE(r->mode >= 0 && r->mode <= 2, DRM_E_ARG, "bad mode");
    DRM_REQUIRE(r->steps >= 1 && r->n >= 1, DRM_E_SHAPE, "steps and n must be positive");
    DRM_REQUIRE(w.A >= 1 && w.S >= 1 && w.D >= 4 && w.Hd >= 4, DRM_E_SHAPE, "bad RSSM sizes");
    DRM_REQUIRE(w.D % 4 == 0 && w.Hd % 4 == 0, DRM_E_SHAPE, "deter and hidden sizes must be multiples of 4");
    DRM_REQUIRE(r->precision == DRM_FP32 || r->precision == DRM_BF16, DRM_E_SHAPE, "unknown precision");
    DRM_REQUIRE(w.W_in && w.b_in && w.W_ih && w.W_hh && w.b_ih && w.b_hh && w.W_p1 && w.b_p1 && w.W_p2 && w.b_p2,
                DRM_E_ARG, "null transition weight");
    DRM_REQUIRE(r->start_stoch && r->start_deter, DRM_E_ARG, "null input");
    // noise: a tensor, or -- fp32 kernels -- generated in the kernel from noise_pair (drm_philox.cuh)
    DRM_REQUIRE(r->eps_prior || r->noise_pair, DRM_E_ARG, "eps_prior is NULL and there is no noise_pair to generate it from");
    DRM_REQUIRE(r->prior_mean && r->prior_std && r->prior_stoch && r->deter, DRM_E_ARG, "null output");
    if (r->mode == DRM_MODE_POLICY) {
        const drm_mlp& a = r->actor;
        DRM_REQUIRE(a.n_linear >= 2 && a.n_linear <= DRM_MAX_LINEAR, DRM_E_SHAPE, "actor needs 2..8 Linear layers");
        DRM_REQUIRE(a.hidden >= 4 && a.hidden % 4 == 0, DRM_E_SHAPE, "actor hidden size must be a multiple of 4");
        DRM_REQUIRE(a.in_dim == w.S + w.D, DRM_E_SHAPE, "actor in_dim != S + D");
        const int aout = r->act_dist == DRM_DIST_TANH_NORMAL ? 2 * w.A : w.A;
        DRM_REQUIRE(r->act_dist >= 0 && r->act_dist <= 2, DRM_E_ARG, "bad action distribution");
        DRM_REQUIRE(a.out_dim == aout, DRM_E_SHAPE, "actor out_dim does not match the action distribution");
        for (int l = 0; l < a.n_linear; ++l) DRM_REQUIRE(a.W[l] && a.b[l], DRM_E_ARG, "null actor weight");
        DRM_REQUIRE((r->noise_act || r->noise_pair) && r->action_out, DRM_E_ARG, "null policy io");
    } else {
        DRM_REQUIRE(r->action_in, DRM_E_ARG, "null action_in");
    }
    if (r->mode == DRM_MODE_REPRESENTATION) {
        DRM_REQUIRE(w.E >= 4 && w.E % 4 == 0, DRM_E_SHAPE, "obs embed size must be a multiple of 4");
        DRM_REQUIRE(w.W_q1 && w.b_q1 && w.W_q2 && w.b_q2, DRM_E_ARG, "null posterior weight");
        DRM_REQUIRE(r->obs_embed && (r->eps_post || r->noise_pair) && r->post_mean && r->post_std && r->post_stoch, DRM_E_ARG,
                    "null representation io");
    }
    return DRM_OK;
}

static void dims(const drm_rollout* r, RollDev& p) {
    const drm_rssm& w = r->rssm;
    p.mode = r->mode; p.T = r->steps; p.n = r->n; p.act_dist = r->act_dist; p.training = r->training;
    p.A = w.A; p.S = w.S; p.D = w.D; p.Hd = w.Hd; p.E = w.E;
    p.A4 = rup(w.A, 4); p.S4 = rup(w.S, 4); p.MS4 = rup(2 * w.S, 4); p.AS4 = rup(w.A + w.S, 4);
    if (r->mode == DRM_MODE_POLICY) {
        p.an = r->actor.n_linear; p.ah = r->actor.hidden; p.Aout = r->actor.out_dim;
    } else {
        p.an = 0; p.ah = 0; p.Aout = 0;
    }
    p.Aout4 = rup(p.Aout > 0 ? p.Aout : 1, 4);
    int U = 4 * w.D;
    if (2 * p.ah > U) U = 2 * p.ah;
    if (w.Hd > U) U = w.Hd;
    p.U = U;
    p.mean_scale = r->act_mean_scale; p.raw_init_std = r->act_raw_init_std; p.act_min_std = r->act_min_std;
    p.temperature = r->act_temperature; p.min_std = r->min_std;
}

struct SaveWidths { int w[DRM_MAX_LINEAR + 8]; int cnt; };

static size_t carve_save(const drm_rollout* r, const RollDev& d, float* base, RollSave* sv) {
    const size_t rows = (size_t)r->steps * r->n;
    size_t off = 0;
    auto take = [&](int width) -> float* {
        float* ptr = base ? base + off : nullptr;
        off += rup_sz(rows * width, 4);
        return ptr;
    };
    RollSave s{};
    if (r->mode == DRM_MODE_POLICY) {
        for (int l = 0; l < d.an - 1; ++l) s.h[l] = take(d.ah);
        s.raw = take(d.Aout);
    }
    s.x = take(d.Hd);
    s.r = take(d.D); s.z = take(d.D); s.nn = take(d.D); s.hn = take(d.D);
    s.hid = take(d.Hd);
    if (r->mode == DRM_MODE_REPRESENTATION) s.hidq = take(d.Hd);
    if (sv) *sv = s;
    return off;
}

static size_t carve_dy(const drm_rollout* r, const RollDev& d, float* base, RollDy* dy) {
    const size_t rows = (size_t)r->steps * r->n;
    size_t off = 0;
    auto take = [&](int width) -> float* {
        float* ptr = base ? base + off : nullptr;
        off += rup_sz(rows * width, 4);
        return ptr;
    };
    RollDy y{};
    y.p2 = take(2 * d.S); y.p1 = take(d.Hd); y.ih = take(3 * d.D); y.hh = take(3 * d.D); y.in = take(d.Hd);
    if (r->mode == DRM_MODE_REPRESENTATION) { y.q2 = take(2 * d.S); y.q1 = take(d.Hd); }
    if (r->mode == DRM_MODE_POLICY) {
        for (int l = 0; l < d.an; ++l) y.a[l] = take(l == d.an - 1 ? d.Aout : d.ah);
    }
    if (dy) *dy = y;
    return off;
}

// Packs the weights (forward or backward layout) at `base` and fills the pointers in p.
static size_t pack_weights(const drm_rollout* r, RollDev& p, float* base, bool backward, PackJobs* jobs_out) {
    const drm_rssm& w = r->rssm;
    const int A = w.A, S = w.S, D = w.D, Hd = w.Hd;
    Packer pk(base);
    if (!backward) {
        p.Win = pk.fwd(w.W_in, A + S, 0, Hd, 0, A);
        pk.fwd(w.W_in, A + S, 0, Hd, A, S);
        p.ld_in = rup(Hd, 32);
        p.Wrz = pk.fwd(w.W_ih, Hd, 0, 2 * D, 0, Hd);
        pk.fwd(w.W_hh, D, 0, 2 * D, 0, D);
        p.ld_rz = rup(2 * D, 32);
        p.Wni = pk.fwd(w.W_ih, Hd, 2 * D, D, 0, Hd);
        p.Wnh = pk.fwd(w.W_hh, D, 2 * D, D, 0, D);
        p.ld_n = rup(D, 32);
        p.Wp1 = pk.fwd(w.W_p1, D, 0, Hd, 0, D); p.ld_p1 = rup(Hd, 32);
        p.Wp2 = pk.fwd(w.W_p2, Hd, 0, 2 * S, 0, Hd); p.ld_p2 = rup(2 * S, 32);
        if (r->mode == DRM_MODE_REPRESENTATION) {
            p.Wq1 = pk.fwd(w.W_q1, D + w.E, 0, Hd, 0, D); p.ld_q1 = rup(Hd, 32);
            p.Wq2 = pk.fwd(w.W_q2, Hd, 0, 2 * S, 0, Hd); p.ld_q2 = rup(2 * S, 32);
        }
        if (r->mode == DRM_MODE_POLICY) {
            const drm_mlp& a = r->actor;
            p.Wa[0] = pk.fwd(a.W[0], S + D, 0, a.hidden, 0, S);
            pk.fwd(a.W[0], S + D, 0, a.hidden, S, D);
            p.ld_a[0] = rup(a.hidden, 32);
            for (int l = 1; l < a.n_linear; ++l) {
                const int nout = l == a.n_linear - 1 ? a.out_dim : a.hidden;
                p.Wa[l] = pk.fwd(a.W[l], a.hidden, 0, nout, 0, a.hidden);
                p.ld_a[l] = rup(nout, 32);
            }
        }
    } else {
        p.Win = pk.bwd(w.W_in, A + S, 0, Hd, 0, A + S); p.ld_in = rup(A + S, 32);
        p.Wih = pk.bwd(w.W_ih, Hd, 0, 3 * D, 0, Hd); p.ld_ih = rup(Hd, 32);
        p.Whh = pk.bwd(w.W_hh, D, 0, 3 * D, 0, D); p.ld_hh = rup(D, 32);
        p.Wp1 = pk.bwd(w.W_p1, D, 0, Hd, 0, D); p.ld_p1 = rup(D, 32);
        p.Wp2 = pk.bwd(w.W_p2, Hd, 0, 2 * S, 0, Hd); p.ld_p2 = rup(Hd, 32);
        if (r->mode == DRM_MODE_REPRESENTATION) {
            p.Wq1 = pk.bwd(w.W_q1, D + w.E, 0, Hd, 0, D); p.ld_q1 = rup(D, 32);
            p.Wq2 = pk.bwd(w.W_q2, Hd, 0, 2 * S, 0, Hd); p.ld_q2 = rup(Hd, 32);
        }
        if (r->mode == DRM_MODE_POLICY) {
            const drm_mlp& a = r->actor;
            for (int l = 1; l < a.n_linear; ++l) {
                const int nout = l == a.n_linear - 1 ? a.out_dim : a.hidden;
                p.Wa[l] = pk.bwd(a.W[l], a.hidden, 0, nout, 0, a.hidden);
                p.ld_a[l] = rup(a.hidden, 32);
            }
        }
    }
    if (jobs_out) *jobs_out = pk.jobs;
    return rup_sz(pk.off, 4);
}

static void fill_io(const drm_rollout* r, RollDev& p) {
    const drm_rssm& w = r->rssm;
    p.b_in = w.b_in; p.b_ih = w.b_ih; p.b_hh = w.b_hh; p.b_p1 = w.b_p1; p.b_p2 = w.b_p2; p.b_q1 = w.b_q1; p.b_q2 = w.b_q2;
    for (int l = 0; l < DRM_MAX_LINEAR; ++l) p.ba[l] = (r->mode == DRM_MODE_POLICY && l < r->actor.n_linear) ? r->actor.b[l] : nullptr;
    p.Wa_last = r->mode == DRM_MODE_POLICY ? r->actor.W[r->actor.n_linear - 1] : nullptr;
    p.start_stoch = r->start_stoch; p.start_deter = r->start_deter; p.action_in = r->action_in;
    p.eps_prior = r->eps_prior; p.eps_post = r->eps_post; p.noise_act = r->noise_act; p.nz = r->noise_pair;
    p.prior_mean = r->prior_mean; p.prior_std = r->prior_std; p.prior_stoch = r->prior_stoch; p.deter = r->deter;
    p.post_mean = r->post_mean; p.post_std = r->post_std; p.post_stoch = r->post_stoch; p.action_out = r->action_out;
}

static size_t fwd_ws_floats(const drm_rollout* r, const RollDev& d) {
    RollDev tmp = d;
    size_t f = pack_weights(r, tmp, nullptr, false, nullptr);
    if (r->mode == DRM_MODE_REPRESENTATION)
        f += rup_sz((size_t)r->steps * r->n * d.Hd, 4) + pack_size(d.E, d.Hd);
    return f;
}
static size_t bwd_ws_floats(const drm_rollout* r, const RollDev& d) {
    RollDev tmp = d;
    size_t f = pack_weights(r, tmp, nullptr, true, nullptr);
    f += carve_dy(r, d, nullptr, nullptr);
    if (r->mode == DRM_MODE_REPRESENTATION) f += pack_size(d.Hd, d.E);
    return f;
}

// The bf16 request is honoured where the tensor-core path exists (imagination rollouts of supported
// shapes); every other rollout runs the fp32 kernels, which satisfy the bf16 tolerance a fortiori.
static bool use_tc(const drm_rollout* r) { return r->precision == DRM_BF16 && tc::rollout_tc_supported(r); }

}  // namespace drm

using namespace drm;

extern "C" {

size_t drm_rollout_save_bytes(const drm_rollout* r) {
    if (!r) return 0;
    if (use_tc(r)) return tc::rollout_tc_save_bytes(r);
    RollDev d{}; dims(r, d);
    return carve_save(r, d, nullptr, nullptr) * sizeof(float) + 64;
}

size_t drm_rollout_workspace_bytes(const drm_rollout* r) {
    if (!r) return 0;
    if (use_tc(r)) {
        const size_t f = tc::rollout_tc_fwd_workspace_bytes(r), b = tc::rollout_tc_bwd_workspace_bytes(r);
        return (f > b ? f : b) + 256;
    }
    RollDev d{}; dims(r, d);
    const size_t f = fwd_ws_floats(r, d), b = bwd_ws_floats(r, d);
    return (f > b ? f : b) * sizeof(float) + 256;
}

int drm_rollout_tiles(const drm_rollout* r) {
    if (!r || validate(r) != DRM_OK) return 0;
    return use_tc(r) ? tc::rollout_tc_tiles(r) : 0;  // 0: fp32 kernels, no per-step signal (flags are set wholesale after the launch)
}

int drm_rollout_fwd(const drm_rollout* r, void* stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    int rc = validate(r); if (rc) return rc;
    rc = check_device(); if (rc) return rc;
    if (use_tc(r)) {
        DRM_REQUIRE(r->eps_prior && r->noise_act, DRM_E_ARG,
                    "the bf16 tensor-core rollout takes materialised noise tensors (drm_noise_fill)");
        return tc::rollout_tc_fwd(r, stream);
    }
    RollDev p{}; dims(r, p);
    DRM_REQUIRE(r->workspace && r->workspace_bytes >= fwd_ws_floats(r, p) * sizeof(float), DRM_E_WORKSPACE,
                "workspace too small");
    if (r->training)
        DRM_REQUIRE(r->save && r->save_bytes >= carve_save(r, p, nullptr, nullptr) * sizeof(float), DRM_E_WORKSPACE,
                    "save buffer too small");
    float* ws = (float*)r->workspace;
    PackJobs jobs;
    size_t off = pack_weights(r, p, ws, false, &jobs);
    rc = launch_pack(jobs, stream); if (rc) return rc;
    fill_io(r, p);
    if (r->training) carve_save(r, p, (float*)r->save, &p.sv);
    if (r->mode == DRM_MODE_REPRESENTATION) {
        // hoisted, non-recurrent half of the posterior's first Linear: obs_embed . W_q1[:, D:]^T for all T*n rows
        float* embproj = ws + off;
        off += rup_sz((size_t)r->steps * r->n * p.Hd, 4);
        rc = launch_linear_nobias(r->obs_embed, p.E, r->steps * r->n, p.E, r->rssm.W_q1, p.D + p.E, p.D, p.Hd, embproj,
                                  ws + off, stream);
        if (rc) return rc;
        p.embproj = embproj;
    }
    rc = launch_rollout_fwd(p, stream); if (rc) return rc;
    // fp32 kernels do not signal per step: every flag passes any target once the whole rollout is in global memory
    if (r->step_flags) DRM_CUDA_OK(cudaMemsetAsync(r->step_flags, 0x7F, sizeof(int32_t) * r->steps, stream));
    return DRM_OK;
}

int drm_rollout_bwd(const drm_rollout* r, const drm_rollout_grad* g, void* stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    int rc = validate(r); if (rc) return rc;
    DRM_REQUIRE(g != nullptr, DRM_E_ARG, "null grad block");
    rc = check_device(); if (rc) return rc;
    DRM_REQUIRE(g->phase >= DRM_PHASE_ALL && g->phase <= DRM_PHASE_RUN, DRM_E_ARG, "bad phase");
    if (use_tc(r)) {
        DRM_REQUIRE(r->eps_prior && r->noise_act, DRM_E_ARG,
                    "the bf16 tensor-core rollout takes materialised noise tensors (drm_noise_fill)");
        return tc::rollout_tc_bwd(r, g, stream);
    }
    if (g->phase == DRM_PHASE_PREPARE) return DRM_OK;  // fp32: everything happens in RUN
    DRM_REQUIRE(!g->g_wait_flags, DRM_E_ARG, "g_wait_flags is honoured by the bf16 policy rollout only");
    RollDev p{}; dims(r, p);
    const drm_rssm& w = r->rssm;
    const int A = w.A, S = w.S, D = w.D, Hd = w.Hd, T = r->steps, n = r->n;
    const bool policy = r->mode == DRM_MODE_POLICY, repr = r->mode == DRM_MODE_REPRESENTATION;
    DRM_REQUIRE(r->workspace && r->workspace_bytes >= bwd_ws_floats(r, p) * sizeof(float), DRM_E_WORKSPACE,
                "workspace too small");
    DRM_REQUIRE(r->save && r->save_bytes >= carve_save(r, p, nullptr, nullptr) * sizeof(float), DRM_E_WORKSPACE,
                "save buffer missing: run drm_rollout_fwd with training=1 first");
    float* ws = (float*)r->workspace;
    PackJobs jobs;
    size_t off = pack_weights(r, p, ws, true, &jobs);
    rc = launch_pack(jobs, stream); if (rc) return rc;
    fill_io(r, p);
    carve_save(r, p, (float*)r->save, &p.sv);
    off += carve_dy(r, p, ws + off, &p.dy);
    p.g_prior_mean = g->g_prior_mean; p.g_prior_std = g->g_prior_std; p.g_prior_stoch = g->g_prior_stoch;
    p.g_deter = g->g_deter; p.g_post_mean = g->g_post_mean; p.g_post_std = g->g_post_std; p.g_post_stoch = g->g_post_stoch;
    p.g_action_out = g->g_action_out;
    p.d_start_stoch = policy ? nullptr : g->d_start_stoch;
    p.d_start_deter = policy ? nullptr : g->d_start_deter;
    p.d_action_in = policy ? nullptr : g->d_action_in;
    p.want_rssm_dy = g->want_rssm_grad ? 1 : 0;
    p.want_actor_dy = (policy && g->want_actor_grad) ? 1 : 0;
    p.want_q1_dy = (repr && (g->want_rssm_grad || g->d_obs_embed)) ? 1 : 0;

    // zero-fill the weight-gradient sinks (accumulated with atomics below)
    const drm_rssm_grad& gw = g->rssm;
    if (g->want_rssm_grad) {
        DRM_REQUIRE(gw.dW_in && gw.db_in && gw.dW_ih && gw.dW_hh && gw.db_ih && gw.db_hh && gw.dW_p1 && gw.db_p1 &&
                        gw.dW_p2 && gw.db_p2, DRM_E_ARG, "null transition gradient sink");
        DRM_CUDA_OK(cudaMemsetAsync(gw.dW_in, 0, sizeof(float) * Hd * (A + S), stream));
        DRM_CUDA_OK(cudaMemsetAsync(gw.db_in, 0, sizeof(float) * Hd, stream));
        DRM_CUDA_OK(cudaMemsetAsync(gw.dW_ih, 0, sizeof(float) * 3 * D * Hd, stream));
        DRM_CUDA_OK(cudaMemsetAsync(gw.dW_hh, 0, sizeof(float) * 3 * D * D, stream));
        DRM_CUDA_OK(cudaMemsetAsync(gw.db_ih, 0, sizeof(float) * 3 * D, stream));
        DRM_CUDA_OK(cudaMemsetAsync(gw.db_hh, 0, sizeof(float) * 3 * D, stream));
        DRM_CUDA_OK(cudaMemsetAsync(gw.dW_p1, 0, sizeof(float) * Hd * D, stream));
        DRM_CUDA_OK(cudaMemsetAsync(gw.db_p1, 0, sizeof(float) * Hd, stream));
        DRM_CUDA_OK(cudaMemsetAsync(gw.dW_p2, 0, sizeof(float) * 2 * S * Hd, stream));
        DRM_CUDA_OK(cudaMemsetAsync(gw.db_p2, 0, sizeof(float) * 2 * S, stream));
        if (repr) {
            DRM_REQUIRE(gw.dW_q1 && gw.db_q1 && gw.dW_q2 && gw.db_q2, DRM_E_ARG, "null posterior gradient sink");
            DRM_CUDA_OK(cudaMemsetAsync(gw.dW_q1, 0, sizeof(float) * Hd * (D + w.E), stream));
            DRM_CUDA_OK(cudaMemsetAsync(gw.db_q1, 0, sizeof(float) * Hd, stream));
            DRM_CUDA_OK(cudaMemsetAsync(gw.dW_q2, 0, sizeof(float) * 2 * S * Hd, stream));
            DRM_CUDA_OK(cudaMemsetAsync(gw.db_q2, 0, sizeof(float) * 2 * S, stream));
        }
    }
    if (p.want_actor_dy) {
        const drm_mlp& a = r->actor;
        for (int l = 0; l < a.n_linear; ++l) {
            const int nout = l == a.n_linear - 1 ? a.out_dim : a.hidden;
            const int nin = l == 0 ? a.in_dim : a.hidden;
            DRM_REQUIRE(g->actor.dW[l] && g->actor.db[l], DRM_E_ARG, "null actor gradient sink");
            DRM_CUDA_OK(cudaMemsetAsync(g->actor.dW[l], 0, sizeof(float) * nout * nin, stream));
            DRM_CUDA_OK(cudaMemsetAsync(g->actor.db[l], 0, sizeof(float) * nout, stream));
        }
    }

    rc = launch_rollout_bwd(p, stream); if (rc) return rc;

    const int M = T * n;           // all (step,row) pairs
    const int M1 = (T - 1) * n;    // pairs whose previous state is a rollout output
    // the state fed INTO step t: start state for t=0, otherwise the carried output of step t-1
    const float* carried_stoch = repr ? r->post_stoch : r->prior_stoch;
    const float* act = policy ? r->action_out : r->action_in;
#define WG(...) do { rc = launch_wgrad(__VA_ARGS__); if (rc) return rc; } while (0)
    if (g->want_rssm_grad) {
        WG(p.dy.p2, 2 * S, p.sv.hid, Hd, M, 2 * S, Hd, gw.dW_p2, Hd, 0, gw.db_p2, stream);
        WG(p.dy.p1, Hd, r->deter, D, M, Hd, D, gw.dW_p1, D, 0, gw.db_p1, stream);
        WG(p.dy.ih, 3 * D, p.sv.x, Hd, M, 3 * D, Hd, gw.dW_ih, Hd, 0, gw.db_ih, stream);
        WG(p.dy.hh, 3 * D, r->start_deter, D, n, 3 * D, D, gw.dW_hh, D, 0, gw.db_hh, stream);
        if (M1 > 0) WG(p.dy.hh + (size_t)n * 3 * D, 3 * D, r->deter, D, M1, 3 * D, D, gw.dW_hh, D, 0, gw.db_hh, stream);
        WG(p.dy.in, Hd, act, A, M, Hd, A, gw.dW_in, A + S, 0, gw.db_in, stream);
        WG(p.dy.in, Hd, r->start_stoch, S, n, Hd, S, gw.dW_in, A + S, A, nullptr, stream);
        if (M1 > 0) WG(p.dy.in + (size_t)n * Hd, Hd, carried_stoch, S, M1, Hd, S, gw.dW_in, A + S, A, nullptr, stream);
        if (repr) {
            WG(p.dy.q2, 2 * S, p.sv.hidq, Hd, M, 2 * S, Hd, gw.dW_q2, Hd, 0, gw.db_q2, stream);
            WG(p.dy.q1, Hd, r->deter, D, M, Hd, D, gw.dW_q1, D + w.E, 0, gw.db_q1, stream);
            WG(p.dy.q1, Hd, r->obs_embed, w.E, M, Hd, w.E, gw.dW_q1, D + w.E, D, nullptr, stream);
        }
    }
    if (repr && g->d_obs_embed) {
        rc = launch_linear_dx(p.dy.q1, M, Hd, w.W_q1, D + w.E, D, w.E, g->d_obs_embed, w.E, ws + off, stream);
        if (rc) return rc;
    }
    if (p.want_actor_dy) {
        const drm_mlp& a = r->actor;
        const int ah = a.hidden;
        // layer 0: input feat = [stoch | deter] of the state fed into the step (detached)
        WG(p.dy.a[0], ah, r->start_stoch, S, n, ah, S, g->actor.dW[0], S + D, 0, g->actor.db[0], stream);
        WG(p.dy.a[0], ah, r->start_deter, D, n, ah, D, g->actor.dW[0], S + D, S, nullptr, stream);
        if (M1 > 0) {
            WG(p.dy.a[0] + (size_t)n * ah, ah, r->prior_stoch, S, M1, ah, S, g->actor.dW[0], S + D, 0, g->actor.db[0], stream);
            WG(p.dy.a[0] + (size_t)n * ah, ah, r->deter, D, M1, ah, D, g->actor.dW[0], S + D, S, nullptr, stream);
        }
        for (int l = 1; l < a.n_linear; ++l) {
            const int nout = l == a.n_linear - 1 ? a.out_dim : ah;
            WG(p.dy.a[l], nout, p.sv.h[l - 1], ah, M, nout, ah, g->actor.dW[l], ah, 0, g->actor.db[l], stream);
        }
    }
#undef WG
    return DRM_OK;
}

}  // extern "C"
