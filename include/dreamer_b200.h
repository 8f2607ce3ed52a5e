/*
 * dreamer_b200.h -- C ABI of libdreamer_b200.so: the RSSM rollout hot path of
 * juliusfrost/dreamer-pytorch as hand-written CUDA for sm_100a (B200).
 *
 * Every entry point is launch-only and asynchronous on the caller's stream, never allocates,
 * frees or retains a pointer past the call, and returns 0 or a negative drm_status.  All
 * pointers are DEVICE pointers to fp32 data unless stated otherwise.  Tensors are contiguous,
 * time-major [time, batch, feature] exactly as the reference lays them out.
 *
 * Reference interfaces replaced (paths relative to the reference checkout):
 *   drm_rollout_fwd / drm_rollout_bwd
 *       DRM_MODE_POLICY          RSSMRollout.rollout_policy          dreamer/models/rnns.py:224-244
 *                                + AgentModel.policy                 dreamer/models/agent.py:101-117
 *                                + ActionDecoder.forward             dreamer/models/action.py:51-66
 *       DRM_MODE_TRANSITION      RSSMRollout.rollout_transition      dreamer/models/rnns.py:207-222
 *                                (T=1: RSSMTransition.forward        dreamer/models/rnns.py:98-107)
 *       DRM_MODE_REPRESENTATION  RSSMRollout.rollout_representation  dreamer/models/rnns.py:179-205
 *                                (T=1: RSSMRepresentation.forward    dreamer/models/rnns.py:149-159)
 *   drm_mlp_fwd / drm_mlp_bwd    DenseModel.forward (the nn.Sequential) dreamer/models/dense.py:28-43
 *                                ActionDecoder.feedforward_model        dreamer/models/action.py:37-52
 *   drm_returns_loss             Dreamer.compute_return + actor/value loss
 *   drm_clip_adam                clip_grad_norm_ + Adam.step of one optimizer group (dreamer_algo.py:182-194)
 *                                dreamer/algos/dreamer_algo.py:314-337,398-422
 *   drm_kl_loss                  KL(post||prior), free nats, entropies
 *                                dreamer/algos/dreamer_algo.py:270-273,342-343
 */
#ifndef DREAMER_B200_H
#define DREAMER_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define DRM_ABI_VERSION 1
#define DRM_MAX_LINEAR 8 /* Linear layers per MLP (reference uses 4) */

typedef enum drm_status {
    DRM_OK = 0,
    DRM_E_SHAPE = -1,     /* unsupported / inconsistent dimensions */
    DRM_E_ALIGN = -2,     /* pointer not 16-byte aligned */
    DRM_E_ARCH = -3,      /* device is not sm_100 */
    DRM_E_WORKSPACE = -4, /* workspace / save buffer too small or NULL */
    DRM_E_CUDA = -5,      /* CUDA runtime error, see drm_last_error() */
    DRM_E_ARG = -6        /* NULL or inconsistent argument */
} drm_status;

typedef enum drm_mode { DRM_MODE_POLICY = 0, DRM_MODE_TRANSITION = 1, DRM_MODE_REPRESENTATION = 2 } drm_mode;
typedef enum drm_act_dist { DRM_DIST_TANH_NORMAL = 0, DRM_DIST_ONE_HOT = 1, DRM_DIST_RELAXED_ONE_HOT = 2 } drm_act_dist;
typedef enum drm_precision { DRM_FP32 = 0, DRM_BF16 = 1 } drm_precision;

/* An MLP: n_linear Linear layers, ELU (alpha=1) between them, none after the last.
 * W[i] is row-major [out_i, in_i] (torch nn.Linear.weight), b[i] is [out_i]. */
typedef struct drm_mlp {
    int32_t n_linear;
    int32_t in_dim;
    int32_t hidden;
    int32_t out_dim;
    const float* W[DRM_MAX_LINEAR];
    const float* b[DRM_MAX_LINEAR];
} drm_mlp;

/* Gradient sinks for an MLP: fresh buffers, OVERWRITTEN (not accumulated). NULL = not wanted. */
typedef struct drm_mlp_grad {
    float* dW[DRM_MAX_LINEAR];
    float* db[DRM_MAX_LINEAR];
} drm_mlp_grad;

/* RSSM weights (rnns.py:72-88,134-139).  A=action, S=stoch, D=deter, Hd=hidden, E=obs embed. */
typedef struct drm_rssm {
    int32_t A, S, D, Hd, E;
    const float* W_in;  /* _rnn_input_model.0.weight        [Hd, A+S]  */
    const float* b_in;  /*                                  [Hd]       */
    const float* W_ih;  /* _cell.weight_ih                  [3D, Hd]   gate order r,z,n */
    const float* W_hh;  /* _cell.weight_hh                  [3D, D]    */
    const float* b_ih;  /*                                  [3D]       */
    const float* b_hh;  /*                                  [3D]       */
    const float* W_p1;  /* _stochastic_prior_model.0.weight [Hd, D]    */
    const float* b_p1;
    const float* W_p2;  /* _stochastic_prior_model.2.weight [2S, Hd]   */
    const float* b_p2;
    const float* W_q1;  /* _stochastic_posterior_model.0.weight [Hd, D+E] (REPRESENTATION only) */
    const float* b_q1;
    const float* W_q2;  /* _stochastic_posterior_model.2.weight [2S, Hd] */
    const float* b_q2;
} drm_rssm;

typedef struct drm_rssm_grad { /* OVERWRITTEN; NULL = not wanted; all-or-nothing per group */
    float *dW_in, *db_in, *dW_ih, *dW_hh, *db_ih, *db_hh, *dW_p1, *db_p1, *dW_p2, *db_p2;
    float *dW_q1, *db_q1, *dW_q2, *db_q2;
} drm_rssm_grad;

/* One rollout of `steps` sequential steps over `n` independent rows. */
typedef struct drm_rollout {
    int32_t mode;      /* drm_mode */
    int32_t steps;     /* T or H */
    int32_t n;         /* rows: B (observation rollout) or N = B*T (imagination) */
    int32_t act_dist;  /* drm_act_dist (POLICY) */
    int32_t precision; /* drm_precision */
    int32_t training;  /* 1: fill `save` for drm_rollout_bwd */
    float act_mean_scale;   /* 5.0      action.py:22,56 */
    float act_raw_init_std; /* ln(e^5-1) action.py:35   */
    float act_min_std;      /* 1e-4     action.py:20,57 */
    float act_temperature;  /* 0.1      action.py:65    */
    float min_std;          /* 0.1      rnns.py:104,155 */
    drm_rssm rssm;
    drm_mlp actor; /* POLICY: in_dim=S+D, out_dim=2A (tanh_normal) or A */
    /* inputs */
    const float* start_stoch; /* [n,S] */
    const float* start_deter; /* [n,D] */
    const float* action_in;   /* [steps,n,A]  TRANSITION / REPRESENTATION */
    const float* obs_embed;   /* [steps,n,E]  REPRESENTATION */
    const float* eps_prior;   /* [steps,n,S]  N(0,1) */
    const float* eps_post;    /* [steps,n,S]  REPRESENTATION */
    const float* noise_act;   /* [steps,n,A]  POLICY: N(0,1) (tanh_normal) or Gumbel(0,1) */
    /* outputs */
    float* prior_mean;  /* [steps,n,S] */
    float* prior_std;   /* [steps,n,S] */
    float* prior_stoch; /* [steps,n,S] */
    float* deter;       /* [steps,n,D]  shared by prior and posterior */
    float* post_mean;   /* [steps,n,S]  REPRESENTATION */
    float* post_std;
    float* post_stoch;
    float* action_out;  /* [steps,n,A]  POLICY */
    /* scratch owned by the caller */
    void* save;         /* drm_rollout_save_bytes(); written by fwd when training, read by bwd */
    size_t save_bytes;
    void* workspace;    /* drm_rollout_workspace_bytes(); contents undefined after the call */
    size_t workspace_bytes;
    /* optional (POLICY, may be NULL): int32 [steps], zeroed by the caller -- and the zero-fill must be stream-ordered
       BEFORE the launch of every consumer that polls the flags (a consumer on another stream waits on an event
       recorded after the fill), or it can see the previous pass' counts.  The forward adds 1 to step_flags[t]
       (device-scope release) whenever one row tile has stored every output of step t, so that a consumer kernel
       running BESIDE the rollout (drm_mlp_call.wait_flags) can start on step t while step t+1 is computed.
       drm_rollout_tiles() is the count each flag reaches. */
    int32_t* step_flags;
    /* Production noise (last sentence; the reference draws torch.randn per step, rnns.py:104,155 /
       action.py:58-65).  Device pointer to uint64 {seed, offset} (drm_noise_next), or NULL.  When it is set, any of
       eps_prior / eps_post / noise_act may be NULL: the fp32 kernels then GENERATE that value in the kernel -- Philox4x32-10
       keyed on (seed, offset, stream, element index), forward and again backward, nothing goes through HBM.  The value of
       element e is exactly what drm_noise_fill writes at index e for the same pair and stream (0 eps_prior, 1 eps_post,
       2 noise_act; noise_act is Gumbel(0,1) for the categorical policies, N(0,1) otherwise).  The bf16 tensor-core
       kernels take materialised tensors only (DRM_E_ARG otherwise): fill them with drm_noise_fill. */
    const unsigned long long* noise_pair;
} drm_rollout;

/* Backward calls in two halves.  A backward pass first re-packs the (unchanged) weights for the transposed products and
 * zero-fills the weight-gradient sinks -- work that depends on nothing the forward pass computes.  PREPARE issues only
 * that part (the caller may put it on another stream, beside the forward pass); RUN issues the rest and requires a
 * completed PREPARE with the same workspace, weights and sinks.  ALL (0, the default) does both.  The fp32 kernels
 * treat PREPARE as a no-op and RUN as ALL. */
#define DRM_PHASE_ALL 0
#define DRM_PHASE_PREPARE 1
#define DRM_PHASE_RUN 2

/* Upstream gradients (NULL = zero) and gradient sinks (NULL = not wanted) of a rollout. */
typedef struct drm_rollout_grad {
    const float *g_prior_mean, *g_prior_std, *g_prior_stoch, *g_deter; /* [steps,n,.] */
    const float *g_post_mean, *g_post_std, *g_post_stoch;
    const float* g_action_out;
    float* d_start_stoch; /* [n,S]  (NULL for POLICY: the start state is detached, rnns.py:236) */
    float* d_start_deter; /* [n,D] */
    float* d_action_in;   /* [steps,n,A] */
    float* d_obs_embed;   /* [steps,n,E] */
    drm_rssm_grad rssm;
    int32_t want_rssm_grad;  /* 0: weights frozen (FreezeParameters), skip all RSSM weight grads */
    int32_t want_actor_grad;
    drm_mlp_grad actor;
    int32_t phase;   /* DRM_PHASE_*: lets the weight packing + zero-filling of the sinks be issued early (another stream) */
    /* bf16 POLICY path, optional: g_* rows of step t are read only once g_wait_flags[t] >= g_wait_target (the
     * step_rows_done counters of the head kernels producing them on other streams, target = n). */
    int32_t g_wait_target;
    const int32_t* g_wait_flags;
    /* bf16 POLICY path, phase RUN only, optional: a second CUDA stream for the actor weight-gradient GEMMs.  They then follow
     * the rollout step by step (device counters in the workspace) instead of starting when it has finished.  The caller
     * orders that stream after the PREPARE call, keeps the SMs the rollout does not use free for it (anything else queued on
     * it runs first), and joins it afterwards. */
    void* wgrad_stream;
} drm_rollout_grad;

/* Batched MLP over M rows; the input is the concatenation of two column segments
 * (e.g. stoch [M,S] and deter [M,D]) so that `feat` never has to be materialised. */
typedef struct drm_mlp_call {
    drm_mlp mlp;
    int32_t M;
    int32_t k0, k1;            /* k0 + k1 == mlp.in_dim; k1 may be 0 */
    int32_t precision;
    int32_t training;          /* fwd: keep hidden activations in `save` */
    const float* x0; int64_t ld0; /* [M,k0], row stride ld0 elements */
    const float* x1; int64_t ld1;
    float* y;                  /* [M,out_dim] */
    void* save; size_t save_bytes;           /* drm_mlp_save_bytes() */
    void* workspace; size_t workspace_bytes; /* drm_mlp_workspace_bytes() */
    /* optional producer/consumer gating (NULL = none): rows [t*wait_rows, (t+1)*wait_rows) of x0/x1 may be read once
       wait_flags[t] >= wait_target (acquire).  The producer must already be running or be guaranteed SMs:
       max_ctas (0 = all SMs) bounds this call's grid so that both kernels stay resident. */
    const int32_t* wait_flags; int32_t wait_target; int32_t wait_rows; int32_t wait_steps; int32_t max_ctas;

} drm_mlp_call;

typedef struct drm_mlp_bwd_call {
    drm_mlp_call fwd;          /* same description as the forward call (y unused) */
    const float* dy;           /* [M,out_dim] */
    const float* dy_scale;     /* device scalar multiplied into dy, or NULL (=1) */
    float* dx0; float* dx1;    /* [M,k0] / [M,k1] with the forward's strides; NULL = not wanted */
    int32_t accumulate_dx;     /* 1: dx += , 0: dx = */
    int32_t want_wgrad;
    drm_mlp_grad grad;
    int32_t phase;   /* DRM_PHASE_* */
    /* Kernels running beside each other (bf16 path only; all optional).  The data-gradient kernels of several heads that
     * accumulate into the same dx, and the reverse-time rollout that consumes dx step by step, can be launched on
     * different streams and hand rows over through device counters instead of stream order:
     *   reverse_tiles   process the 128-row tiles from the last to the first (the rollout backward wants step T-1 first)
     *   tile_done[i]    set to 1 once rows [128 i, 128 i + 128) of dx are final for this call            (zeroed by the caller)
     *   tile_wait[i]    spin until non-zero before touching those dx rows (the previous head's tile_done)
     *   step_rows_done  [step_count] counters, += rows of step s (s = row / step_rows) finished            (zeroed by the caller)
     * fwd.max_ctas bounds the grid; the caller must keep all gated kernels resident together. */
    int32_t reverse_tiles;
    int32_t* tile_done;
    const int32_t* tile_wait;
    int32_t* step_rows_done;
    int32_t step_rows, step_count;
} drm_mlp_bwd_call;

/* lambda-returns + actor loss + value loss over imagined heads, forward AND the gradients of
 * both scalar losses w.r.t. the head outputs (dreamer_algo.py:314-337,398-422). */
typedef struct drm_returns_loss_call {
    int32_t H, N;
    float discount, lambda_;
    const float* reward;  /* [H,N] reward head mean */
    const float* value;   /* [H,N] value head mean  */
    const float* pcont;   /* [H,N] pcont logits or NULL (constant discount) */
    float* returns;       /* [H-1,N] */
    float* disc;          /* [H-1,N] cumulative discount */
    float* losses;        /* [2]: actor_loss, value_loss  (must be zero-filled by the caller) */
    float* d_reward;      /* [H,N] d actor_loss / d reward */
    float* d_value;       /* [H,N] d actor_loss / d value  */
    float* d_pcont;       /* [H,N] d actor_loss / d pcont logits, or NULL */
    float* d_value_pred;  /* [H,N] d value_loss / d value_pred (row H-1 zero) */
} drm_returns_loss_call;

/* KL(post || prior) mean, free-nats clamp and entropies, forward and gradient. */
typedef struct drm_kl_call {
    int32_t M, S;         /* M = T*B rows */
    float free_nats;
    const float *post_mean, *post_std, *prior_mean, *prior_std; /* [M,S] */
    float* out;           /* [4]: div (clamped), raw mean KL, prior entropy, post entropy; zero-filled by caller */
    /* gradient of the clamped div (unit upstream), written by drm_kl_bwd; NULL = not wanted */
    float *d_post_mean, *d_post_std, *d_prior_mean, *d_prior_std;
} drm_kl_call;

/* Image log-likelihood of the observation decoder: Independent(Normal(mean, 1), 3).log_prob(obs)
 * (dreamer/models/observation.py:117 as consumed by dreamer_algo.py:264), one fused reduction per image:
 *     logp[row] = -0.5 * sum_i (obs[row,i] - mean[row,i])^2 - D * 0.5 ln(2 pi)
 * and its gradient  dmean[row,i] = g[row] * (obs[row,i] - mean[row,i]).  HBM-bound. */
typedef struct drm_image_logprob_call {
    int64_t rows;          /* images (T*B) */
    int32_t D;             /* elements per image (C*H*W), multiple of 4 */
    int32_t mean_bf16;     /* 1: mean / dmean are bf16 (an autocast decoder), 0: fp32 */
    const void* mean;      /* [rows, D] */
    const float* obs;      /* [rows, D] fp32, already normalised */
    float* logp;           /* [rows]   (fwd) */
    const float* g;        /* [rows]   upstream gradient of logp (bwd) */
    void* dmean;           /* [rows, D] same dtype as mean (bwd) */
} drm_image_logprob_call;
/* Pointwise glue of the conv stack (dreamer/models/observation.py:18-25,83-93: Conv2d / ConvTranspose2d + ReLU) on
 * channels-last fp32 activations viewed as [rows = N*H*W, C]; the convolution itself is called without bias (cuDNN).
 *   fwd: x = act(x + bias[c]) in place                     (replaces torch's broadcast add + ReLU: two passes -> one)
 *   bwd: g_pre = g * (y > 0),  dbias[c] = sum_rows g_pre    (replaces threshold_backward + the bias-gradient reduction)
 * C must be a multiple of 4 whose quarter divides 256 (the reference's 32 / 64 / 128 / 256 channels). */
typedef struct drm_bias_act_call {
    int64_t rows;
    int32_t C;
    int32_t relu;       /* 1: ReLU, 0: bias only (then bwd writes no g_pre, only dbias) */
    float* x;           /* fwd: [rows, C] in place */
    const float* bias;  /* fwd: [C] */
    const float* g;     /* bwd: upstream gradient [rows, C] */
    const float* y;     /* bwd: forward output [rows, C] (relu = 1) */
    float* g_pre;       /* bwd: out [rows, C] (relu = 1) */
    float* dbias;       /* bwd: out [C] or NULL; zero-filled by the call */
} drm_bias_act_call;
int drm_bias_act_fwd(const drm_bias_act_call* c, void* stream);
int drm_bias_act_bwd(const drm_bias_act_call* c, void* stream);
int drm_image_logprob_fwd(const drm_image_logprob_call* c, void* stream);
int drm_image_logprob_bwd(const drm_image_logprob_call* c, void* stream);

/* Environment-time policy head: SampleDist.mode() of the tanh-Normal action distribution
 * (dreamer/models/distribution.py:24-35 over dreamer/models/action.py:54-61), one launch. */
typedef struct drm_policy_mode_call {
    int32_t n, A, samples;    /* rows, action dims, Monte-Carlo draws (the reference uses 100) */
    float mean_scale, raw_init_std, min_std; /* 5, ln(e^5 - 1), 1e-4 (action.py:20-22,35) */
    const float* raw;    /* [n, 2A] output of the actor MLP (mean | std pre-activations) */
    const float* noise;  /* [samples, n, A] N(0,1) */
    float* action;       /* [n, A] the draw with the highest log-probability */
    float* logp;         /* [samples, n] or NULL */
    int32_t* index;      /* [n] winning draw, or NULL */
} drm_policy_mode_call;
int drm_policy_mode(const drm_policy_mode_call* c, void* stream);

/* Replay -> device: sequence gather on a ring buffer resident in HBM (rlpyt/replays/sequence/n_step.py:68-106,
 * rlpyt/utils/misc.py:38-56), dst[j, i, :] = src[(t0[i] + t_offset + j) mod T_cap, b[i], :]. */
typedef struct drm_replay_gather_call {
    int64_t T_cap;       /* time length of the ring */
    int32_t B_env;       /* environments (second ring dimension) */
    int32_t T_seq;       /* sequence length to extract */
    int32_t batch;       /* sequences */
    int32_t t_offset;    /* 0, or -1 for "previous" fields (all_action, all_reward) */
    int64_t row_bytes;   /* bytes of one [t, b] entry */
    const void* src;     /* [T_cap, B_env, row_bytes] */
    void* dst;           /* [T_seq, batch, row_bytes] */
    const int64_t* t0;   /* [batch] start times (device) */
    const int64_t* b;    /* [batch] environment indices (device) */
} drm_replay_gather_call;
int drm_replay_gather(const drm_replay_gather_call* c, void* stream);

/* Production noise: counter-based N(0,1) / Gumbel(0,1) streams (Philox4x32-10) keyed on (seed, offset, stream, element).
 * Replaces the reference's per-step torch.randn / OneHotCategorical.sample draws (rnns.py:104,155, action.py:58-65,
 * distribution.py:27) with one launch per rollout -- or none, see drm_rollout.noise_pair.
 *   state: device uint64[4] {seed, offset, ticket (0), unused}, one per device, owned by the caller.
 *   drm_noise_next   pair_out = {seed, offset}; offset += 1.  One tiny launch, recordable into a CUDA graph (every replay
 *                    then draws fresh noise without the host).
 *   drm_noise_fill   writes up to 4 blocks (e.g. noise_act and eps_prior of one imagination pass) in ONE launch, keyed
 *                    either on a frozen `pair` or on `state` -- then the offset is bumped by the last CTA to finish and
 *                    the {seed, offset} used is stored to pair_out (may be NULL). */
#define DRM_MAX_NOISE_BLOCKS 4
#define DRM_NOISE_NORMAL 0
#define DRM_NOISE_GUMBEL 1
typedef struct drm_noise_fill_call {
    unsigned long long* state;        /* exactly one of state / pair */
    const unsigned long long* pair;
    unsigned long long* pair_out;     /* with `state`: receives the pair used, or NULL */
    int32_t n_blocks;
    int32_t pad_;
    float* out[DRM_MAX_NOISE_BLOCKS];
    int64_t numel[DRM_MAX_NOISE_BLOCKS];
    int32_t stream_id[DRM_MAX_NOISE_BLOCKS]; /* 0 eps_prior, 1 eps_post, 2 noise_act, 3 SampleDist.mode draws, .. 255 */
    int32_t kind[DRM_MAX_NOISE_BLOCKS];      /* DRM_NOISE_* */
} drm_noise_fill_call;
int drm_noise_next(unsigned long long* state, unsigned long long* pair_out, void* stream);
int drm_noise_fill(const drm_noise_fill_call* c, void* stream);
/* Diagnostics / known-answer tests: the raw Philox words of groups [group0, group0 + n_groups) of a stream -> out[4 n]. */
int drm_noise_words(const unsigned long long* pair, int stream_id, unsigned long long group0, int n_groups,
                    unsigned int* out, void* stream);

/* ---- multi-GPU: one-shot all-reduce (mean) of the flat gradient bucket over NVLink peer memory ------------
 * The reference is single-GPU; a data-parallel run of it would average the gradients of dreamer_algo.py:174-180 with
 * torch's NCCL all-reduce.  Here the backward kernels write the actor / value weight gradients straight into one flat
 * fp32 bucket in symmetric (peer-mapped) memory and ONE kernel averages it in place with P2P loads: two flag
 * hand-shakes + one pass of 16-byte loads (one-shot; from 4 ranks on reduce-scatter + all-gather over the same
 * loads, three hand-shakes), recordable into the step's CUDA graph.  All ranks call it with the same
 * numel / ctas, in the same order, each on its own device.  Results are bit-identical on every rank (rank-order sums).
 *   buf[r]      rank r's bucket as mapped into THIS process (buf[rank] = the local one), 16-byte aligned
 *   signal[r]   rank r's signal pad (uint32 [3][ctas][DRM_MAX_RANKS], zero-filled once before the first call)
 *   staging     local scratch of numel floats;  epoch: local uint64[2] = {0, 0} before the first call */
#define DRM_MAX_RANKS 8
#define DRM_AR_MAX_CTAS 64
typedef struct drm_allreduce_call {
    int32_t world, rank;
    int32_t ctas;      /* CTAs per rank (same on every rank) */
    float scale;       /* 1 / world for the mean */
    int64_t numel;     /* floats, multiple of 4 */
    int32_t algo;      /* 0 auto, 1 one-shot, 2 two-shot (reduce-scatter + all-gather; needs world >= 3).  Auto picks two-shot
                          for world >= 4 and buckets >= 4 MB: measured at 8 GPUs on the 1 MB actor+value bucket, the
                          third hand-shake costs more than the 7 MB of peer reads it saves (1.153 vs 1.140 ms per step) */
    int32_t pad_;
    void* buf[DRM_MAX_RANKS];
    void* signal[DRM_MAX_RANKS];
    float* staging;
    unsigned long long* epoch;
} drm_allreduce_call;
int drm_allreduce(const drm_allreduce_call* c, void* stream);

int drm_version(void);
/* Once per device, before the first compute call (the ctypes binding does it at load time): checks the architecture and
 * loads every kernel of the library (CUDA would otherwise load a kernel lazily at its first launch, and that load may
 * wait for the device to go idle -- which a kernel spinning on the counters of a producer never lets happen). */
int drm_init(void);
/* SMs of the current device: the co-residency budget of kernels that run beside each other (max_ctas, wgrad_stream). */
int drm_num_sms(void);
const char* drm_last_error(void);
size_t drm_sizeof(const char* struct_name); /* ABI self-check for FFI bindings */

/* Diagnostics.  drm_launch_count: kernels launched by this library since load.
 * drm_prof_enable(1) resets and starts bracketing every kernel launch with a cudaEvent pair on
 * the launching stream; drm_prof_read sums the elapsed time of one slot (it synchronises on the
 * recorded events).  Slots: 0 weight pack, 1 rollout fwd, 2 rollout bwd, 3 MLP fwd, 4 MLP bwd,
 * 5 weight-gradient GEMM, 6 loss kernels. */
unsigned long long drm_launch_count(void);
void drm_prof_enable(int on);
int drm_prof_read(int slot, double* total_ms, int* count);

size_t drm_rollout_save_bytes(const drm_rollout* r);
size_t drm_rollout_workspace_bytes(const drm_rollout* r);
/* Row tiles the forward of `r` works on (what every step_flags entry counts up to); 0 when this call runs on the
   fp32 kernels, which publish no per-step progress (every flag is then set past any target after the launch). */
int drm_rollout_tiles(const drm_rollout* r);
int drm_rollout_fwd(const drm_rollout* r, void* stream);
int drm_rollout_bwd(const drm_rollout* r, const drm_rollout_grad* g, void* stream);

size_t drm_mlp_save_bytes(const drm_mlp_call* c);
size_t drm_mlp_workspace_bytes(const drm_mlp_call* c);
int drm_mlp_fwd(const drm_mlp_call* c, void* stream);
int drm_mlp_bwd(const drm_mlp_bwd_call* c, void* stream);

/* drm_mlp_fwd_noalloc -- NO COUNTERPART IN THE REFERENCE; added for callers that
 * cannot call cudaMalloc (code under CUDA-graph capture, tight serving loops, harnesses that forbid it).
 * Same result as drm_mlp_fwd for an fp32 inference forward (precision = DRM_FP32, training = 0, wait_flags = NULL), but
 * c->workspace / c->save are IGNORED: the packed-weight scratch is a static __device__ array of
 * DRM_NOALLOC_SCRATCH_BYTES inside the library (held by every GPU that loads it, until exit).  The call allocates
 * nothing (no cudaMalloc / cudaFree / allocating library routine).  Returns DRM_E_WORKSPACE, with the sizes in
 * drm_last_error(), when drm_mlp_noalloc_scratch_bytes(c) exceeds the limit (the reference's widest head,
 * 230 -> 3 x 300 -> 1, needs 1.05 MiB).  All calls on a device share the array: calls that overlap on the GPU (different
 * streams or threads) race -- keep them in one stream order.  Nothing has to be cleared between calls (the packing
 * launch rewrites every float the MLP kernel reads).  drm_prof_enable(1) makes the diagnostics create CUDA events;
 * leave it off where event creation is forbidden too. */
#define DRM_NOALLOC_SCRATCH_BYTES (2u << 20)
size_t drm_mlp_noalloc_scratch_bytes(const drm_mlp_call* c);
int drm_mlp_fwd_noalloc(const drm_mlp_call* c, void* stream);

int drm_returns_loss(const drm_returns_loss_call* c, void* stream);

/* ---- optimizer tail of one parameter group: clip_grad_norm_ + Adam.step ---------------------------
 * Replaces, for one optimizer group, dreamer_algo.py:182-194:
 *     grad_norm = torch.nn.utils.clip_grad_norm_(params, grad_clip)      (torch/nn/utils/clip_grad.py)
 *     optimizer.step()                                                    (torch/optim/adam.py, foreach path,
 *                                                                          amsgrad=False, weight_decay=0, maximize=False)
 * in two launches (sum of squares; clip + update) instead of ~16, without a host sync.  Arithmetic, per element:
 *     g *= min(1, max_norm / (total_norm + 1e-6));  m += (1-beta1) (g - m);  v = beta2 v + (1-beta2) g g;
 *     p -= step_size * m / (sqrt(v) / bias2_sqrt + eps),   step_size = lr / (1 - beta1^step), bias2_sqrt = sqrt(1 - beta2^step)
 * (step_size and bias2_sqrt are computed by the caller in double, as torch does).  Gradients are scaled in place, like
 * clip_grad_norm_ does.  norm_out receives the total 2-norm BEFORE clipping (clip_grad_norm_'s return value). */
#define DRM_MAX_OPT_TENSORS 64
typedef struct drm_clip_adam_call {
    int32_t n_tensors;
    int32_t pad_;
    float max_norm;   /* <= 0: no clipping (the norm is still reported) */
    float beta1, beta2, eps;
    float one_minus_beta1, one_minus_beta2;  /* computed in double by the host (1 - 0.999f in fp32 is off by 5e-5) */
    float step_size, bias2_sqrt;
    /* Device-side step counter (optional, for CUDA-graph capture where host scalars would be frozen): when step_dev is
     * non-NULL the call increments *step_dev and derives step_size = lr / (1 - beta1^t) and bias2_sqrt = sqrt(1 - beta2^t)
     * from it on the device (in double, from beta1_d / beta2_d); step_size / bias2_sqrt above are then ignored. */
    float lr;
    int32_t pad2_;
    double beta1_d, beta2_d;
    int32_t* step_dev;
    float* param[DRM_MAX_OPT_TENSORS];
    float* grad[DRM_MAX_OPT_TENSORS];
    float* exp_avg[DRM_MAX_OPT_TENSORS];
    float* exp_avg_sq[DRM_MAX_OPT_TENSORS];
    int64_t numel[DRM_MAX_OPT_TENSORS];
    float* norm_out;   /* device scalar */
    void* workspace;   /* >= 16 bytes: sum-of-squares accumulator */
    size_t workspace_bytes;
    /* Duplicated parameters.  The reference's world-model group lists the transition parameters TWICE (the transition
     * module is also a sub-module of the representation module, dreamer_algo.py:112-118 / agent.py:57-68; torch warns).
     * torch then -- in its sequential semantics (CPU, foreach=False) -- counts such a gradient twice in the norm, scales
     * it by the clip coefficient twice, and runs two Adam updates on it per step, its per-parameter step count advancing
     * by two.  mult[i] (0 or 1: listed once) is the number of listings of tensor i; for mult > 1 the bias corrections
     * are derived on the device in double from the update count (*step_dev, or `update` when step_dev is NULL):
     * steps (update - 1) * mult + 1 .. update * mult. */
    int32_t mult[DRM_MAX_OPT_TENSORS];
    int32_t update;    /* 1-based count of this optimizer step (host side), used when step_dev is NULL */
    int32_t pad3_;
} drm_clip_adam_call;
int drm_clip_adam(const drm_clip_adam_call* c, void* stream);
int drm_kl_fwd(const drm_kl_call* c, void* stream);
int drm_kl_bwd(const drm_kl_call* c, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* DREAMER_B200_H */
