/*
 * nonmyopia_b200 - C ABI of the B200-native lookahead-acquisition hot path.
 *
 * The reference (sangttruong/nonmyopia, synthetic_experiment/) is pure Python and
 * has no FFI of its own; its de-facto boundary is the Python class API of
 * acqfs.py / actor.py / amortized_network.py.  The host-side mirror
 * of those classes lives in nonmyopia_b200/{acqfs,actor,amortized_network}.py and
 * reaches the GPU only through the entry points declared here (ctypes binding in
 * nonmyopia_b200/_lib.py + torch.autograd.Function nodes; there is no C++ torch.library
 * shim).  Each entry point names the reference code it replaces.
 *
 * Conventions
 *  - every pointer is a DEVICE pointer unless the name says host; tensors are dense
 *    row-major; `dtype` is NMB_F64 or NMB_F32 and applies to every floating tensor
 *    of that call;
 *  - `stream` is a cudaStream_t passed as void* (0 = legacy default stream);
 *  - functions are asynchronous on `stream`, never allocate or free device memory
 *    (workspaces are caller-provided; nmb_*_workspace_bytes tell the size) and
 *    return 0 on success, a negative NMB_E* code for bad arguments, or a positive
 *    cudaError_t; nmb_last_error() gives the message for the calling thread.
 *  - no CPU fallback exists: without a CUDA device every compute entry point fails.
 */
#ifndef NONMYOPIA_B200_H
#define NONMYOPIA_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define NMB_F64 0
#define NMB_F32 1

/* stationary kernels of the GP surrogate (reference main.py:86-92) */
#define NMB_KERNEL_RBF      0
#define NMB_KERNEL_MATERN12 1
#define NMB_KERNEL_MATERN32 2
#define NMB_KERNEL_MATERN52 3

#define NMB_EINVAL  (-1)   /* bad argument (shape, dtype, null pointer) */
#define NMB_ELIMIT  (-2)   /* a size exceeds what the kernels were built for */
#define NMB_ENODEV  (-3)   /* no CUDA device */

#define NMB_MAX_DIM 16     /* input dimension d <= 16 (reference envs: 2,4,6,8) */

const char* nmb_last_error(void);
int         nmb_version(void);
/* number of SMs of the current device (grid sizing), <0 on error */
int         nmb_device_sm_count(void);

/* ------------------------------------------------------------------------
 * Thompson-sample path  f(x) = c + feat_scale * [sin(z), cos(z)] . w_s + k(x, X_tr) . v_s
 * with z = W (x / lengthscale).  Replaces the BoTorch MatheronPath call
 * `self.f(X)` at reference acqfs.py:362 and :409-411.
 * The host prepares the constants once per BO step (dump_model, acqfs.py:258-266).
 * ---------------------------------------------------------------------- */
typedef struct nmb_path {
    int32_t d;            /* input dim */
    int32_t n_freq;       /* M/2 spectral frequencies */
    int32_t n_train;      /* n GP training points */
    int32_t n_paths;      /* S jointly drawn paths */
    int32_t kernel_id;    /* NMB_KERNEL_* */
    int32_t _pad;
    double  feat_scale;   /* sqrt(outputscale * 2 / M) */
    double  amplitude;    /* outputscale (1 for a bare kernel) */
    double  mean_const;   /* constant mean c */
    const void* w_phase;  /* [n_freq, d]  W[m,k] / (lengthscale[k] * pi): phase in half turns */
    const void* w_feat;   /* [n_paths, 2*n_freq] feature weights (sin block, then cos block) */
    const void* inv_ls;   /* [d] 1 / lengthscale (isotropic kernels repeat the value) */
    const void* train_x;  /* [n_train, d] */
    const void* v_update; /* [n_paths, n_train] pathwise update weights K^-1 (y - prior - eps) */
} nmb_path_t;

/* y[i] = f_{s(i)}(x[i]) and dydx[i,:] = grad_x f (either output may be NULL).
 * Point i uses path s(i) = i / points_per_path (points_per_path = N for one path).
 * The backward of the op is gx = gy * dydx (elementwise), so one launch serves
 * forward and backward. */
int nmb_path_eval(const nmb_path_t* path, int dtype, const void* x, int64_t n_points,
                  int64_t points_per_path, void* y, void* dydx, void* stream);

/* ------------------------------------------------------------------------
 * Dense kernel matrix  K[i][j] = amplitude * kappa(x1_i, x2_j) (+ diag_add where i == j),
 * x1 [n1][d], x2 [n2][d], K [n1][ldk].  Replaces the GPyTorch kernel evaluations behind
 * SingleTaskGP(...).posterior and draw_matheron_paths (reference main.py:479-484,
 * acqfs.py:264,366,414).  diag_add = noise builds K + sigma^2 I.
 * ---------------------------------------------------------------------- */
int nmb_kernel_matrix(int dtype, int kernel_id, int d, const void* x1, int64_t n1, const void* x2, int64_t n2,
                      const void* inv_ls, double amplitude, double diag_add, void* K, int64_t ldk, void* stream);

/* ------------------------------------------------------------------------
 * Amortized policy network (reference amortized_network.py:15-139, continuous branch):
 *   Linear(d+1,H) ELU Linear(H,H) ELU -> GRUCell(H,H) -> cat -> Linear(2H,H) ELU
 *   Linear(H,H) ELU Linear(H,d*n_a) -> sigmoid [-> random planar rotation about x]
 * Parameters travel as ONE flat buffer in the reference's state_dict order
 * (embedding.{weight,bias}, prepro.2.{weight,bias}, rnn.{weight_ih,weight_hh,bias_ih,
 * bias_hh}, postpro_A.{0,3,6}.{weight,bias}, postpro_X.{0,3,6}.{weight,bias}), torch
 * [out][in] layout ("theta"), plus a copy with every weight matrix transposed to
 * [in][out] ("theta_t") that the forward GEMMs stream.  H <= 64, H % 4 == 0, n_a == 1.
 * ---------------------------------------------------------------------- */
int64_t nmb_policy_param_count(int d, int H, int n_actions);
/* theta_t <- per-matrix transpose of theta (biases copied) */
int nmb_policy_transpose(int dtype, int d, int H, int n_actions, const void* theta, void* theta_t, void* stream);

/* spotlight cost (reference acqfs.py:583-633; presets main.py:219-237) */
typedef struct nmb_cost {
    double radius;             /* the host passes the float32-rounded radius, as the reference's buffer holds it */
    double k, max_noise, p_norm, discount, discount_threshold;
    double prev_cost;          /* cost already spent before this BO step (actor.py:482-486) */
} nmb_cost_t;

/* ------------------------------------------------------------------------
 * Accumulated spotlight cost along every root-to-leaf chain of a lookahead tree: the L + n_a
 * qCostFunction calls of reference acqfs.py:421-441 (free-parameter HES-TS, HES-MC and the
 * baselines; the amortized rollout has the same arithmetic fused in).
 * Level s holds counts[s] = n_r * prod(nf[:s]) points, restart index fastest (the reference's
 * [*nf_rev[:s], n_r, 1, d] flattened), stored back to back in `pts`; node i of level s has
 * parent i % counts[s-1]; level-0 node i moves away from root[i % n_root].
 *   fwd: acc[leaf] = sum over the chain of  mult * (max(k (|move|_p - r), 0) + noise * max_noise);
 *        noise [sum counts] (one N(0,1) draw per node, NULL = 0), mult [sum counts] (out, saved).
 *   bwd: g_pts [sum counts][d], g_root [n_root][d] (optional) from g_acc [counts[n_levels-1]];
 *        deterministic (no atomics).
 * ---------------------------------------------------------------------- */
int nmb_tree_cost_fwd(int dtype, int d, int n_levels, const int64_t* counts_host, const nmb_cost_t* cost,
                      const void* root, int64_t n_root, const void* pts, const void* noise, void* acc, void* mult,
                      void* stream);
int nmb_tree_cost_bwd(int dtype, int d, int n_levels, const int64_t* counts_host, const nmb_cost_t* cost,
                      const void* root, int64_t n_root, const void* pts, const void* mult, const void* g_acc,
                      void* g_pts, void* g_root, void* stream);

/* ------------------------------------------------------------------------
 * Fused HES-TS-AM rollout: for every chain (restart x path) run L design steps and
 * the action step  policy -> [rotation] -> Thompson path -> spotlight cost, i.e. the
 * whole of qMultiStepHEntropySearch.forward with enable_ts and an amortized map
 * (reference acqfs.py:277-456 with :361-362, :408-411; SURVEY App. C dataflow).
 * ---------------------------------------------------------------------- */
typedef struct nmb_ts_problem {
    int32_t dtype, d, H, L;
    int64_t n_chains;          /* N = restarts x paths */
    int64_t chains_per_path;   /* chain i evaluates path i / chains_per_path */
    const nmb_path_t* path;    /* host pointer */
    nmb_cost_t cost;
    const void* prev_x;        /* [N, d]  last observed design, expanded (actor.py:458-460) */
    const void* prev_y;        /* [N]     */
    const void* prev_h;        /* [N, H]  hidden state + hidden_noise (actor.py:472-481) */
    uint64_t seed;             /* in-kernel RNG (used only where rot_q / cost_noise are NULL) */
    const int64_t* step_ptr;   /* device counter mixed into the in-kernel RNG, may be NULL */
    const int32_t* done_flag;  /* device flag: non-zero => kernels return immediately, may be NULL */
    int64_t chain_offset;      /* global index of this rank's chain 0 (restarts sharded over GPUs): the
                                  in-kernel RNG is keyed by the GLOBAL chain index, so a sharded run draws the same
                                  numbers per restart as the unsharded one */
} nmb_ts_problem_t;

typedef struct nmb_ts_rollout {
    const void* theta;         /* [n_params] torch layout (used by backward) */
    const void* theta_t;       /* [n_params] transposed layout (used by forward) */
    int32_t noisy_design;      /* enable_noise: rotate the design outputs (acqfs.py:320-326) */
    int32_t noisy_action;      /* the action call always rotates in the reference (acqfs.py:390-395) */
    const void* rot_q;         /* [L+1][N][4] host-built float32-rounded 2x2 blocks (amortized_network.py:144-153) or NULL */
    const int32_t* rot_dims;   /* [L+1][N][2] coordinate pairs (amortized_network.py:156) */
    const void* cost_noise;    /* [L+1][N] N(0,1) draws of the L+1 cost calls (acqfs.py:624) or NULL */
    uint32_t rng_stream;
    uint32_t _pad;
    void* X;                   /* out [L][N][d]   design points x_1..x_L */
    void* A;                   /* out [N][d]      action */
    void* Y;                   /* out [L+1][N]    f(x_1..x_L), f(a) */
    void* loss;                /* out [N]         -f(a)  (top-k loss, n_actions = 1) */
    void* cost;                /* out [N]         accumulated spotlight cost */
    void* save;                /* workspace for backward (nmb_ts_save_bytes) or NULL for a no-grad rollout */
    /* Host-stream ring (optional; takes precedence over rot_q / cost_noise and over the in-kernel RNG): the draws of
     * ring_len consecutive Adam iterations, produced ahead of time on the host in the reference's generator order
     * (nmb_host_rng_ts_am_stream) and uploaded while earlier iterations run; iteration `*step_ptr` reads slot
     * step % ring_len, so ONE captured CUDA graph serves every iteration with the reference's exact random stream. */
    const void* ring_cs;       /* [ring_len][L+1][N][2] float32 (cos, sin) of the rotation angle (amortized_network.py:144-153) */
    const int32_t* ring_dims;  /* [ring_len][L+1][N][2] coordinate pairs */
    const void* ring_noise;    /* [ring_len][L+1][N]    N(0,1) draws of the cost calls, problem dtype */
    int32_t ring_len;          /* 0: no ring */
    int32_t _pad2;
    /* Imitation mode (optional; Actor.reset_parameters, reference actor.py:164-220): y_obs [L][N] non-NULL => this
     * launch runs only level `level_begin` (level_end = level_begin + 1) of the rollout, with no Thompson path and no
     * cost: the observation fed to level l > 0 is y_obs[l-1] - a draw of the surrogate's marginal posterior at the
     * previous output (nmb_gp_whiten_fwd + nmb_gp_marginal_sample between the launches), detached like the
     * reference's `.sample()` - its input point is X[l-1] and its hidden state the saved HS[l].  The saved rollout
     * (dy/dx = 0) is then differentiated by nmb_ts_rollout_bwd / nmb_policy_wgrad unchanged, seeded through g_X / g_A. */
    const void* y_obs;
    int32_t level_begin, level_end;
} nmb_ts_rollout_t;

int64_t nmb_ts_save_bytes(int dtype, int d, int H, int L, int64_t n_chains);
/* element offset of the hidden-state block HS[L+2][N][H] inside a save workspace */
int64_t nmb_ts_save_hidden_offset(int d, int H, int L, int64_t n_chains);

/* forward for n_rollouts (1 or 2) rollouts in ONE launch (grid.y = rollout) */
int nmb_ts_rollout_fwd(const nmb_ts_problem_t* prob, const nmb_ts_rollout_t* rollouts, int n_rollouts, void* stream);

/* backward of one saved rollout.  Seeds: g_loss[N], g_cost[N] (required), g_X [L][N][d],
 * g_A [N][d] (optional extra adjoints of the returned points, e.g. the cosine term of
 * actor.py:594-609).  Fills the pre-activation adjoints in `save` for nmb_policy_wgrad.
 * g_prev (optional) receives [N][d+1+H] adjoints of (prev_x, prev_y, prev_h). */
int nmb_ts_rollout_bwd(const nmb_ts_problem_t* prob, const nmb_ts_rollout_t* rollout, const void* g_loss,
                       const void* g_cost, const void* g_X, const void* g_A, void* g_prev, void* stream);

/* Test hook: the in-kernel ("device" rng mode) draws of Adam iteration `step`, exactly as the rollout kernels make
 * them (same device functions): cs [n_calls][N][2] = (cos, sin) of the rotation angle (amortized_network.py:136,
 * 144-153), dims [n_calls][N][2] = the coordinate pair (amortized_network.py:156), noise [n_calls][N] = the N(0,1)
 * draw of the cost call (acqfs.py:624).  Keyed by (seed, step, rng_stream, call, chain_offset + chain). */
int nmb_ts_device_draws(uint64_t seed, int64_t step, unsigned rng_stream, int n_calls, int64_t n_chains,
                        int64_t chain_offset, int d, double* cs, int32_t* dims, double* noise, void* stream);

/* Weight gradients of a saved+back-propagated rollout: split-K GEMMs over all nodes.
 * partials: [n_splits_max][n_params]; seg_splits[20] (host, out) = how many leading
 * rows of `partials` hold data for each parameter segment.  Deterministic (no atomics). */
int64_t nmb_policy_wgrad_max_splits(int L, int64_t n_chains);
int nmb_policy_wgrad(const nmb_ts_problem_t* prob, const void* save, void* partials, int32_t* seg_splits_host,
                     void* stream);
/* grad[i] = sum_s partials[s][i] over the segment's split count */
int nmb_policy_reduce_partials(int dtype, int d, int H, int n_actions, const void* partials,
                               const int32_t* seg_splits_host, void* grad, void* stream);

/* ------------------------------------------------------------------------
 * Device-resident optimisation loop state (reference actor.py:489-662: AdamW loop, best-restart
 * bookkeeping, early stop).  Lives in device memory so that one Adam iteration is a fixed
 * sequence of launches (CUDA-graph capturable) with no host synchronisation.
 * ---------------------------------------------------------------------- */
/* The device buffer that holds the loop state must be NMB_LOOP_STATE_BYTES long and zero-initialised: the struct
 * below occupies the first bytes, the rest is scratch for the multi-CTA reduction of nmb_ts_objective. */
#define NMB_LOOP_STATE_BYTES 4096
typedef struct nmb_loop_state {
    int64_t step;          /* iterations completed (ep) */
    int32_t done;          /* loop has stopped (early stop): kernels become no-ops */
    int32_t done_next;     /* set by the bookkeeping of this iteration, latched into done by the optimiser step */
    int32_t stall;         /* early-stop counter (actor.py:636-639) */
    int32_t improved;      /* this iteration took a snapshot (actor.py:623-636) */
    int32_t chosen_idx;    /* argmin of this iteration's total loss */
    int32_t n_snapshots;
    double  best_min;      /* min over restarts of best_loss + best_cost + best_KL */
    double  loss_mean, cost_mean;   /* acqf_loss.mean(), acqf_cost.mean() of the last iteration (actor.py:618-619) */
    double  cand_total;    /* this iteration's min over restarts of the total loss (offset 56) */
    double  cand_lck;      /* this iteration's min over restarts of loss + cost + KL (offset 64) */
} nmb_loop_state_t;

/* Per-restart objective of one iteration + bookkeeping (one CTA):
 *   KL[c]    = sum_l (1 - cos(X_orig[l][c], X[l][c])) + (1 - cos(A_orig[c], A[c]))   (actor.py:594-609; 0 if X_orig NULL)
 *   total[c] = loss[c] + |budget - cost[c]| + KL[c]                                   (actor.py:620)
 *   seeds of d mean_c(total) / d{loss, cost, X, A}  (the mean is over n_global restarts)
 *   chosen = argmin total; snapshot decision and early-stop counter (actor.py:622-639).
 * past_initial = (iteration > n_initial_points).  patience = early-stop threshold (reference: 50,
 * actor.py:655; < 0 disables early stopping).  hist_loss/hist_cost [max_iter][N] and hist_X optional
 * (the per-iteration records of actor.py:561-564,580-581). */
int nmb_ts_objective(int dtype, int d, int L, int64_t n_chains, int64_t n_global, double budget, int past_initial,
                     int patience, const void* loss, const void* cost, const void* X, const void* A,
                     const void* X_orig, const void* A_orig, void* KL, void* total, void* g_loss, void* g_cost,
                     void* g_X, void* g_A, void* hist_loss, void* hist_cost, void* hist_X /* [max_iter][L][N][d] */,
                     int64_t hist_capacity, int defer_decision, nmb_loop_state_t* state, void* stream);
/* The snapshot / early-stop decision of actor.py:622-639 from state->cand_total / cand_lck.  With restarts
 * sharded over ranks: nmb_ts_objective(defer_decision = 1), all-reduce(MIN) the two candidates in place
 * (16 bytes at offset 56 of the state), then this. */
int nmb_loop_decide(nmb_loop_state_t* state, int past_initial, int patience, void* stream);

/* if (state->improved) dst[i] <- src[i] for up to 8 (src, dst, n_elems) triples (snapshot, actor.py:624-635) */
int nmb_snapshot_if_improved(int dtype, int n_copies, const void* const* src_host, void* const* dst_host,
                             const int64_t* n_elems_host, const nmb_loop_state_t* state, void* stream);

/* AdamW step (torch.optim.AdamW semantics, actor.py:489,646-648) on the flat parameter buffer with the
 * split reduction of nmb_policy_wgrad fused in; also refreshes theta_t, latches done, advances step.
 * seg_splits_host[s] == 0 => the segment received no gradient and is skipped, like a None grad.
 * For free tensors use nmb_adamw_flat.  bias_corr: device [capacity][2] = (1-beta1^t, sqrt(1-beta2^t)), t = step+1. */
typedef struct nmb_adamw {
    double lr, beta1, beta2, eps, weight_decay;
} nmb_adamw_t;
int nmb_policy_adamw_step(int dtype, int d, int H, int n_actions, const nmb_adamw_t* hp, void* theta, void* theta_t,
                          void* exp_avg, void* exp_avg_sq, const void* partials, const int32_t* seg_splits_host,
                          nmb_loop_state_t* state, int advance_state, void* stream);
int nmb_adamw_flat(int dtype, int64_t n, const nmb_adamw_t* hp, void* param, void* exp_avg, void* exp_avg_sq,
                   const void* grad, nmb_loop_state_t* state, int advance_state, void* stream);

/* ------------------------------------------------------------------------
 * Cholesky root of the training covariance and its explicit inverse: the caches behind SingleTaskGP.posterior that
 * GPyTorch builds once per model (psd_safe_cholesky of K + s2 I and the inverse root it multiplies by; reached from
 * reference acqfs.py:258-266, 364-373).  `batch` row-major matrices, one CTA each (one warp for n <= 32):
 * A [n][lda] is overwritten by L (lower, strict upper triangle zeroed), Linv (may be NULL) receives L^-1.
 * info[b] follows LAPACK potrf: 0, or k > 0 when the leading minor of order k is not positive definite (the caller
 * applies GPyTorch's jitter rule and retries on a fresh copy).  n <= 800 (GPyTorch's max_cholesky_size).
 * ---------------------------------------------------------------------- */
int nmb_cholesky_root(int dtype, int n, int batch, void* A, int64_t lda, int64_t stride_a, void* Linv, int64_t ldi,
                      int64_t stride_i, int32_t* info, void* stream);

/* ------------------------------------------------------------------------
 * Exact marginal log-likelihood of the surrogate and its hyper-parameter gradient: the arithmetic of
 * `fit_gpytorch_mll(ExactMarginalLogLikelihood(likelihood, SingleTaskGP(...)))` in the reference's BO loop
 * (main.py:472-486; BoTorch / GPyTorch restated).  FP64.  Inputs: the training data,
 * 1 / lengthscale [d], the mean constant, and L / L^-1 of K + s2 I from nmb_kernel_matrix + nmb_cholesky_root.
 * out [5 + d] (device): 0 quad = (y-c)^T Khat^-1 (y-c), 1 log det Khat, 2 sum(alpha), 3 tr G, 4 tr(G kappa),
 * 5 + k tr(G dkappa / dlengthscale_k), with alpha = Khat^-1 (y-c), G = alpha alpha^T - Khat^-1, kappa the unit-amplitude
 * kernel, so that d(-log p)/d theta = -1/2 tr(G dKhat/d theta) and d(-log p)/dc = -sum(alpha).  Deterministic.
 * work: nmb_gp_mll_workspace_elems doubles.  n <= 800.
 * ---------------------------------------------------------------------- */
int64_t nmb_gp_mll_workspace_elems(int n, int d);
int nmb_gp_mll_value_grad(int kernel_id, int d, int n, const double* X, const double* inv_ls, const double* y,
                          double mean_const, const double* L, int64_t ldl, const double* Linv, int64_t ldi, double* work,
                          double* out, void* stream);

/* ------------------------------------------------------------------------
 * Allocation-free variant of nmb_gp_mll_value_grad.  NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops,
 * harnesses that forbid it).  Same inputs, same `out`, same kernels; the workspace is a file-scope
 * `static __device__` array inside the library (523 KB, held by every GPU that loads the library until exit) whose
 * device address is looked up once per device with cudaGetSymbolAddress.  Calls no cudaMalloc / cudaFree / allocating
 * library routine.  Limit: n <= 256 (NMB_ELIMIT beyond, with a message naming the limit); d <= NMB_MAX_DIM.
 * Every call on a device shares the one array, which keeps the last call's contents: each call overwrites everything it
 * reads, so nothing needs clearing, but two calls that overlap on the GPU (different streams, or threads without
 * ordering) RACE - serialise them, or use nmb_gp_mll_value_grad with separate caller workspaces.
 * ---------------------------------------------------------------------- */
int nmb_gp_mll_value_grad_static(int kernel_id, int d, int n, const double* X, const double* inv_ls, const double* y,
                                 double mean_const, const double* L, int64_t ldl, const double* Linv, int64_t ldi,
                                 double* out, void* stream);

/* ------------------------------------------------------------------------
 * Restart shards over NVLink peer memory (one process per GPU; SURVEY 8e C2 + C3): the per-iteration exchange of
 * the sharded Actor.query loop - policy-gradient SUM and the MIN of the two snapshot-decision candidates
 * (actor.py:622-639, 641-648) - written against peer-mapped buffers, fused with the AdamW update, instead of two
 * NCCL all-reduces.  Every rank allocates nmb_peer_exchange_bytes(dtype, n) zero-initialised bytes in memory its
 * peers can map (host side: torch.distributed._symmetric_memory), and passes all ranks' addresses of it.
 * Call order per iteration, on every rank, same stream: nmb_peer_publish -> nmb_peer_wait_decide ->
 * (nmb_snapshot_if_improved) -> nmb_peer_adamw_step.  CUDA-graph capturable; no-ops once state->done is set.
 * ---------------------------------------------------------------------- */
#define NMB_MAX_PEERS 16
#define NMB_PEER_HEADER_BYTES 256
#define NMB_PEER_TIMEOUT_STALL (-777)   /* state->stall after a peer failed to publish within ~2 s (loop stopped) */
typedef struct nmb_peer_exchange {
    int32_t rank, world;
    int64_t n;                    /* gradient elements per rank (policy parameter count) */
    void* base[NMB_MAX_PEERS];    /* base[r]: rank r's exchange buffer as mapped into THIS process */
} nmb_peer_exchange_t;
int64_t nmb_peer_exchange_bytes(int dtype, int64_t n);
/* seg_splits_host == NULL: grad [n] is this rank's gradient; otherwise grad is the [splits][n] partials buffer of
 * nmb_policy_wgrad and its seg_splits, reduced on the way out (d, H, n_actions give the parameter layout).
 * Reads state->cand_total / cand_lck (nmb_ts_objective with defer_decision = 1). */
int nmb_peer_publish(const nmb_peer_exchange_t* px, int dtype, const void* grad, int d, int H, int n_actions,
                     const int32_t* seg_splits_host, const nmb_loop_state_t* state, void* stream);
/* waits for every rank's publish, cand_* = MIN over ranks, then the decision of nmb_loop_decide */
int nmb_peer_wait_decide(const nmb_peer_exchange_t* px, int dtype, nmb_loop_state_t* state, int past_initial,
                         int patience, void* stream);
/* theta <- AdamW(theta, sum over ranks of the published gradients); advances the loop state and the exchange epoch */
int nmb_peer_adamw_step(const nmb_peer_exchange_t* px, int dtype, int d, int H, int n_actions, const nmb_adamw_t* hp,
                        void* theta, void* theta_t, void* exp_avg, void* exp_avg_sq, nmb_loop_state_t* state,
                        void* stream);
/* The three calls above as ONE launch (what the sharded Actor.query loop uses): CTA b reduces its 256-element slice of
 * the nmb_policy_wgrad partials into the rank's slot, raises a per-slice flag in every peer, waits for the same slice
 * of every peer, sums it in rank order and applies AdamW; the last CTA to finish takes the MIN of the candidates, the
 * decision of nmb_loop_decide, and advances the exchange epoch and the loop state.  Call nmb_snapshot_if_improved
 * AFTER it (the snapshot copies rollout outputs, never the weights).  Needs (n + 255) / 256 <= 8 x SM count
 * (NMB_ELIMIT otherwise: fall back to the three calls).  Same buffer, same epoch counter as the three-call form. */
int nmb_peer_exchange_adamw_step(const nmb_peer_exchange_t* px, int dtype, const void* partials, int d, int H,
                                 int n_actions, const int32_t* seg_splits_host, const nmb_adamw_t* hp, void* theta,
                                 void* theta_t, void* exp_avg, void* exp_avg_sq, nmb_loop_state_t* state,
                                 int past_initial, int patience, void* stream);

/* ------------------------------------------------------------------------
 * Exact-GP posterior over a fantasy tree (HES-MC and the qEI/qPI/qSR/qUCB/qKG baselines):
 * replaces model.posterior -> sampler -> condition_on_observations per level and the leaf
 * posterior + sampler (reference acqfs.py:363-373, 413-415; BoTorch/GPyTorch arithmetic of
 * SURVEY App. A.3-A.5) without materialising any per-node (n+s)x(n+s) root.
 *
 * With u(x) = L^-1 k(X_tr, x) for the base root L (one shared triangular product per query
 * point), conditioning on the q = 1 ancestors z_0..z_{s-1} of a node is the tail recursion
 *   w_t(x) = (k(z_t, x) - u(z_t).u(x) - sum_{t'<t} w_t'(z_t) w_t'(x)) / c_t,
 *   c_t    = sqrt(var_t(z_t) + noise),   var_s(x) = k(x,x) - |u(x)|^2 - sum_t w_t(x)^2,
 *   mean_s(x) = c + k(x, X_tr).alpha + sum_t w_t(x) b_t,   b_t = sigma_t(z_t) z_t^base[f_t] / c_t
 * (SURVEY App. A.4 "structural identity").  Tree layout as in nmb_tree_cost_*.
 * ---------------------------------------------------------------------- */
#define NMB_MAX_LEVELS 16
#define NMB_TILE_STRIDE 68   /* doubles per tile row: 64 columns + 4 pad (bank-conflict-free fragment loads) */
typedef struct nmb_gp {
    int32_t d, n, kernel_id;
    int32_t ldu;            /* row stride of U / gU (>= n, multiple of 4) */
    int32_t ldl;            /* leading dimension of linv / linv_t: n rounded up to a multiple of 8 */
    int32_t _pad;
    double amplitude, noise, mean_const;
    const void* train_x;    /* [n][d] */
    const void* inv_ls;     /* [d] */
    const void* linv;       /* [ldl][ldl] inverse Cholesky root, row-major, lower triangular, zero padded */
    const void* linv_t;     /* [ldl][ldl] its transpose, zero padded */
    const void* alpha;      /* [n] (K + s2 I)^-1 (y - c) */
    /* FP64 tensor-pipe kernels: the same two matrices tiled by 64-column blocks, [ceil(ldl/64)][ldl][NMB_TILE_STRIDE = 68]
     * (tile[b][k][c] = M[k][64 b + c] for c < 64 and 64 b + c < ldl, else 0): one bulk copy per operand chunk */
    const void* linv_tiles;
    const void* linv_t_tiles;
} nmb_gp_t;

typedef struct nmb_tree {
    int32_t n_levels;                 /* design levels + the action level */
    int32_t obs_noise_fantasy;        /* 1: fantasies drawn with observation noise (qKG) */
    int64_t counts[NMB_MAX_LEVELS];   /* nodes per level, counts[0] = n_restarts */
    const void* base[NMB_MAX_LEVELS]; /* device: base samples of level s < n_levels-1, [counts[s+1]/counts[s]] */
    double action_base_mean;          /* mean of the action sampler's base samples (n_actions = 1) */
} nmb_tree_t;

/* U[p][:] = L^-1 (amp k(X_tr, x_p)), m0[p] = amp k(x_p, X_tr).alpha, q[p] = |U[p]|^2 for n_pts points */
int nmb_gp_whiten_fwd(const nmb_gp_t* gp, int dtype, const void* pts, int64_t n_pts, void* U, void* m0, void* q,
                      void* stream);
/* g_pts[p][:] += d/dx_p of ( gU[p].u(x_p) + g_m0[p] m0(x_p) ) */
int nmb_gp_whiten_bwd(const nmb_gp_t* gp, int dtype, const void* pts, int64_t n_pts, const void* gU,
                      const void* g_m0, void* g_pts, void* stream);

/* per-node state of a tree evaluation; every array is caller-allocated, [total] = sum(counts) */
typedef struct nmb_tree_state {
    const void* pts;    /* [total][d] all levels back to back */
    void *U, *m0, *q;   /* whitened vectors [total][ldu], base mean part, squared norm */
    void *W;            /* [total][n_levels] tails w_t(x) */
    void *var, *mean, *cst, *sig, *bv;   /* [total]: posterior variance / mean, c, sigma, b of the node's branch */
    void *leaf_loss;    /* [counts[n_levels-1]] -(mean + sigma * action_base_mean) */
    /* backward */
    void *Gw, *Gdot;    /* [total][n_levels] */
    void *Gc, *Gb, *Gq, *Gm0;   /* [total] */
    void *gU;           /* [total][ldu] */
    void *g_pts;        /* [total][d] */
    void *Gscr;         /* [nmb_mc_level_bwd_scratch_elems] per-level scratch of the deterministic ancestor reduction */
} nmb_tree_state_t;

/* forward of one level (levels must be issued in order 0..n_levels-1 after nmb_gp_whiten_fwd on all points) */
int nmb_mc_level_fwd(const nmb_gp_t* gp, const nmb_tree_t* tree, const nmb_tree_state_t* st, int dtype, int level,
                     void* stream);
/* elements (of the call's dtype) of nmb_tree_state_t.Gscr */
int64_t nmb_mc_level_bwd_scratch_elems(const nmb_tree_t* tree, int d);
/* backward of one level (issue levels n_levels-1..0; Gw/Gc/Gb/g_pts zeroed by the caller first); deterministic:
 * a node's contributions to its ancestors go through Gscr and are summed in a fixed order (no atomics);
 * g_leaf_sig == NULL: g_leaf [counts[n_levels-1]] is the adjoint of leaf_loss; otherwise g_leaf / g_leaf_sig are
 * the adjoints of the leaf nodes' posterior mean / sigma (baselines) */
int nmb_mc_level_bwd(const nmb_gp_t* gp, const nmb_tree_t* tree, const nmb_tree_state_t* st, int dtype, int level,
                     const void* g_leaf, const void* g_leaf_sig, void* stream);
/* gU of every node from Gq / Gdot (own, ancestors', descendants' terms) */
int nmb_mc_gather_gu(const nmb_gp_t* gp, const nmb_tree_t* tree, const nmb_tree_state_t* st, int dtype, void* stream);
/* out[r] = mean over the leaves of restart r (leaf l belongs to restart l % n_r); reference acqfs.py:443-447 */
int nmb_tree_restart_mean(int dtype, const void* leaf, int64_t n_leaf, int64_t n_r, void* out, void* stream);

/* Monte-Carlo utilities of the myopic baselines at q = 1: value[r] = mean_s u(mean_r + sigma_r z_s)
 * and its derivatives w.r.t. mean and sigma.  which: 0 qEI relu(y - best_f), 1 qPI sigmoid((y - best_f)/1e-3),
 * 2 qSR y, 3 qUCB mu^ + sqrt(beta pi/2) |y - mu^| with mu^ = mean_s y. */
int nmb_mc_utility(int dtype, int which, double best_f, double beta, const void* mean, const void* sig, int64_t n,
                   const void* base, int n_samples, void* value, void* d_mean, void* d_sig, void* stream);

/* ------------------------------------------------------------------------
 * Amortized pre-training of Actor.reset_parameters (reference actor.py:164-220) - the pieces that are not already
 * the query loop's kernels (nmb_ts_rollout_fwd in imitation mode, nmb_ts_rollout_bwd, nmb_policy_wgrad,
 * nmb_policy_adamw_step):
 *   nmb_gp_marginal_sample   y[i] = mean_const + m0[i] + sqrt(amplitude - q[i]) z[i]: one draw of the surrogate's
 *       marginal posterior per point (`model(x.unsqueeze(-2)).sample()`, actor.py:186-191) from the outputs of
 *       nmb_gp_whiten_fwd; the 1x1 "Cholesky" follows GPyTorch's jitter rule.  z: explicit [n], or slot
 *       (*step_ptr % ring_len, level) of a host-stream ring z_ring [ring_len][n_levels][n] (the reference's
 *       generator order, nmb_host_rng_randn_f64), or NULL/NULL = in-kernel Philox keyed by (seed, *step_ptr,
 *       level, chain_offset + i).  done_flag (optional): non-zero => no-op.
 *   nmb_imitation_objective  loss = mean |target - out| over the (L+1) N d rollout outputs (actor.py:198), its
 *       adjoint seeds g_X [L][N][d], g_A [N][d] (= sign(out - target) / numel; g_loss, g_cost [N] are zeroed), the
 *       per-epoch record hist[step] and the early-stop rule of actor.py:200-219 in `state` (best loss in best_min,
 *       counter in stall, done_next once stall >= patience; patience < 0 disables).  One CTA, fixed order.
 * ---------------------------------------------------------------------- */
int nmb_gp_marginal_sample(int dtype, double amplitude, double mean_const, const void* m0, const void* q, int64_t n,
                           const void* z, const void* z_ring, int ring_len, int n_levels, int level, uint64_t seed,
                           int64_t chain_offset, const int64_t* step_ptr, const int32_t* done_flag, void* y, void* stream);
int nmb_imitation_objective(int dtype, int d, int L, int64_t n_chains, const void* X, const void* A, const void* target,
                            void* g_X, void* g_A, void* g_loss, void* g_cost, void* hist, int64_t max_epochs,
                            int patience, nmb_loop_state_t* state, void* stream);

/* ------------------------------------------------------------------------
 * FP64 / FP32 pipe micro-benchmarks used for the roofline denominators
 * (MEASURED_PEAKS.json has no FP64 entry).
 * which: 0 = dependent-chain-free FMA throughput, 1 = sincospi, 2 = exp.
 * Writes the achieved operations/second (FMA counted as 2 flops for which=0,
 * one op per call otherwise) to *host_result. Synchronous.
 * ---------------------------------------------------------------------- */
int nmb_pipe_peak(int dtype, int which, double* host_result);

/* Device-side evaluation of the straight-line FP64 transcendentals the rollout kernels use (csrc/fastmath.cuh), for
 * accuracy tests against libm: y0 = table-driven variant, y1 = Taylor variant, n doubles each.
 * which: 0 exp, 1 expm1, 2 sincospi (y0 = sin, y1 = cos, table), 3 sigmoid, 4 tanh, 5 reciprocal (2 / 3 Newton
 * steps), 6 elu, 7 sincospi (Taylor).  They stand in for torch's exp / sigmoid / tanh / elu / sin / cos inside
 * amortized_network.py:97-133 and the Matheron path features (acqfs.py:362). */
int nmb_fastmath_probe(int which, const void* x, void* y0, void* y1, int64_t n, void* stream);

/* ------------------------------------------------------------------------
 * Host-side replica of torch's CPU generator streams (pure host code, no CUDA call): the random numbers the
 * reference draws inside every Adam iteration of Actor.query, produced in bulk, in the reference's order, bit for
 * bit, so that the CUDA-graph loop can read them from a device ring ("same inputs and seeds => identical next
 * query").  `state` is the 5056-byte blob of torch.get_rng_state() (ATen CPUGeneratorImplState: mt19937 words,
 * position, cached normal) and is advanced in place.
 *   nmb_host_rng_rand_f32      torch.rand(n)                       amortized_network.py:136
 *   nmb_host_rng_randperm      n_perms x torch.randperm(d) [int64] amortized_network.py:156
 *   nmb_host_rng_randn_f64     n_calls x torch.randn(sizes[c], dtype=float64), back to back   acqfs.py:624
 *   nmb_host_rng_ts_am_stream  n_iters Adam iterations of HES-TS-AM (actor.py:568-593): per iteration the main
 *       rollout's action-call rotation draw (u_main [it][N] float32 in [0,1), dims_main [it][N][2]) and its L+1 cost
 *       noises (nz_main [it][L+1][N]), then the frozen rollout's L+1 rotation draws (u_orig [it][L+1][N],
 *       dims_orig [it][L+1][N][2]) and its L+1 cost noises (nz_orig [it][L+1][N]).  states_out (optional,
 *       [n_iters][5056]) receives the generator state after every iteration, so an early stop can leave the
 *       generator exactly where the reference would have.  n_threads: host threads for the uniform -> normal
 *       transform (the word stream itself is serial).
 * ---------------------------------------------------------------------- */
int64_t nmb_host_rng_state_bytes(void);
int nmb_host_rng_rand_f32(uint8_t* state, int64_t n, float* out);
int nmb_host_rng_randperm(uint8_t* state, int n_perms, int d, int64_t* out);
int nmb_host_rng_randn_f64(uint8_t* state, int n_calls, const int64_t* sizes, double* out, int n_threads);
int nmb_host_rng_ts_am_stream(uint8_t* state, int n_iters, int L, int64_t N, int d, float* u_main, int32_t* dims_main,
                              double* nz_main, float* u_orig, int32_t* dims_orig, double* nz_orig,
                              uint8_t* states_out, int n_threads);

#ifdef __cplusplus
}
#endif
#endif /* NONMYOPIA_B200_H */
