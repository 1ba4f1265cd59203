/*
 * secem.h -- C ABI of libsecem.so: the B200 (sm_100a) implementation of safe-exploration's safe CEM-MPC hot path.
 *
 * The reference (oscarkey/safe-exploration @ 32f0582a + constrained-cem-mpc @ 4e207630) is pure Python and has no
 * FFI; its plug-in boundaries on this path are the Python classes `CemSSM`, `DynamicsFunc`/`Constraint`,
 * `ConstrainedCemMpc` and `SafeMPC`.  The Python mirror of those classes lives in
 * `safe_exploration_b200/`; this header is what it binds with `ctypes`, and what a maintainer of the reference would
 * bind from `safe_exploration/ssm_cem/` and `constrained_cem_mpc/`.  Each entry point cites the
 * reference interface it replaces (paths relative to the reference root).
 *
 * Conventions
 *   - every function returns 0 on success, <0 on error; `se_last_error()` returns the message (thread-local).
 *   - all `dev_*` / `void*` array arguments are CALLER-OWNED DEVICE pointers, row-major contiguous, element type
 *     selected by the handle's `se_dtype` (float or double); `host_*` arguments are host pointers to double.
 *   - every call takes a `cudaStream_t` (passed as void*) and is asynchronous on it: no call synchronises the device
 *     unless its comment says so.  Handles are not thread-safe; one handle per GPU.
 *   - no torch / C++ types cross this boundary.
 */
#ifndef SECEM_H_
#define SECEM_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define SE_MAX_NS 6   /* state dimension  (reference: pendulum 2, cart-pole 4, lunar lander 6) */
#define SE_MAX_NU 3   /* action dimension */
#define SE_MAX_FACES 16 /* faces of the safety polytope (pendulum 4, cart-pole 9) */
#define SE_MAX_ACT_FACES 8
#define SE_MAX_OBS_FACES 8
#define SE_MAX_T 64   /* horizon */

typedef enum { SE_F32 = 0, SE_F64 = 1 } se_dtype;

/* objective selector: safe_exploration/safempc_cem.py:304-312 */
typedef enum {
  SE_OBJ_NEG_VARIANCE = 0, /* default: -sum_j sigma[:, j]                         (safempc_cem.py:310) */
  SE_OBJ_ABS_STATE1 = 1,   /* |theta* - p[:, 1]|  (InvertedPendulum, environments.py:505-510)          */
  SE_OBJ_NEG_LAST_STATE = 2 /* -p[:, -1]           (LunarLander, lunarlander.py:111-113)                */
} se_objective;

/* which variance-GEMM implementation runs (all produce the same quantity) */
typedef enum {
  SE_GEMM_AUTO = 0,
  SE_GEMM_SIMT = 1, /* FP32 FFMA / FP64 DFMA register-tiled */
  SE_GEMM_MMA = 2,  /* FP64: DMMA m8n8k4 (mma.sync) */
  SE_GEMM_TC = 3    /* FP64 models on the 5th-gen tensor cores: tcgen05.mma kind::i8 over balanced radix-256 digit
                       planes of both operands with exact int32 accumulation in TMEM (Ozaki scheme), recombined in
                       int64 / double.  SE_GEMM_AUTO picks it for float64 models of 256 .. 16384 training points */
} se_gemm_impl;

typedef struct se_gp_s* se_gp_t;
typedef struct se_cem_s* se_cem_t;

const char* se_last_error(void);
int se_version(void);

/* ---------------------------------------------------------------------------------------------------------------
 * Exact-GP state-space model.  Replaces GpCemSSM (safe_exploration/ssm_cem/gp_ssm_cem.py:17-133) behind the CemSSM
 * interface (safe_exploration/ssm_cem/ssm_cem.py:11-131).
 * ------------------------------------------------------------------------------------------------------------- */

/* Build / refresh the device-side model.  Replaces GpCemSSM._update_model (gp_ssm_cem.py:96-101): the reference hands
 * (X, Y^T) to GPyTorch which lazily builds its prediction caches; here the caller passes the materialised caches.
 *   dev_x           [n, d]        training inputs, d = n_s + n_u
 *   dev_alpha       [n_s, n]      (K_o + noise_o I)^-1 y_o
 *   dev_linv        [n_s, n, n]   L_o^-1 (lower triangular), K_o + noise_o I = L_o L_o^T
 *   host_lengthscale[n_s, d], host_outputscale[n_s], host_noise[n_s]  materialised (not raw) hyper-parameters of
 *                   ScaleKernel(RBFKernel(ard)) + GaussianLikelihood (gp_ssm_cem.py:33, 45-57)
 * The handle keeps its own padded copies; the caller may free its arrays once the stream has passed this call. */
int se_gp_create(se_gp_t* out, int n_s, int n_u, int n, se_dtype dtype, const void* dev_x, const void* dev_alpha,
                 const void* dev_linv, const double* host_lengthscale, const double* host_outputscale,
                 const double* host_noise, void* stream);
int se_gp_destroy(se_gp_t gp);
/* workspace cap (bytes) for the materialised kernel rows K*[n_s, N_chunk, n]; default 6 GiB */
int se_gp_set_workspace_limit(se_gp_t gp, uint64_t bytes);
int se_gp_set_gemm_impl(se_gp_t gp, se_gemm_impl impl);
/* digit planes per operand of SE_GEMM_TC.  0 (default) = chosen from max_o outputscale_o / noise_o for a relative
 * variance error <= 1e-8 (never fewer than 6); measured against the float64 DMMA path at outputscale / noise = 10^3:
 * 6 planes 1.7e-9 (21 int8 MMAs per K step), 7 planes 1.5e-11 (28), 8 planes 7e-12 (36, float64 rounding level),
 * 5 planes 3e-7 (15).  4 and 3 planes (10 / 6 MMAs) are the float32-TOLERANCE settings of BASELINE's "within 1e-4
 * relative in fp32": the truncation is ~1e-7 / ~3e-5 of the prior scale outputscale + noise (float32 arithmetic itself
 * reaches ~2e-7 of it at best), mean and Jacobian stay float64. */
int se_gp_set_tc_digits(se_gp_t gp, int digits);
/* which variance-GEMM implementation the handle resolved to, and the digit planes in use (0 until the first call) */
int se_gp_get_gemm_info(se_gp_t gp, int* impl_out, int* tc_digits_out);

/* mean [N, n_s], variance [N, n_s] (likelihood noise INCLUDED, gp_ssm_cem.py:93) and, if dev_jac != NULL, the
 * Jacobian of the mean w.r.t. z, [N, n_s, d].  Replaces GpCemSSM.predict_without_jacobians / predict_with_jacobians
 * (gp_ssm_cem.py:59-88) and compute_jacobian_fast (safe_exploration/ssm_pytorch/utilities.py:54-85).
 *   dev_z [N, d] = [states | actions] */
int se_gp_predict(se_gp_t gp, const void* dev_z, int64_t N, void* dev_mean, void* dev_var, void* dev_jac,
                  void* stream);

/* ---------------------------------------------------------------------------------------------------------------
 * Model refresh: what happens between episodes in CemSafeMPC.update_model
 * (safe_exploration/safempc_cem.py:314-327) -> CemSSM.update_model (ssm_cem.py:72-99) -> GpCemSSM._update_model /
 * _train_model (gp_ssm_cem.py:96-129).  In the reference the arithmetic is GPyTorch's; here the covariance assembly
 * (se_gp_kernel_matrix), the factorisation / inverse factor (se_gp_factor), the inverse covariance (se_gp_kinv) and the
 * marginal-log-likelihood gradient (se_gp_mll_grad) are all kernels of this library.  float64 only.
 * ------------------------------------------------------------------------------------------------------------- */

/* dev_k [n_s, n, n] = outputscale_o exp(-1/2 sum_c ((x_ic - x_jc) / lengthscale_oc)^2) + noise_o [i == j]: the
 * covariance of ScaleKernel(RBFKernel(ard_num_dims=d, batch_size=n_s)) plus the GaussianLikelihood noise
 * (gp_ssm_cem.py:33, 45-57) on the training inputs dev_x [n, d], for all n_s outputs in one launch. */
int se_gp_kernel_matrix(int n_s, int d, int64_t n, const double* dev_x, const double* host_lengthscale,
                        const double* host_outputscale, const double* host_noise, double* dev_k, void* stream);

/* Gradient of the exact marginal log likelihood MLL_o = -1/2 y_o^T alpha_o - sum_i log L_o,ii - n/2 log 2 pi of each
 * output's GP with respect to its MATERIALISED hyper-parameters:
 *   dev_grad [n_s, d + 2] = d MLL_o / d (lengthscale_o1 .. lengthscale_od, outputscale_o, noise_o)
 *                         = 1/2 sum_ij (alpha alpha^T - (K_o + noise_o I)^-1)_ij  d K_o,ij / d theta
 * -- what loss.backward() computes in GpCemSSM._train_model (gp_ssm_cem.py:116-121) up to the chain rule through
 * GPyTorch's softplus parametrisation and ExactMarginalLogLikelihood's 1/n, which the host applies.
 *   dev_alpha [n_s, n], dev_kinv [n_s, n, n] (symmetric; only the lower triangle is read)
 *   dev_workspace: se_gp_mll_grad_workspace(n_s, d, n) bytes of per-tile partial sums (reduced in a fixed order:
 *   the result is deterministic). */
int64_t se_gp_mll_grad_workspace(int n_s, int d, int64_t n);
int se_gp_mll_grad(int n_s, int d, int64_t n, const double* dev_x, const double* dev_alpha, const double* dev_kinv,
                   const double* host_lengthscale, const double* host_outputscale, const double* host_noise,
                   double* dev_workspace, double* dev_grad, void* stream);

/* On-device factor refresh: W_o = L_o^-1 for K_o + noise_o I = L_o L_o^T, built block row by block
 * row of 128 points from FP64 tensor-core (DMMA) products and one small Cholesky per block -- the arithmetic GPyTorch
 * does lazily behind GpCemSSM._update_model (ssm_cem/gp_ssm_cem.py:96-101).  Because a block row only needs the rows
 * above it, a model that only GREW (CemSSM.update_model(replace_old=False), ssm_cem.py:72-99; the reference's
 * update_cholesky idea, ssm_pytorch/utilities.py:122-165) is extended from its old factor: pass the first block row that
 * changed.  No cuSOLVER / cuBLAS call.
 *   n_pad          a multiple of 128 >= n; rows / columns >= n of dev_k hold the identity (1 on the diagonal)
 *   dev_k          [n_s, n_pad, n_pad] covariance (only rows >= 128 first_block are read, lower triangle + diagonal blocks)
 *   dev_w, dev_wt  [n_s, n_pad, n_pad] W and W^T; rows (of W) / columns (of W^T) below 128 first_block must hold the old
 *                  factor, everything else must be ZERO on entry
 *   dev_diag       [n_s, n_pad] diagonal of L (log det K = 2 sum log diag)
 *   dev_workspace  se_gp_factor_workspace(n_s, n_pad) bytes
 *   dev_info       int32[n_s], zeroed by the caller: 1 + the first row whose pivot was not positive (0 = success) */
int64_t se_gp_factor_workspace(int n_s, int64_t n_pad);
int se_gp_factor(int n_s, int64_t n_pad, int first_block, const double* dev_k, double* dev_w, double* dev_wt,
                 double* dev_diag, double* dev_workspace, int32_t* dev_info, void* stream);
/* dev_kinv [n_s, n_pad, n_pad] = W^T W = (K + noise I)^-1, LOWER triangle only (64 x 128 tiles that touch it; the rest is
 * left untouched), from dev_wt = W^T as written by se_gp_factor: what se_gp_mll_grad reads.  The reference gets the same
 * quantity implicitly from autograd through GPyTorch's factorisation (gp_ssm_cem.py:116-121). */
int se_gp_kinv(int n_s, int64_t n_pad, const double* dev_wt, double* dev_kinv, void* stream);

/* se_gp_mll_grad without a caller workspace and without any allocation.  NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight
 * serving loops, harnesses that forbid it).  The per-tile partial sums live in a file-scope
 * `static __device__ __align__(256)` array of SE_MLL_GRAD_STATIC_MAX_BYTES inside libsecem.so (every GPU that loads
 * the library holds it until the process exits); the entry point resolves its device address once per device with
 * cudaGetSymbolAddress and calls no cudaMalloc / cudaFree / allocating library routine.  Same arguments and result
 * (bit for bit: same kernels, same fixed reduction order) as se_gp_mll_grad minus dev_workspace.
 *   - sizes with se_gp_mll_grad_workspace(n_s, d, n) > SE_MLL_GRAD_STATIC_MAX_BYTES are refused (< 0, message in
 *     se_last_error()); 4 MiB covers n = 5000, n_s = 4, d = 5 (0.7 MB) up to n = 8000, n_s = 6, d = 9.
 *   - every call on a device shares the one array (it keeps the last call's contents; nothing in it has to start at
 *     zero, each call overwrites every slot it reads).  Calls that OVERLAP on the same GPU -- two streams, or two
 *     host threads -- race on it: serialise them (one stream, or events between streams). */
#define SE_MLL_GRAD_STATIC_MAX_BYTES (4 << 20)
int64_t se_gp_mll_grad_static_max_bytes(void);
int se_gp_mll_grad_static(int n_s, int d, int64_t n, const double* dev_x, const double* dev_alpha,
                          const double* dev_kinv, const double* host_lengthscale, const double* host_outputscale,
                          const double* host_noise, double* dev_grad, void* stream);

/* ---------------------------------------------------------------------------------------------------------------
 * Problem constants: what CemSafeMPC._dynamics_func and construct_constraints close over
 * (safe_exploration/safempc_cem.py:135-146, 162-200, 288-312).  All host pointers, double.
 * ------------------------------------------------------------------------------------------------------------- */
typedef struct {
  int n_s, n_u;
  const double* k_fb;    /* [n_u, n_s]  LQR feedback gain (safempc_simple.py:1118-1129)                   */
  const double* a;       /* [n_s, n_s]  prior linear model (zeros when use_prior_model=False)              */
  const double* b;       /* [n_s, n_u]                                                                     */
  const double* l_mu;    /* [n_s]       Lipschitz constants (environments.py:477-482, 936-941)             */
  const double* l_sigma; /* [n_s]                                                                          */
  double c_safety;       /* beta_safety                                                                    */
  int m_safe;            /* faces of the safety polytope h_mat x <= h_vec (0: no terminal constraint)      */
  const double* h_mat;   /* [m_safe, n_s]                                                                  */
  const double* h_vec;   /* [m_safe]                                                                       */
  int m_act;             /* faces of the action polytope (0: no action constraint)                         */
  const double* act_a;   /* [m_act, n_u]                                                                   */
  const double* act_b;   /* [m_act]                                                                        */
  double act_penalty;    /* 3   (constrained_cem_mpc.py:103)                                               */
  double state_penalty;  /* 10  (safempc_cem.py:132)                                                       */
  int objective;         /* se_objective                                                                   */
  double objective_theta;
  int all_states_constraint; /* 0 = reference behaviour (terminal ellipsoid only, quirk Q1); 1 = check every
                                ellipsoid of the trajectory (intended semantics)                           */
  /* Obstacle polytope for the INTERMEDIATE ellipsoids (steps 1 .. T-1), the constraint the reference's CasADi solver
   * poses (safempc_simple.py:373-383: lin_ellipsoid_safety_distance(p_i, q_i, h_mat_obs, h_obs) <= 0 for i < H-1) and
   * that construct_constraints drops on the CEM path (safempc_cem.py:137 reads h_mat_obs, h_obs and never uses them).
   * Every intermediate ellipsoid that is not strictly inside costs state_penalty.  m_obs = 0: reference behaviour. */
  int m_obs;               /* faces of h_mat_obs x <= h_vec_obs, at most SE_MAX_OBS_FACES                  */
  const double* h_mat_obs; /* [m_obs, n_s]                                                                 */
  const double* h_vec_obs; /* [m_obs]                                                                      */
  /* _fix_zeros_nans (gp_reachability_pytorch.py:234-243) sets EVERY x <= 0 of a batch to 1e-5 when the batch holds an
   * exact zero.  se_reach_onestep / se_reach_multistep / se_reach_step apply that rule to the batch they are given
   * (they scan it first).  Inside se_cem_solve the "batch" of the reference (all candidates of one step) never exists
   * in one place, so the rule is a switch: 0 = only exact zeros are replaced, a negative variance surfaces as NaN at
   * the next sqrt and the solve reports it (flags[0]) -- what the reference does when the batch holds no exact zero;
   * 1 = every variance <= 0 becomes 1e-5 and is counted in flags[1] -- what it does when one is present. */
  int clamp_nonpositive;
} se_problem;

/* ---------------------------------------------------------------------------------------------------------------
 * Reachability.  Replaces safe_exploration/gp_reachability_pytorch.py.
 * ------------------------------------------------------------------------------------------------------------- */

/* One ellipsoid step for N candidates: onestep_reachability (gp_reachability_pytorch.py:18-181), including
 * compute_remainder_overapproximations_pytorch (utils.py:152-194), ellipsoid_from_rectangle_pytorch and
 * sum_two_ellipsoids_pytorch (utils_ellipsoid.py:102-140, 282-309) and _fix_zeros_nans (:234-243).
 *   dev_p [N, n_s]; dev_q [N, n_s, n_s] or NULL (point states); dev_u [N, n_u] (= k_ff)
 *   dev_p1 [N, n_s], dev_q1 [N, n_s, n_s], dev_sigma [N, n_s]
 *   dev_flags: int32[4] device counters, incremented: [0] #NaN variances, [1] #variances / bounds replaced by 1e-5
 *              (the reference raises / warns on these; the Python wrapper reads them after the call); [2] is scratch
 *              ("the batch of the last step held an exact zero") */
int se_reach_onestep(se_gp_t gp, const se_problem* prob, const void* dev_p, const void* dev_q, const void* dev_u,
                     int64_t N, void* dev_p1, void* dev_q1, void* dev_sigma, int32_t* dev_flags, void* stream);

/* The ellipsoid algebra of one step alone, for a posterior the CALLER evaluated: the interface fallback for a CemSSM
 * that is not the B200 GP (e.g. the MC-dropout models of ssm_cem/dropout_ssm_cem.py): the host calls
 * ssm.predict_with_jacobians / predict_without_jacobians (ssm_cem.py:36-58) on device tensors and hands the result
 * here.  Same arithmetic as se_reach_onestep after its GP stage (gp_reachability_pytorch.py:76-181).
 *   dev_mean [N, n_s], dev_var [N, n_s], dev_jac [N, n_s, n_s + n_u] (ignored when dev_q == NULL: point branch) */
int se_reach_step(se_dtype dtype, const se_problem* prob, const void* dev_p, const void* dev_q, const void* dev_u,
                  const void* dev_mean, const void* dev_var, const void* dev_jac, int64_t N, void* dev_p1,
                  void* dev_q1, void* dev_sigma, int32_t* dev_flags, void* stream);

/* T steps for N ellipsoids: the loop of RolloutFunction._sample_trajectory (constrained_cem_mpc.py:175-183) with the
 * dynamics of CemSafeMPC._dynamics_func, equivalently the NumPy multistep_reachability
 * (safe_exploration/gp_reachability.py:168-221) batched over N.
 *   dev_u [N, T, n_u]; host_k_fb_steps: NULL (use prob->k_fb for every step) or [T-1, n_u, n_s] per-step gains
 *   (gp_reachability.py:213-216); outputs dev_p_all [N, T, n_s] / dev_q_all [N, T, n_s, n_s] / dev_sigma_all
 *   [N, T, n_s] may be NULL; dev_pT [N, n_s], dev_qT [N, n_s, n_s] are the final ellipsoids. */
int se_reach_multistep(se_gp_t gp, const se_problem* prob, const void* dev_p0, const void* dev_q0, const void* dev_u,
                       int64_t N, int T, const double* host_k_fb_steps, void* dev_p_all, void* dev_q_all,
                       void* dev_sigma_all, void* dev_pT, void* dev_qT, int32_t* dev_flags, void* stream);

/* d [N, m] = h_mat p + c sqrt(diag(h_mat q h_mat^T)) - h_vec and inside[N] (uint8) = all(d < 0):
 * lin_ellipsoid_safety_distance / is_ellipsoid_inside_polytope (gp_reachability_pytorch.py:184-231).
 * dev_d or dev_inside may be NULL. */
int se_safety_distance(se_dtype dtype, int n_s, int m, const double* host_h_mat, const double* host_h_vec,
                       double c_safety, const void* dev_p, const void* dev_q, int64_t N, void* dev_d,
                       uint8_t* dev_inside, void* stream);

/* ---------------------------------------------------------------------------------------------------------------
 * Constrained CEM.  Replaces ConstrainedCemMpc (constrained_cem_mpc.py:195-280) + RolloutFunction (:130-192) with the
 * CemSafeMPC dynamics / constraints plugged in (safempc_cem.py:102-146, 288-312).
 * ------------------------------------------------------------------------------------------------------------- */
typedef struct {
  int num_instances;   /* M independent MPC problems solved side by side (1 for ConstrainedCemMpc)              */
  int num_rollouts;    /* N  population per instance (global, over all ranks)                                   */
  int time_horizon;    /* T                                                                                     */
  int num_elites;      /* E                                                                                     */
  int num_iterations;  /* I                                                                                     */
  int pick_best;       /* 0 = reference get_actions (first feasible rollout of the last iteration, quirk Q3);
                          1 = arg-min objective among feasible                                                  */
  /* sharding: this rank owns candidates [first_rollout, first_rollout + local_rollouts) of EVERY local instance;
     world_size == 1: first_rollout = 0, local_rollouts = num_rollouts */
  int first_rollout;
  int local_rollouts;
  /* instance sharding (BASELINE config "batched independent MPC solves"): global index of this handle's instance 0.
     The candidate RNG is keyed by (seed, iteration, GLOBAL instance id, rollout id), so a batch of instances dealt out
     to several GPUs draws exactly the candidates of the same batch solved on one.  0 for a single handle. */
  int first_instance;
} se_cem_config;

typedef struct {
  void* actions;        /* [M, T, n_u]  chosen action sequence (undefined where found == 0)                      */
  int32_t* found;       /* [M]          1 if a feasible rollout exists in the last iteration                     */
  int32_t* best_id;     /* [M]          global rollout id that was returned, -1 if none                          */
  void* means;          /* [I, M, T, n_u] sampling mean AFTER each iteration's refit, or NULL                    */
  void* stds;           /* [I, M, T, n_u] or NULL                                                                */
  int32_t* elite_ids;   /* [I, M, E]    global rollout ids in selection order, or NULL                           */
  void* objective;      /* [I, M, N_local] or NULL                                                               */
  void* constraint;     /* [I, M, N_local] or NULL                                                               */
  void* rollout_actions;/* [I, M, N_local, T, n_u] or NULL                                                       */
  void* trajectories;   /* [I, M, N_local, T+1, n_s + n_s^2] or NULL  ([p | vec(q)], PQFlattener layout)         */
  int32_t* flags;       /* int32[4] counters as in se_reach_onestep, or NULL                                     */
} se_cem_outputs;

int se_cem_create(se_cem_t* out, se_gp_t gp, const se_problem* prob, const se_cem_config* cfg);
int se_cem_destroy(se_cem_t cem);

/* Attach an NCCL communicator (ncclComm_t as void*) for population sharding: one all-gather of the per-candidate
 * (constraint, objective) costs per CEM iteration.  The reference has no counterpart (single process). */
int se_cem_set_comm(se_cem_t cem, void* nccl_comm, int rank, int world_size);

/* NCCL bootstrap helpers (the library resolves NCCL from the libnccl.so.2 the process already loaded, e.g. torch's):
 * rank 0 calls se_nccl_unique_id, the 128 bytes travel to the other ranks by any means (the Python host uses
 * torch.distributed.broadcast), then every rank calls se_nccl_comm_init.  No reference counterpart. */
int se_nccl_unique_id(char* out128);
int se_nccl_comm_init(void** comm_out, int world_size, const char* id128, int rank);
int se_nccl_comm_destroy(void* comm);

/* One MPC solve == ConstrainedCemMpc.get_actions (constrained_cem_mpc.py:259-280) for every instance.
 *   dev_p0    [M, n_s]   initial (point) states
 *   dev_noise NULL -> candidates are drawn on-device (Philox4x32-10 keyed by (seed, iteration, global rollout id));
 *             else [I, M, N_global, T, n_u] standard-normal noise: actions = means + stds * noise, the RNG-fixed
 *             equivalent of Normal(means, stds).sample((N,)) (constrained_cem_mpc.py:167) */
int se_cem_solve(se_cem_t cem, const void* dev_p0, const void* dev_noise, uint64_t seed, const se_cem_outputs* out,
                 void* stream);

/* The sampler alone: writes what se_cem_solve would draw.  dev_noise_out [M, N_count, T, n_u]. */
int se_cem_noise(se_dtype dtype, uint64_t seed, int iteration, int num_instances, int num_rollouts_global,
                 int first_rollout, int count, int T, int n_u, void* dev_noise_out, void* stream);

/* Elite selection + refit alone (constrained_cem_mpc.py:234-257) on given costs/actions, for parity tests of the
 * integer part: dev_objective/dev_constraint [N], dev_actions [N, T, n_u] -> dev_elite_ids [E] (int32, selection
 * order: ascending (cost, index)), dev_mean/dev_std [T, n_u]. */
int se_cem_select_refit(se_dtype dtype, const void* dev_objective, const void* dev_constraint,
                        const void* dev_actions, int N, int T, int n_u, int E, int32_t* dev_elite_ids,
                        void* dev_mean, void* dev_std, void* stream);

/* The return value of get_actions alone (constrained_cem_mpc.py:272-280) on given costs [N]: dev_best_id[0] = the
 * first rollout with constraint cost 0 (quirk Q3; pick_best != 0: the one of lowest objective, ties by index), -1 if
 * none.  With se_cem_noise and se_cem_select_refit this is what the interface fallback of B200ConstrainedCemMpc
 * (arbitrary Python DynamicsFunc / Constraint objects) runs on the device. */
int se_cem_pick(se_dtype dtype, const void* dev_objective, const void* dev_constraint, int N, int pick_best,
                int32_t* dev_best_id, void* stream);

/* ---------------------------------------------------------------------------------------------------------------
 * Measurement helpers (bench.py): launch counter and pipe-peak micro-benchmarks.
 * ------------------------------------------------------------------------------------------------------------- */
uint64_t se_launch_count(void);      /* kernels launched by this library since load / last reset */
void se_launch_count_reset(void);
/* Runs a register-resident FMA loop on every SM and returns the measured FLOP/s (synchronises the device).
 * kind: 0 = FP32 FFMA, 1 = FP64 DFMA, 2 = FP64 DMMA m8n8k4, 3 = int8 tcgen05.mma (OP/s = 2 x MAC/s, operands resident; burst),
 *       4 = the same sustained for ~1.5 s (what the pipe delivers under the 1 kW power cap),
 *       5 / 6 = burst / sustained with the variance GEMM's own instruction shape (128 x 64 x 32, A tile reused),
 *       7 / 8 / 9 / 10 = burst with two issuing warps (N = 64, N = 256), N = 128 with one and with two issuers:
 *       micro-benchmarks behind DESIGN.md's account of where the variance GEMM's time goes */
int se_measure_peak(int kind, double* flops_out);
/* Device time (ms, CUDA events recorded by the library on the launch stream) spent in the variance GEMM kernel and
 * in the kernel-row kernel since the last reset, and their launch counts; timing must be enabled first
 * (it inserts events, so leave it off outside benchmarks).  Synchronises the stream. */
int se_gp_timing_enable(se_gp_t gp, int enable);
int se_gp_timing_read(se_gp_t gp, double* gemm_ms, int64_t* gemm_launches, double* kstar_ms,
                      int64_t* kstar_launches, void* stream);

/* Test hook: runs the tcgen05 tile primitive of SE_GEMM_TC on one 128 x 64 tile.  dev_a_img [nk][digits][128 x 32 B],
 * dev_b_img [nk][digits][64 x 32 B] are int8 digit planes in the shared-memory operand order (8-row x 16-byte core
 * matrices; lbo / sbo = the two strides of the UMMA descriptor); dev_out [digits][128][64] int32 receives the level sums
 * D_l = sum_{t+u=l} A_t B_u^T. */
int se_debug_tc_probe(const int8_t* dev_a_img, const int8_t* dev_b_img, int nk, int digits, int lbo, int sbo,
                      int32_t* dev_out, void* stream);

/* Test hook (runs on the HOST, no GPU needed): the general small-matrix eigenvalue routine of the remainder bound
 * (csrc/eig_small.cuh; the device code calls the same function for a non-symmetric shape matrix, utils.py:175-185,
 * 651-665): host_out[0] = the largest real eigenvalue of the n x n row-major matrix host_m, NaN for a complex pair. */
int se_debug_eig_max(const double* host_m, int n, double* host_out);

/* Test hook: out[i] = the branch-free double exp() of the model-refresh gradient kernel (csrc/common.cuh exp_nb, polynomial from
 * tools/exp_poly.py), for finite dev_x[i] <= 700.  Lets a test pin its accuracy against a libm exp. */
int se_debug_exp(const double* dev_x, int64_t n, double* dev_out, void* stream);
/* The same for the table-based variant (csrc/common.cuh exp_tab) of the kernel-row kernel of the tensor-core path. */
int se_debug_exp_table(const double* dev_x, int64_t n, double* dev_out, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* SECEM_H_ */
