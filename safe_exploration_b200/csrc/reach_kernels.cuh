// reach_kernels.cuh -- per-candidate ellipsoid algebra (a13-a22 in SURVEY.md 8a), one thread per candidate with the
// n_s x n_s matrices held in registers (all loops have compile-time bounds).
//
// Restates safe_exploration/gp_reachability_pytorch.py:18-231, utils.py:152-194, utils_ellipsoid.py:102-140, 282-309
// and the constraint / objective epilogue of safempc_cem.py:102-132, 304-312 + constrained_cem_mpc.py:100-110.
#pragma once
#include "common.cuh"
#include "eig_small.cuh"

namespace se {

// largest eigenvalue of the symmetric NS x NS matrix m (cyclic Jacobi, destroys m)
template <typename T, int NS>
__device__ __forceinline__ T sym_lambda_max(T (&m)[NS][NS]) {
  if (NS == 1) return m[0][0];
  const T eps = sizeof(T) == 8 ? T(1e-300) : T(1e-37);
#pragma unroll 1
  for (int sweep = 0; sweep < 12; ++sweep) {
    T off = T(0), diag = T(0);
#pragma unroll
    for (int i = 0; i < NS; ++i) {
      diag += m[i][i] * m[i][i];
#pragma unroll
      for (int j = i + 1; j < NS; ++j) off += m[i][j] * m[i][j];
    }
    const T tol = sizeof(T) == 8 ? T(1e-30) : T(1e-13);
    if (off <= tol * diag || off < eps) break;
#pragma unroll
    for (int p = 0; p < NS - 1; ++p)
#pragma unroll
      for (int q = p + 1; q < NS; ++q) {
        const T apq = m[p][q];
        if (Num<T>::abs_(apq) > eps) {
          const T theta = (m[q][q] - m[p][p]) / (T(2) * apq);
          const T t = (theta >= T(0) ? T(1) : T(-1)) / (Num<T>::abs_(theta) + Num<T>::sqrt_(theta * theta + T(1)));
          const T c = T(1) / Num<T>::sqrt_(t * t + T(1));
          const T s = t * c;
          m[p][p] -= t * apq;
          m[q][q] += t * apq;
          m[p][q] = m[q][p] = T(0);
#pragma unroll
          for (int r = 0; r < NS; ++r) {
            if (r != p && r != q) {
              const T arp = m[r][p], arq = m[r][q];
              m[r][p] = m[p][r] = c * arp - s * arq;
              m[r][q] = m[q][r] = s * arp + c * arq;
            }
          }
        }
      }
  }
  T best = m[0][0];
#pragma unroll
  for (int i = 1; i < NS; ++i) best = m[i][i] > best ? m[i][i] : best;
  return best;
}

// r^2 = lambda_max(q B), B = I + k_fb^T k_fb  (utils.py:175-185).  For NS <= 2 the characteristic polynomial of the
// (possibly non-symmetric) product q B is solved in closed form, like the reference's general `eig`; for NS >= 3 a
// symmetric q (every shape matrix the path itself produces: asymmetry at rounding level) makes the product similar to
// the symmetric C^T q C (B = C C^T), whose spectrum Jacobi resolves to full relative accuracy, and a q that is NOT
// symmetric (an entry point may be handed one, test_utils.py:35-59) takes the general route of the reference: the
// largest real eigenvalue of q B by shifted QR (eig_small.cuh), NaN for a complex pair where the reference asserts.
template <typename T, int NS>
__device__ __forceinline__ T remainder_r_sqr(const T (&q)[NS][NS], const ProbDev<T>& pr) {
  if (NS == 1) return q[0][0] * pr.bmat[0][0];
  if (NS == 2) {
    T m00 = q[0][0] * pr.bmat[0][0] + q[0][1] * pr.bmat[1][0];
    T m01 = q[0][0] * pr.bmat[0][1] + q[0][1] * pr.bmat[1][1];
    T m10 = q[1][0] * pr.bmat[0][0] + q[1][1] * pr.bmat[1][0];
    T m11 = q[1][0] * pr.bmat[0][1] + q[1][1] * pr.bmat[1][1];
    const T half_tr = T(0.5) * (m00 + m11);
    const T half_diff = T(0.5) * (m00 - m11);
    const T disc = half_diff * half_diff + m01 * m10;  // = (tr/2)^2 - det, cancellation-free form
    return half_tr + Num<T>::sqrt_(disc);              // NaN if the pair is complex (reference asserts)
  }
  T tmp[NS][NS], m[NS][NS];
  T asym = T(0), scale = T(0);
#pragma unroll
  for (int i = 0; i < NS; ++i)
#pragma unroll
    for (int j = 0; j < NS; ++j) {
      scale = Num<T>::abs_(q[i][j]) > scale ? Num<T>::abs_(q[i][j]) : scale;
      if (j > i) asym = Num<T>::abs_(q[i][j] - q[j][i]) > asym ? Num<T>::abs_(q[i][j] - q[j][i]) : asym;
    }
  if (asym > (sizeof(T) == 8 ? T(1e-12) : T(1e-5)) * scale) {
    T mg[NS][NS];  // its own array: this one goes to local memory (the routine is not inlined), `m` below stays in registers
#pragma unroll
    for (int i = 0; i < NS; ++i)
#pragma unroll
      for (int j = 0; j < NS; ++j) {
        T s = T(0);
#pragma unroll
        for (int k = 0; k < NS; ++k) s += q[i][k] * pr.bmat[k][j];
        mg[i][j] = s;
      }
    return max_real_eig_general<T, NS>(mg, NS);
  }
#pragma unroll
  for (int i = 0; i < NS; ++i)
#pragma unroll
    for (int j = 0; j < NS; ++j) {
      T s = T(0);
#pragma unroll
      for (int k = j; k < NS; ++k) s += T(0.5) * (q[i][k] + q[k][i]) * pr.bchol[k][j];  // (q_sym C)[i][j]
      tmp[i][j] = s;
    }
#pragma unroll
  for (int i = 0; i < NS; ++i)
#pragma unroll
    for (int j = i; j < NS; ++j) {
      T s = T(0);
#pragma unroll
      for (int k = i; k < NS; ++k) s += pr.bchol[k][i] * tmp[k][j];  // (C^T tmp)[i][j]
      m[i][j] = m[j][i] = s;
    }
  return sym_lambda_max<T, NS>(m);
}

// _fix_zeros_nans (gp_reachability_pytorch.py:234-243): NaN is counted in flags[0] (the caller raises); an exact zero
// is counted in flags[1] and replaced by 1e-5.  The reference's rule is batch-wide: when the batch holds an exact zero
// it sets EVERY x <= 0 to 1e-5, so negative entries are clamped too -- `clamp_neg` says whether that applies (the
// se_reach_* entry points scan the batch first, var_zero_scan_kernel; se_cem_solve takes it from
// se_problem.clamp_nonpositive).  Without it a negative value turns into NaN at the next sqrt and is reported there,
// which is what the reference does when no exact zero is present.
template <typename T>
__device__ __forceinline__ T fix_zero_nan(T x, int32_t* flags, bool clamp_neg = false) {
  if (x != x) {
    if (flags) atomicAdd(&flags[0], 1);
    return x;
  }
  if (x == T(0) || (clamp_neg && x < T(0))) {
    if (flags) atomicAdd(&flags[1], 1);
    return T(1e-5);
  }
  return x;
}

// flags[2] |= 1 when var[0 .. count) holds an exact zero (the batch-wide condition of _fix_zeros_nans)
template <typename T>
__global__ void var_zero_scan_kernel(const T* __restrict__ var, int64_t count, int32_t* __restrict__ flags) {
  const int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  const bool z = i < count && var[i] == T(0);
  if (__any_sync(0xffffffffu, z) && (threadIdx.x & 31) == 0) atomicOr(&flags[2], 1);
}

// support-function distance of the ellipsoid (p, q) to every face; returns true iff all d < 0
// (gp_reachability_pytorch.py:184-231; c_safety = 1 on the CEM path, quirk Q2)
template <typename T, int NS, int MAXF>
__device__ __forceinline__ bool inside_faces(const T (&p)[NS], const T (&q)[NS][NS], const T (&h_mat)[MAXF][SE_MAX_NS],
                                             const T (&h_vec)[MAXF], int m, T c_safety, T* d_out) {
  bool inside = true;
  for (int f = 0; f < m; ++f) {
    T hp = T(0), hqh = T(0);
#pragma unroll
    for (int i = 0; i < NS; ++i) {
      hp += h_mat[f][i] * p[i];
      T qh = T(0);
#pragma unroll
      for (int j = 0; j < NS; ++j) qh += q[i][j] * h_mat[f][j];
      hqh += h_mat[f][i] * qh;
    }
    const T d = hp + c_safety * Num<T>::sqrt_(hqh) - h_vec[f];
    if (d_out) d_out[f] = d;
    if (d >= T(0)) inside = false;  // (d_safety >= 0).sum() == 0, gp_reachability_pytorch.py:230-231
  }
  return inside;
}
template <typename T, int NS>
__device__ __forceinline__ bool inside_polytope(const T (&p)[NS], const T (&q)[NS][NS], const ProbDev<T>& pr,
                                                T c_safety, T* d_out) {
  return inside_faces<T, NS, SE_MAX_FACES>(p, q, pr.h_mat, pr.h_vec, pr.m_safe, c_safety, d_out);
}

template <typename T>
struct StepIO {
  // state (in / out, updated in place unless out pointers are given)
  const T* p_in;   // [N, NS]
  const T* q_in;   // [N, NS, NS] (ignored when point != 0)
  T* p_out;        // [N, NS]
  T* q_out;        // [N, NS, NS]
  const T* u;      // actions, row stride u_stride
  int64_t u_stride;
  const T* mean;   // [N, NS]
  const T* var;    // [N, NS]
  const T* jac;    // [N, NS, D] (unused in the point branch)
  T* sigma_out;    // [N, NS] or null
  int point;       // 1 = all states are points (q = None branch)
  // CEM bookkeeping (null when called from se_reach_*)
  T* objective;    // [N] running sum
  T* constraint;   // [N] written at the last step (and accumulated on every step if all_states)
  int first_step, last_step;
  const T* act_all;     // [N, T, NU] full action sequences for the action constraint (last step only)
  int T_len;
  // optional dumps
  T* traj;         // [N, T+1, NS+NS*NS] or null
  int t_index;     // step index (trajectory slot t_index + 1)
  T* p_all;        // [N, T, NS] or null (multistep API)
  T* q_all;        // [N, T, NS, NS]
  T* sigma_all;    // [N, T, NS]
  int32_t* flags;
  int batch_zero_scan;  // 1: flags[2] holds "this batch has an exact-zero variance" (set by var_zero_scan_kernel)
  int64_t first;   // first candidate of this launch (all arrays are indexed by the absolute candidate id)
  // Fused variance reduction (se_cem_solve): when `partial` is set, the kernel forms var = s + noise - sum_ib partial
  // itself (what gp_var_kernel does, same order) instead of reading `var`, and, when the kernel-row launch was split
  // (nseg > 0), also the mean / Jacobian from the per-segment sums in kpart.  Rows are local: candidate c <-> c - first.
  const T* partial;  // [ns][nb][rows_pad]
  int nb, rows_pad;
  const T* kpart;    // [nseg][rows_pad][ns][1 + d]
  int nseg;
  T gp_outputscale[SE_MAX_NS], gp_noise[SE_MAX_NS], gp_inv_l2[SE_MAX_NS][SE_MAX_D];
};

// One ellipsoid step on register operands (gp_reachability_pytorch.py:18-181): (p, q | point) + GP posterior at
// [p, u] -> (p1, q1), sg = the variances after _fix_zeros_nans.  Shared by reach_step_kernel (one thread per candidate)
// and the fused small-model rollout kernel (rollout_kernels.cuh).
template <typename T, int NS, int NU>
__device__ __forceinline__ void ellipsoid_step_core(const ProbDev<T>& pr, const T (&p)[NS], const T (&q)[NS][NS],
                                                    bool point, const T (&u)[NU], const T (&mu)[NS],
                                                    const T (&var)[NS], const T (&jac)[NS][NS + NU], int32_t* flags,
                                                    T (&p1)[NS], T (&q1)[NS][NS], T (&sg)[NS],
                                                    bool clamp_neg = false) {
  clamp_neg = clamp_neg || pr.clamp_nonpos != 0;
#pragma unroll
  for (int i = 0; i < NS; ++i) sg[i] = fix_zero_nan<T>(var[i], flags, clamp_neg);  // sigm_0 (:76, :117)

  // p_0 = mu_0 + a p + b u   (:93-94, :132)
#pragma unroll
  for (int i = 0; i < NS; ++i) {
    T s = mu[i];
#pragma unroll
    for (int k = 0; k < NS; ++k) s += pr.a[i][k] * p[k];
#pragma unroll
    for (int k = 0; k < NU; ++k) s += pr.b[i][k] * u[k];
    p1[i] = s;
  }

  if (point) {
    // q_1 = diag(n_s (c_safety sqrt(sigm_0))^2)   (:82-91, utils_ellipsoid.py:302-309)
#pragma unroll
    for (int i = 0; i < NS; ++i)
#pragma unroll
      for (int j = 0; j < NS; ++j) q1[i][j] = T(0);
#pragma unroll
    for (int i = 0; i < NS; ++i) {
      const T rk = fix_zero_nan<T>(pr.c_safety * Num<T>::sqrt_(sg[i]), flags);
      q1[i][i] = T(NS) * rk * rk;
    }
  } else {
    T h[NS][NS];
    // H = a + a_mu + (b_mu + b) k_fb   (:127-131)
#pragma unroll
    for (int i = 0; i < NS; ++i)
#pragma unroll
      for (int j = 0; j < NS; ++j) {
        T s = pr.a[i][j] + jac[i][j];
#pragma unroll
        for (int k = 0; k < NU; ++k) s += (jac[i][NS + k] + pr.b[i][k]) * pr.k_fb[k][j];
        h[i][j] = s;
      }
    // Q_0 = H q H^T   (:134)
    T hq[NS][NS];
#pragma unroll
    for (int i = 0; i < NS; ++i)
#pragma unroll
      for (int j = 0; j < NS; ++j) {
        T s = T(0);
#pragma unroll
        for (int k = 0; k < NS; ++k) s += h[i][k] * q[k][j];
        hq[i][j] = s;
      }
    T tr_q0 = T(0);
#pragma unroll
    for (int i = 0; i < NS; ++i)
#pragma unroll
      for (int j = 0; j < NS; ++j) {
        T s = T(0);
#pragma unroll
        for (int k = 0; k < NS; ++k) s += hq[i][k] * h[j][k];
        q1[i][j] = s;  // holds Q_0 for now
        if (i == j) tr_q0 += s;
      }
    // Lagrange remainder boxes   (:142-162, utils.py:152-194)
    const T r_sqr = remainder_r_sqr<T, NS>(q, pr);
    const T r = Num<T>::sqrt_(r_sqr);
    T d_sig[NS], d_mu[NS];
    T tr_sig = T(0), tr_mu = T(0);
#pragma unroll
    for (int i = 0; i < NS; ++i) {
      const T ub_mean = pr.l_mu[i] * r_sqr;
      const T ub_sigma = pr.l_sigma[i] * r;
      const T bse = fix_zero_nan<T>(pr.c_safety * (Num<T>::sqrt_(sg[i]) + ub_sigma), flags);
      d_sig[i] = T(NS) * bse * bse;
      d_mu[i] = T(NS) * ub_mean * ub_mean;
      tr_sig += d_sig[i];
      tr_mu += d_mu[i];
    }
    // Q_L = Q_sigma (+) Q_mu ; q_1 = Q_L (+) Q_0   (:169-172, utils_ellipsoid.py:133-138; no guards, quirk Q8)
    const T c1 = Num<T>::sqrt_(tr_sig / tr_mu);
    T tr_l = T(0);
    T d_l[NS];
#pragma unroll
    for (int i = 0; i < NS; ++i) {
      d_l[i] = (T(1) + T(1) / c1) * d_sig[i] + (T(1) + c1) * d_mu[i];
      tr_l += d_l[i];
    }
    const T c2 = Num<T>::sqrt_(tr_l / tr_q0);
    const T f_l = T(1) + T(1) / c2, f_0 = T(1) + c2;
#pragma unroll
    for (int i = 0; i < NS; ++i)
#pragma unroll
      for (int j = 0; j < NS; ++j) q1[i][j] = f_0 * q1[i][j] + (i == j ? f_l * d_l[i] : T(0));
  }
}

// cost of one step for the CEM objective (safempc_cem.py:304-312; environments.py:505-510; lunarlander.py:111-113)
template <typename T, int NS>
__device__ __forceinline__ T step_objective(const ProbDev<T>& pr, const T (&p1)[NS], const T (&sg)[NS]) {
  if (pr.objective == SE_OBJ_NEG_VARIANCE) {
    T cost = T(0);
#pragma unroll
    for (int i = 0; i < NS; ++i) cost -= sg[i];  // -sum of variances incl. noise (quirk Q7)
    return cost;
  }
  if (pr.objective == SE_OBJ_ABS_STATE1) return Num<T>::abs_(pr.objective_theta - p1[NS > 1 ? 1 : 0]);
  return -p1[NS - 1];
}

// constraint cost added by the ellipsoid (p1, q1) produced by step t (last = terminal step): safety polytope on the
// terminal (or every) ellipsoid, obstacle polytope on the intermediate ones, action box at the end
template <typename T, int NS, int NU>
__device__ __forceinline__ T step_constraint(const ProbDev<T>& pr, const T (&p1)[NS], const T (&q1)[NS][NS], bool last,
                                             const T* act_row, int t_len) {
  T cc = T(0);
  if (pr.m_safe > 0 && (last || pr.all_states)) {
    const bool inside = inside_polytope<T, NS>(p1, q1, pr, T(1), nullptr);
    if (!inside) cc += pr.state_penalty;  // 10 * (1 - inside), safempc_cem.py:132
  }
  if (pr.m_obs > 0 && !last) {
    // intermediate ellipsoids against the obstacle polytope (safempc_simple.py:373-383, as a penalty)
    if (!inside_faces<T, NS, SE_MAX_OBS_FACES>(p1, q1, pr.hobs_mat, pr.hobs_vec, pr.m_obs, T(1), nullptr))
      cc += pr.state_penalty;
  }
  if (last && pr.m_act > 0) {
    // ActionConstraint: penalty * #{t : some face has A u_t - b > 0}  (constrained_cem_mpc.py:29-41, 100-110)
    int bad = 0;
    for (int t = 0; t < t_len; ++t) {
      bool out = false;
      for (int f = 0; f < pr.m_act; ++f) {
        T s = -pr.act_b[f];
#pragma unroll
        for (int k = 0; k < NU; ++k) s += pr.act_a[f][k] * act_row[t * NU + k];
        if (s > T(0)) out = true;
      }
      bad += out ? 1 : 0;
    }
    cc += pr.act_penalty * T(bad);
  }
  return cc;
}

// The ellipsoid step of candidate c with everything around it (variance / mean / Jacobian sums, outputs, CEM
// bookkeeping).  reach_step_kernel runs it with one thread per candidate; the fused kernel-row kernel of small batches
// (gp_kstar_slices8_kernel<..., FUSE>) runs it in every lane that works on the candidate, with `writer` set in exactly
// one of them: the others only need the new centre p1.
template <typename T, int NS, int NU>
__device__ __forceinline__ void reach_step_body(const ProbDev<T>& pr, const StepIO<T>& io, int64_t c, bool writer,
                                                T (&p1)[NS]) {
  constexpr int D = NS + NU;

  T p[NS], u[NU], mu[NS], var[NS], sg[NS];
  T q[NS][NS], jac[NS][D];
  const bool split_sums = io.partial && io.nseg > 0;
  // The sums below add in ascending block / segment order (what gp_var_kernel does), but the loads of a batch are
  // issued together for every output: a thread waits once per 8 blocks instead of once per block -- the kernel is a
  // pure latency chain (one thread per candidate), and these loads were most of it.
  if (io.partial) {  // gp_var_kernel, fused
    const int64_t r = c - io.first;
    T s[NS];
#pragma unroll
    for (int i = 0; i < NS; ++i) s[i] = T(0);
    for (int ib0 = 0; ib0 < io.nb; ib0 += 8) {
      T v[NS][8];
#pragma unroll
      for (int i = 0; i < NS; ++i)
#pragma unroll
        for (int k = 0; k < 8; ++k)
          v[i][k] = ib0 + k < io.nb ? io.partial[((int64_t)i * io.nb + ib0 + k) * io.rows_pad + r] : T(0);
#pragma unroll
      for (int i = 0; i < NS; ++i)
#pragma unroll
        for (int k = 0; k < 8; ++k) s[i] += v[i][k];  // + 0 for the blocks past nb: exact
    }
#pragma unroll
    for (int i = 0; i < NS; ++i) var[i] = io.gp_outputscale[i] - s[i] + io.gp_noise[i];
  } else {
#pragma unroll
    for (int i = 0; i < NS; ++i) var[i] = io.var[c * NS + i];
  }
  if (split_sums) {
    constexpr int W = NS * (1 + D);  // values per (segment, candidate): mean and Jacobian sums of every output
    const T* src = io.kpart + (c - io.first) * W;
    const int64_t seg_stride = (int64_t)io.rows_pad * W;
    T acc[W];
#pragma unroll
    for (int e = 0; e < W; ++e) acc[e] = T(0);
    for (int sg0 = 0; sg0 < io.nseg; sg0 += 2) {
      T v[2][W];
#pragma unroll
      for (int k = 0; k < 2; ++k)
#pragma unroll
        for (int e = 0; e < W; ++e) {
          const bool want = sg0 + k < io.nseg && (!io.point || e % (1 + D) == 0);  // point branch: means only
          v[k][e] = want ? src[(sg0 + k) * seg_stride + e] : T(0);
        }
#pragma unroll
      for (int k = 0; k < 2; ++k)
#pragma unroll
        for (int e = 0; e < W; ++e) acc[e] += v[k][e];
    }
#pragma unroll
    for (int i = 0; i < NS; ++i) {
      mu[i] = acc[i * (1 + D)];
#pragma unroll
      for (int j = 0; j < D; ++j) jac[i][j] = -acc[i * (1 + D) + 1 + j] * io.gp_inv_l2[i][j];
    }
  } else {
#pragma unroll
    for (int i = 0; i < NS; ++i) {
      mu[i] = io.mean[c * NS + i];
#pragma unroll
      for (int j = 0; j < D; ++j) jac[i][j] = io.point ? T(0) : io.jac[(c * NS + i) * D + j];
    }
  }
#pragma unroll
  for (int i = 0; i < NS; ++i) {
    p[i] = io.p_in[c * NS + i];
#pragma unroll
    for (int j = 0; j < NS; ++j) q[i][j] = io.point ? T(0) : io.q_in[(c * NS + i) * NS + j];
  }
#pragma unroll
  for (int j = 0; j < NU; ++j) u[j] = io.u[c * io.u_stride + j];
  // non-finite inputs poison the mean (kernel-row kernels); the reference's GP returns NaN variances there as well
#pragma unroll
  for (int i = 0; i < NS; ++i)
    if (mu[i] != mu[i]) var[i] = mu[i];
  T q1[NS][NS];
  const bool clamp_neg = io.batch_zero_scan && io.flags && io.flags[2] != 0;
  ellipsoid_step_core<T, NS, NU>(pr, p, q, io.point != 0, u, mu, var, jac, writer ? io.flags : nullptr, p1, q1, sg,
                                 clamp_neg);
  if (!writer) return;

  // ---- write the new ellipsoid ----
#pragma unroll
  for (int i = 0; i < NS; ++i) {
    io.p_out[c * NS + i] = p1[i];
    if (io.sigma_out) io.sigma_out[c * NS + i] = sg[i];
#pragma unroll
    for (int j = 0; j < NS; ++j) io.q_out[(c * NS + i) * NS + j] = q1[i][j];
  }
  if (io.p_all) {
    const int64_t base = c * io.T_len + io.t_index;
#pragma unroll
    for (int i = 0; i < NS; ++i) {
      io.p_all[base * NS + i] = p1[i];
      if (io.sigma_all) io.sigma_all[base * NS + i] = sg[i];
    }
  } else if (io.sigma_all) {
    const int64_t base = c * io.T_len + io.t_index;
#pragma unroll
    for (int i = 0; i < NS; ++i) io.sigma_all[base * NS + i] = sg[i];
  }
  if (io.q_all) {
    const int64_t base = c * io.T_len + io.t_index;
#pragma unroll
    for (int i = 0; i < NS; ++i)
#pragma unroll
      for (int j = 0; j < NS; ++j) io.q_all[(base * NS + i) * NS + j] = q1[i][j];
  }
  if (io.traj) {
    constexpr int S = NS + NS * NS;
    T* dst = io.traj + (c * (io.T_len + 1) + io.t_index + 1) * S;
#pragma unroll
    for (int i = 0; i < NS; ++i) dst[i] = p1[i];
#pragma unroll
    for (int i = 0; i < NS; ++i)
#pragma unroll
      for (int j = 0; j < NS; ++j) dst[NS + i * NS + j] = q1[i][j];
    if (io.first_step) {  // slot 0 = [p_0 | 0]  (PQFlattener.flatten(p, None), safempc_cem.py:40-54)
      T* d0 = io.traj + c * (io.T_len + 1) * S;
#pragma unroll
      for (int i = 0; i < NS; ++i) d0[i] = p[i];
#pragma unroll
      for (int i = 0; i < NS * NS; ++i) d0[NS + i] = T(0);
    }
  }

  // ---- CEM bookkeeping ----
  if (io.objective) {
    const T prev = io.first_step ? T(0) : io.objective[c];
    io.objective[c] = prev + step_objective<T, NS>(pr, p1, sg);
  }
  if (io.constraint) {
    const T cc = io.first_step ? T(0) : io.constraint[c];
    io.constraint[c] = cc + step_constraint<T, NS, NU>(pr, p1, q1, io.last_step != 0,
                                                       io.act_all ? io.act_all + c * io.T_len * NU : nullptr, io.T_len);
  }
}

template <typename T, int NS, int NU>
__global__ void __launch_bounds__(128)
reach_step_kernel(ProbDev<T> pr, StepIO<T> io, int64_t N) {
  const int64_t c = io.first + blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (c >= io.first + N) return;
  T p1[NS];
  reach_step_body<T, NS, NU>(pr, io, c, true, p1);
}

// stand-alone polytope distance (se_safety_distance)
template <typename T, int NS>
__global__ void safety_distance_kernel(ProbDev<T> pr, const T* __restrict__ p, const T* __restrict__ q, int64_t N,
                                       T* __restrict__ d, uint8_t* __restrict__ inside) {
  const int64_t c = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (c >= N) return;
  T pp[NS], qq[NS][NS];
#pragma unroll
  for (int i = 0; i < NS; ++i) {
    pp[i] = p[c * NS + i];
#pragma unroll
    for (int j = 0; j < NS; ++j) qq[i][j] = q[(c * NS + i) * NS + j];
  }
  T dd[SE_MAX_FACES];
  const bool in = inside_polytope<T, NS>(pp, qq, pr, pr.c_safety, dd);
  if (d)
    for (int f = 0; f < pr.m_safe; ++f) d[c * pr.m_safe + f] = dd[f];
  if (inside) inside[c] = in ? 1 : 0;
}

}  // namespace se
