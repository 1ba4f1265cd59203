// rollout_kernels.cuh -- the whole rollout of a candidate in ONE kernel, for small models (n_train <= 128).
//
// RolloutFunction._sample_trajectory (constrained_cem_mpc.py:158-185) calls the dynamics T times per candidate; for the
// model sizes the reference itself runs (n_train ~ 100, N = 20 .. 500, T = 5) a horizon step is ~2 x 10^4 FMA per
// candidate and the four-kernel step (kernel rows -> variance GEMM -> variance -> ellipsoid) is pure launch and
// pipeline latency (ncu, C1: 12.8 + 7.0 + 3.1 + 5.6 us per step for 0.2 us of arithmetic).  Candidates are independent
// until elite selection, so here ONE CTA carries one candidate through all T steps, the ellipsoid in registers.  With
// only N ~ 500 candidates in flight the kernel is a latency chain, so the work of a step is spread as wide as it goes:
//
//   threads 0 .. 127 own training point j          kernel rows k_o(z, x_j) -> shared memory; partial mean / Jacobian
//   5 warps own the 10 lower 32 x 32 blocks         v_i += L^-1[i][j] k_j: 32 independent coalesced loads of L^-T per
//   (jb <= m) of L^-1                               thread (all in flight at once), k_j broadcast from shared memory
//   threads 0 .. 127 own row i                      v_i = sum of its blocks' partials, |v|^2 -> var = s + noise - |v|^2
//   every thread (redundantly, no divergence)       the ellipsoid step, objective and constraint bookkeeping
//
// One launch per CEM iteration replaces 4 T launches.  Arithmetic and reference citations are those of gp_kernels.cuh and
// reach_kernels.cuh (ellipsoid_step_core is shared with reach_step_kernel).
#pragma once
#include "cem_kernels.cuh"
#include "common.cuh"
#include "reach_kernels.cuh"

namespace se {

constexpr int kSmallN = 128;        // n_pad of every model with n_train <= 128 (kTile)
constexpr int kRolloutBlocks = 10;  // lower 32 x 32 blocks of the 128 x 128 factor
constexpr int kRolloutWarps = 5;    // two blocks per warp: 160 threads, 4 CTAs per SM -> 592 candidates in flight

template <typename T>
struct RolloutIO {
  ActionSrc<T> src;   // where the candidates come from: the kernel draws its own action sequence (cem_sample fused in)
  int first, n_local; // this rank's shard of every instance: local candidate cl = inst * n_local + l, global id first + l
  const T* p0;        // [M, NS]      initial point state of every instance
  int T_len;
  T* pack;            // [2][N]       (objective, constraint) planes, the all-gather send buffer (cem_pack_costs fused in)
  T* rec_obj;         // [N] or null  per-iteration records
  T* rec_con;
  T* rec_actions;     // [N, T, NU] or null
  T* traj;            // [N, T+1, NS + NS*NS] or null
  int32_t* flags;
  int64_t N;          // local candidates over all instances
};

// wt[o][j][i] = w[o][i][j]: L^-T, so that a warp reads 32 consecutive ROWS of column j in one transaction
template <typename T>
__global__ void transpose_w_kernel(const T* __restrict__ w, int n_pad, T* __restrict__ wt) {
  __shared__ T tile[32][33];
  const int o = blockIdx.z, i0 = blockIdx.y * 32, j0 = blockIdx.x * 32;
  const T* src = w + (int64_t)o * n_pad * n_pad;
  T* dst = wt + (int64_t)o * n_pad * n_pad;
  for (int r = threadIdx.y; r < 32; r += blockDim.y) tile[r][threadIdx.x] = src[(int64_t)(i0 + r) * n_pad + j0 + threadIdx.x];
  __syncthreads();
  for (int r = threadIdx.y; r < 32; r += blockDim.y) dst[(int64_t)(j0 + r) * n_pad + i0 + threadIdx.x] = tile[threadIdx.x][r];
}

template <typename T>
__device__ __forceinline__ T warp_sum(T v) {
#pragma unroll
  for (int s = 16; s > 0; s >>= 1) v += __shfl_xor_sync(0xffffffffu, v, s);
  return v;
}

template <typename T, int NS, int NU>
__global__ void __launch_bounds__(32 * kRolloutWarps, 4)
rollout_small_kernel(GpDev<T> gp, const T* __restrict__ wt, ProbDev<T> pr, RolloutIO<T> io) {
  constexpr int D = NS + NU, NP = kSmallN, M = NP / 32, NRED = NS * (1 + D);
  static_assert(M == 4 && kRolloutBlocks == M * (M + 1) / 2 && kRolloutBlocks == 2 * kRolloutWarps && kRolloutWarps > M,
                "two lower blocks of a 4 x 4 block triangle per warp; warps 0 .. 3 also own the kernel rows");
  __shared__ T ks[NS][NP];                      // kernel rows of this step
  __shared__ T red[M][NRED];                    // per-warp partial sums of mean and Jacobian
  __shared__ T part[NS][kRolloutBlocks][32];    // per-block partial products L^-1[i][jb block] k[jb block]
  __shared__ T ssq[NS][M];                      // per-row-block sums of v_i^2
  __shared__ T pbuf[NS];                        // centre of the current ellipsoid (warp 0 -> kernel-row threads)
  __shared__ T act[SE_MAX_T * NU];              // this candidate's action sequence
  const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
  const int64_t c = blockIdx.x;
  int32_t* const flags = tid == 0 ? io.flags : nullptr;  // every thread walks the same ellipsoid: count once
  // blocks (jb, m), jb <= m, in the order (0,0) (0,1) (0,2) (0,3) (1,1) (1,2) (1,3) (2,2) (2,3) (3,3): warp w owns 2w, 2w+1
  const int first_of_jb[M] = {0, 4, 7, 9};

  const int inst = (int)(c / io.n_local), gid = io.first + (int)(c % io.n_local);
  // actions = means + stds * eps (constrained_cem_mpc.py:167), drawn here instead of by a separate launch
  for (int e = tid; e < io.T_len * NU; e += blockDim.x) {
    const T a = candidate_action<T>(io.src, inst, gid, e);
    act[e] = a;
    if (io.rec_actions) io.rec_actions[c * io.T_len * NU + e] = a;
  }
  __syncthreads();
  T p[NS], q[NS][NS];
#pragma unroll
  for (int i = 0; i < NS; ++i) {
    p[i] = io.p0[inst * NS + i];
#pragma unroll
    for (int j = 0; j < NS; ++j) q[i][j] = T(0);
  }
  T nh_inv_l2[NS][D];
#pragma unroll
  for (int o = 0; o < NS; ++o)
#pragma unroll
    for (int dd = 0; dd < D; ++dd) nh_inv_l2[o][dd] = T(-0.5) * gp.inv_l2[o][dd];
  if (io.traj && tid == 0) {  // slot 0 = [p_0 | 0]  (PQFlattener.flatten(p, None), safempc_cem.py:40-54)
    T* d0 = io.traj + c * (io.T_len + 1) * (NS + NS * NS);
#pragma unroll
    for (int i = 0; i < NS; ++i) d0[i] = p[i];
#pragma unroll
    for (int i = 0; i < NS * NS; ++i) d0[NS + i] = T(0);
  }

  T obj = T(0), cc = T(0);
#pragma unroll 1
  for (int t = 0; t < io.T_len; ++t) {
    const bool point = t == 0;  // get_action starts from a point state (safempc_cem.py:234-235, quirk Q9)
    T u[NU], z[D];
#pragma unroll
    for (int k = 0; k < NU; ++k) u[k] = act[t * NU + k];
#pragma unroll
    for (int dd = 0; dd < D; ++dd) z[dd] = dd < NS ? (point ? p[dd] : pbuf[dd]) : u[dd - NS];

    // ---- kernel rows, mean, mean-Jacobian (gp_ssm_cem.py:45-94; ssm_pytorch/utilities.py:54-85) ----
    if (tid < NP) {
      const int j = tid;
      T diff[D], sq[D], vals[NRED];
#pragma unroll
      for (int dd = 0; dd < D; ++dd) {
        diff[dd] = z[dd] - gp.xt[dd * NP + j];
        sq[dd] = diff[dd] * diff[dd];
      }
#pragma unroll
      for (int o = 0; o < NS; ++o) {
        T arg = T(0);
#pragma unroll
        for (int dd = 0; dd < D; ++dd) arg += sq[dd] * nh_inv_l2[o][dd];
        const T k = gp.outputscale[o] * Num<T>::exp_(arg);
        ks[o][j] = k;
        const T wk = gp.alpha[o * NP + j] * k;  // alpha is zero on padded training points
        vals[o * (1 + D)] = wk;
#pragma unroll
        for (int dd = 0; dd < D; ++dd) vals[o * (1 + D) + 1 + dd] = wk * diff[dd];
      }
#pragma unroll
      for (int r = 0; r < NRED; ++r) {
        const T v = warp_sum(vals[r]);
        if (lane == 0) red[wid][r] = v;
      }
    }
    __syncthreads();
    // ---- variance s + noise - || L^-1 k ||^2: this warp's 32 x 32 block, rows 32 mb + lane, columns 32 jb .. ----
#pragma unroll
    for (int bi = 0; bi < 2; ++bi) {
      const int blk = 2 * wid + bi;
      const int jb = blk < 4 ? 0 : blk < 7 ? 1 : blk < 9 ? 2 : 3;
      const int mb = blk < 4 ? blk : blk < 7 ? blk - 3 : blk < 9 ? blk - 5 : 3;
#pragma unroll
      for (int o = 0; o < NS; ++o) {
        const T* const wo = wt + ((int64_t)o * NP + 32 * jb) * NP + 32 * mb + lane;
        T wv[32];
#pragma unroll
        for (int jj = 0; jj < 32; ++jj) wv[jj] = wo[jj * NP];  // 32 independent coalesced loads in flight
        T v0 = T(0), v1 = T(0);
#pragma unroll
        for (int jj = 0; jj < 32; jj += 2) {
          v0 += wv[jj] * ks[o][32 * jb + jj];
          v1 += wv[jj + 1] * ks[o][32 * jb + jj + 1];
        }
        part[o][blk][lane] = v0 + v1;
      }
    }
    __syncthreads();
    if (tid < NP) {
      const int m = wid;  // row tid = 32 m + lane: blocks (0, m) .. (m, m)
#pragma unroll
      for (int o = 0; o < NS; ++o) {
        T v = T(0);
#pragma unroll
        for (int b = 0; b < M; ++b)
          if (b <= m) v += part[o][first_of_jb[b] + (m - b)][lane];
        const T s2 = warp_sum(v * v);
        if (lane == 0) ssq[o][m] = s2;
      }
    }
    __syncthreads();
    // ---- ellipsoid step, costs: a serial chain of sqrt / divide, one warp walks it (lanes redundantly) ----
    if (wid == 0) {
      T mean[NS], var[NS], jac[NS][D];
#pragma unroll
      for (int o = 0; o < NS; ++o) {
        mean[o] = (red[0][o * (1 + D)] + red[1][o * (1 + D)]) + (red[2][o * (1 + D)] + red[3][o * (1 + D)]);
#pragma unroll
        for (int dd = 0; dd < D; ++dd) {
          const int r = o * (1 + D) + 1 + dd;
          jac[o][dd] = point ? T(0) : -((red[0][r] + red[1][r]) + (red[2][r] + red[3][r])) * gp.inv_l2[o][dd];
        }
        var[o] = gp.outputscale[o] + gp.noise[o] - ((ssq[o][0] + ssq[o][1]) + (ssq[o][2] + ssq[o][3]));
      }
      T p1[NS], q1[NS][NS], sg[NS];
      ellipsoid_step_core<T, NS, NU>(pr, p, q, point, u, mean, var, jac, flags, p1, q1, sg);
      obj += step_objective<T, NS>(pr, p1, sg);
      cc += step_constraint<T, NS, NU>(pr, p1, q1, t == io.T_len - 1, act, io.T_len);
      if (io.traj && tid == 0) {
        T* dst = io.traj + (c * (io.T_len + 1) + t + 1) * (NS + NS * NS);
#pragma unroll
        for (int i = 0; i < NS; ++i) dst[i] = p1[i];
#pragma unroll
        for (int i = 0; i < NS; ++i)
#pragma unroll
          for (int j = 0; j < NS; ++j) dst[NS + i * NS + j] = q1[i][j];
      }
#pragma unroll
      for (int i = 0; i < NS; ++i) {
        p[i] = p1[i];
        if (lane == 0) pbuf[i] = p1[i];
#pragma unroll
        for (int j = 0; j < NS; ++j) q[i][j] = q1[i][j];
      }
    }
    // publishes pbuf; also keeps a fast warp from overwriting ks / red / part / ssq while warp 0 still reads them
    __syncthreads();
  }
  if (tid == 0) {
    io.pack[c] = obj;
    io.pack[io.N + c] = cc;
    if (io.rec_obj) io.rec_obj[c] = obj;
    if (io.rec_con) io.rec_con[c] = cc;
  }
}

}  // namespace se
