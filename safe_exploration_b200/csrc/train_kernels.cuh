// train_kernels.cuh -- the model-refresh side of the path: covariance assembly and the
// gradient of the exact marginal log likelihood for the n_s independent RBF-ARD GPs of GpCemSSM
// (safe_exploration/ssm_cem/gp_ssm_cem.py:45-57 kernel structure, :103-129 training loop; the arithmetic itself is
// inside gpytorch 0.3.2 in the reference: ScaleKernel(RBFKernel(ard)) + GaussianLikelihood + ExactMarginalLogLikelihood).
//
// Both kernels walk the n x n covariance in 64 x 64 tiles, one 256-thread CTA per tile: thread (ty, tx) owns column
// j0 + tx and rows i0 + ty, ty + 4, ...; the 64 + 64 training inputs of the tile sit in shared memory; every global
// access of a warp is one contiguous 256-byte run of a matrix row.
#pragma once
#include "common.cuh"

namespace se {

constexpr int kTrainTile = 64;
constexpr int kTrainThreads = 256;

struct TrainHyp {
  double neg_half_inv_l2[SE_MAX_NS][SE_MAX_D];  // -1 / (2 l_od^2)
  double inv_l3[SE_MAX_NS][SE_MAX_D];           // 1 / l_od^3
  double lengthscale[SE_MAX_NS][SE_MAX_D];
  double outputscale[SE_MAX_NS];
  double noise[SE_MAX_NS];
};

// K[o][i][j] = s_o exp(-1/2 sum_d ((x_id - x_jd) / l_od)^2) + noise_o [i == j], the full (exactly symmetric) matrix.
// grid (tiles, tiles, n_s).  Bound by the FP64 pipe (the exp) and the 8 B per element it writes.
// K feeds a Cholesky factorisation whose condition number (~ n s / sigma^2) amplifies every ulp of K into the posterior
// variance (measured: a 2-ulp exp moves a variance that is 20x below the prior by 5e-9 relative).  So this kernel -- a
// few per cent of the factorisation's time either way -- does not use the branch-free exp_nb / fused-multiply-add
// argument of the kernel-row kernels: it evaluates the expression in the order a plain float64 array program does,
// scaled inputs x / l first, then differences, squares added in dimension order, exp from the library (<= 1 ulp),
// with no FMA contraction, so that K agrees with such a program (the float64 oracle, the reference's torch code) to
// the last bit of everything except the exponential itself.
template <int D>
__global__ void __launch_bounds__(kTrainThreads) gp_kernel_matrix_kernel(const double* __restrict__ x, int n,
                                                                         TrainHyp hyp, double* __restrict__ k) {
  __shared__ double xi[kTrainTile][D], xj[kTrainTile][D + 1];
  const int o = blockIdx.z, i0 = blockIdx.y * kTrainTile, j0 = blockIdx.x * kTrainTile;
  for (int e = threadIdx.x; e < kTrainTile * D; e += kTrainThreads) {
    const int r = e / D, c = e % D;
    const double l = hyp.lengthscale[o][c];
    xi[r][c] = (i0 + r < n) ? x[(int64_t)(i0 + r) * D + c] / l : 0.0;
    xj[r][c] = (j0 + r < n) ? x[(int64_t)(j0 + r) * D + c] / l : 0.0;
  }
  __syncthreads();
  const int tx = threadIdx.x % kTrainTile, ty = threadIdx.x / kTrainTile;
  const int j = j0 + tx;
  if (j >= n) return;
  double mine[D];
#pragma unroll
  for (int c = 0; c < D; ++c) mine[c] = xj[tx][c];
  const double s = hyp.outputscale[o], nz = hyp.noise[o];
  double* ko = k + (int64_t)o * n * n;
#pragma unroll 4
  for (int r = ty; r < kTrainTile; r += kTrainThreads / kTrainTile) {
    const int i = i0 + r;
    if (i >= n) break;
    double d2 = 0.0;
#pragma unroll
    for (int c = 0; c < D; ++c) {
      const double df = __dsub_rn(xi[r][c], mine[c]);
      d2 = __dadd_rn(d2, __dmul_rn(df, df));
    }
    const double kv = __dmul_rn(s, exp(__dmul_rn(-0.5, d2)));
    ko[(int64_t)i * n + j] = i == j ? __dadd_rn(kv, nz) : kv;
  }
}

// Per-tile partial sums of   sum_ij G_ij dK_ij/dtheta,   G = alpha alpha^T - (K + noise I)^-1   (both symmetric), over
// the lower-triangular tiles only (off-diagonal elements weigh 2):
//   part[o][tile][c < D]  = sum w g e (x_ic - x_jc)^2        e = exp(-1/2 sum_d ((x_id - x_jd)/l_od)^2)
//   part[o][tile][D]      = sum w g e
//   part[o][tile][D + 1]  = sum_i g_ii
// grid (tiles (tiles + 1) / 2, n_s).  Reads the lower triangle of K^-1 once (4 n_s n^2 bytes) and recomputes e.
template <int D>
__global__ void __launch_bounds__(kTrainThreads) gp_mll_grad_tiles_kernel(const double* __restrict__ x,
                                                                          const double* __restrict__ alpha,
                                                                          const double* __restrict__ kinv, int n,
                                                                          TrainHyp hyp, double* __restrict__ part) {
  __shared__ double xi[kTrainTile][D], xj[kTrainTile][D + 1], ai[kTrainTile], aj[kTrainTile];
  __shared__ double red[kTrainThreads / 32][D + 2];
  const int o = blockIdx.y;
  // tile index -> (ti >= tj): ti = floor((sqrt(8 t + 1) - 1) / 2), corrected for rounding
  const int t = blockIdx.x;
  int ti = (int)((sqrt(8.0 * t + 1.0) - 1.0) * 0.5);
  while ((int64_t)(ti + 1) * (ti + 2) / 2 <= t) ++ti;
  while ((int64_t)ti * (ti + 1) / 2 > t) --ti;
  const int tj = t - ti * (ti + 1) / 2;
  const int i0 = ti * kTrainTile, j0 = tj * kTrainTile;
  const double* al = alpha + (int64_t)o * n;
  for (int e = threadIdx.x; e < kTrainTile * D; e += kTrainThreads) {
    const int r = e / D, c = e % D;
    xi[r][c] = (i0 + r < n) ? x[(int64_t)(i0 + r) * D + c] : 0.0;
    xj[r][c] = (j0 + r < n) ? x[(int64_t)(j0 + r) * D + c] : 0.0;
  }
  if (threadIdx.x < kTrainTile) {
    ai[threadIdx.x] = (i0 + threadIdx.x < n) ? al[i0 + threadIdx.x] : 0.0;
    aj[threadIdx.x] = (j0 + threadIdx.x < n) ? al[j0 + threadIdx.x] : 0.0;
  }
  __syncthreads();
  const int tx = threadIdx.x % kTrainTile, ty = threadIdx.x / kTrainTile;
  const int j = j0 + tx;
  double acc[D + 2];
#pragma unroll
  for (int c = 0; c < D + 2; ++c) acc[c] = 0.0;
  if (j < n) {
    double mine[D], w[D];
#pragma unroll
    for (int c = 0; c < D; ++c) {
      mine[c] = xj[tx][c];
      w[c] = hyp.neg_half_inv_l2[o][c];
    }
    const double a_j = aj[tx];
    const double* ko = kinv + (int64_t)o * n * n;
#pragma unroll 4
    for (int r = ty; r < kTrainTile; r += kTrainThreads / kTrainTile) {
      const int i = i0 + r;
      if (i >= n) break;
      if (j > i) continue;  // upper part of a diagonal tile
      const double g = fma(ai[r], a_j, -ko[(int64_t)i * n + j]);
      double df2[D], a = 0.0;
#pragma unroll
      for (int c = 0; c < D; ++c) {
        const double df = xi[r][c] - mine[c];
        df2[c] = df * df;
        a = fma(df2[c], w[c], a);
      }
      const double ge = (i == j ? 1.0 : 2.0) * g * exp_nb(a);
#pragma unroll
      for (int c = 0; c < D; ++c) acc[c] = fma(ge, df2[c], acc[c]);
      acc[D] += ge;
      if (i == j) acc[D + 1] += g;
    }
  }
  // fixed-order reduction: lanes by shuffle, then the 8 warps in index order
#pragma unroll
  for (int c = 0; c < D + 2; ++c) {
    double v = acc[c];
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) v += __shfl_down_sync(0xffffffffu, v, off);
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5][c] = v;
  }
  __syncthreads();
  if (threadIdx.x < D + 2) {
    double v = 0.0;
#pragma unroll
    for (int wv = 0; wv < kTrainThreads / 32; ++wv) v += red[wv][threadIdx.x];
    part[((int64_t)o * gridDim.x + t) * (D + 2) + threadIdx.x] = v;
  }
}

// grad[o][c < D] = 1/2 s_o / l_oc^3 sum_tiles part[.][c]   (d MLL_o / d lengthscale_oc)
// grad[o][D]     = 1/2 sum_tiles part[.][D]                (d MLL_o / d outputscale_o)
// grad[o][D + 1] = 1/2 sum_tiles part[.][D + 1]            (d MLL_o / d noise_o)
// One CTA per output; every thread sums a fixed strided subset of tiles, then a fixed tree: deterministic.
__global__ void __launch_bounds__(256) gp_mll_grad_reduce_kernel(const double* __restrict__ part, int tiles, int d,
                                                                 TrainHyp hyp, double* __restrict__ grad) {
  __shared__ double red[256];
  const int o = blockIdx.x;
  for (int c = 0; c < d + 2; ++c) {
    double v = 0.0;
    for (int t = threadIdx.x; t < tiles; t += 256) v += part[((int64_t)o * tiles + t) * (d + 2) + c];
    red[threadIdx.x] = v;
    __syncthreads();
    for (int off = 128; off > 0; off >>= 1) {
      if (threadIdx.x < off) red[threadIdx.x] += red[threadIdx.x + off];
      __syncthreads();
    }
    if (threadIdx.x == 0) {
      const double scale = c < d ? 0.5 * hyp.outputscale[o] * hyp.inv_l3[o][c] : 0.5;
      grad[o * (d + 2) + c] = scale * red[0];
    }
    __syncthreads();
  }
}

}  // namespace se
