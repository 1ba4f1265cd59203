// factor_kernels.cuh -- on-device factor refresh: L^-1 of the n_s covariance matrices
// K_o + noise_o I = L_o L_o^T, built block row by block row ("bordering") so that a model that only GREW is extended
// from its old factor instead of being refactorised -- the reference's idea in
// safe_exploration/ssm_pytorch/utilities.py:122-165, and what GPyTorch's caches do behind GpCemSSM._update_model
// (ssm_cem/gp_ssm_cem.py:96-101).  Nothing here calls cuSOLVER / cuBLAS.
//
// With W = L^-1 (lower triangular) known for the leading n1 x n1 block and a new block row of FB = 128 points:
//     L21 = K21 W11^T                       (FB x n1)    GEMM, B = rows of W11: K extent ends at the diagonal block
//     S   = K22 - L21 L21^T                 (FB x FB)    split-K GEMM + fixed-order reduction
//     L22 = chol(S),  W22 = L22^-1                       one CTA per output, packed triangles in shared memory
//     T^T = W11^T L21^T                     (n1 x FB)    GEMM, A = rows of W11^T: K extent starts at the row block
//     W21^T = -T^T W22^T                    (n1 x FB)    GEMM with K = FB, written to W^T and (transposed) to W
// Every product is of the form C = alpha A B^T with both operands K-contiguous ("NT"), which is why W^T is kept beside W.
// The GEMM is FP64 DMMA (mma.sync.m8n8k4.f64), 64 x 128 tile, BK = 16, 3-stage cp.async pipeline, 2 CTAs / SM -- the main
// loop of the FP64 variance kernel (gp_kernels.cuh: gp_trisumsq_dmma_kernel) with a storing epilogue.  The same kernel
// forms (K + noise I)^-1 = W^T W = (W^T)(W^T)^T for the marginal-likelihood gradient (lower tiles only).
#pragma once
#include "gp_kernels.cuh"

namespace se {
namespace fac {

constexpr int FB = 128;          // rows per bordering step = columns of a GEMM tile
constexpr int kMaxSplits = 32;   // split-K slices of the FB x FB Schur complement

using Cfg = DmmaCfg<2, 4, 16, 3, 64, 2>;  // 64 x 128 tile, 256 threads, 2 CTAs / SM

enum KMode {
  K_FULL = 0,       // k in [0, K)  (or the blockIdx.z slice of it when splits > 1)
  K_B_LOWER = 1,    // B rows are rows of a lower-triangular matrix: column tile jt needs k < (jt + 1) 128
  K_A_UPPER = 2,    // A rows are rows of an upper-triangular matrix: row tile it needs k >= 64 it
  K_SYM_LOWER = 3   // A = B = rows of an upper-triangular matrix, only tiles that touch the lower triangle:
                    // k >= max(64 it, 128 jt)
};

struct GemmNT {
  const double* a;  // [M x K], row stride lda
  int64_t lda, a_batch;
  const double* b;  // [N x K], row stride ldb
  int64_t ldb, b_batch;
  double* c;        // [M x N], row stride ldc; split-K slice z at c + z c_split
  int64_t ldc, c_batch, c_split;
  double* ct;       // optional transposed copy [N x M] (nullptr: none)
  int64_t ldct, ct_batch;
  double alpha;
  int m_tiles, n_tiles;  // 64-row / 128-column tiles
  int k;                 // K extent, a multiple of 16
  int kmode;
  int splits;
};

// C = alpha A B^T on the tiles of one batch entry.  grid (m_tiles n_tiles, batch, splits).
static __global__ void __launch_bounds__(Cfg::THREADS, Cfg::MINB) dgemm_nt_kernel(GemmNT p) {
  using C = Cfg;
  extern __shared__ __align__(16) unsigned char smem_raw[];
  double* stages = reinterpret_cast<double*>(smem_raw);
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int wm = warp / C::WN, wn = warp % C::WN;
  const int g = lane >> 2, q = lane & 3;
  constexpr int WTM = C::BM / C::WM, WTN = C::BN / C::WN;
  const int it = blockIdx.x / p.n_tiles, jt = blockIdx.x % p.n_tiles, o = blockIdx.y, z = blockIdx.z;

  int k_lo = 0, k_hi = p.k;
  if (p.kmode == K_B_LOWER) k_hi = min(p.k, (jt + 1) * C::BN);
  if (p.kmode == K_A_UPPER) k_lo = it * C::BM;
  if (p.kmode == K_SYM_LOWER) {
    if (it * C::BM + C::BM - 1 < jt * C::BN) return;  // the tile lies above the diagonal
    k_lo = max(it * C::BM, jt * C::BN);
  }
  if (p.splits > 1) {
    const int per = ((p.k + p.splits - 1) / p.splits + C::BK - 1) / C::BK * C::BK;
    k_lo = min(p.k, z * per), k_hi = min(p.k, k_lo + per);
  }
  const int ktiles = (k_hi - k_lo) / C::BK;
  const double* a_base = p.a + o * p.a_batch + (int64_t)it * C::BM * p.lda + k_lo;
  const double* b_base = p.b + o * p.b_batch + (int64_t)jt * C::BN * p.ldb + k_lo;

  auto issue_stage = [&](int kt) {
    double* sa = stages + (kt % C::STAGES) * C::STAGE_DOUBLES;
    double* sb = sa + C::BM * C::LDS;
    const int k0 = kt * C::BK;
#pragma unroll
    for (int c = 0; c < C::CHUNKS_A; ++c) {
      const int id = tid + C::THREADS * c;
      const int row = id / (C::BK / 2), kc = (id % (C::BK / 2)) * 2;
      cp_async16(sa + row * C::LDS + kc, a_base + (int64_t)row * p.lda + k0 + kc);
    }
#pragma unroll
    for (int c = 0; c < C::CHUNKS_B; ++c) {
      const int id = tid + C::THREADS * c;
      const int row = id / (C::BK / 2), kc = (id % (C::BK / 2)) * 2;
      cp_async16(sb + row * C::LDS + kc, b_base + (int64_t)row * p.ldb + k0 + kc);
    }
  };

  double acc[C::MI][C::NI][2];
#pragma unroll
  for (int mi = 0; mi < C::MI; ++mi)
#pragma unroll
    for (int ni = 0; ni < C::NI; ++ni) acc[mi][ni][0] = acc[mi][ni][1] = 0.0;

#pragma unroll
  for (int s = 0; s < C::STAGES - 1; ++s) {
    if (s < ktiles) issue_stage(s);
    cp_async_commit();
  }
  for (int kt = 0; kt < ktiles; ++kt) {
    cp_async_wait<C::STAGES - 2>();
    __syncthreads();
    if (kt + C::STAGES - 1 < ktiles) issue_stage(kt + C::STAGES - 1);
    cp_async_commit();
    const double* sa = stages + (kt % C::STAGES) * C::STAGE_DOUBLES + (wm * WTM + g) * C::LDS + q;
    const double* sb = stages + (kt % C::STAGES) * C::STAGE_DOUBLES + C::BM * C::LDS + (wn * WTN + g) * C::LDS + q;
#pragma unroll
    for (int kk = 0; kk < C::BK; kk += 4) {
      double a[C::MI], b[C::NI];
#pragma unroll
      for (int mi = 0; mi < C::MI; ++mi) a[mi] = sa[mi * 8 * C::LDS + kk];
#pragma unroll
      for (int ni = 0; ni < C::NI; ++ni) b[ni] = sb[ni * 8 * C::LDS + kk];
#pragma unroll
      for (int mi = 0; mi < C::MI; ++mi)
#pragma unroll
        for (int ni = 0; ni < C::NI; ++ni) dmma_m8n8k4(acc[mi][ni][0], acc[mi][ni][1], a[mi], b[ni]);
    }
  }
  cp_async_wait<0>();

  // accumulator fragment (mi, ni): row g (+8 mi), columns 2q, 2q + 1 (+8 ni) of the warp tile
  double* c_base = p.c + o * p.c_batch + z * p.c_split;
  double* ct_base = p.ct ? p.ct + o * p.ct_batch : nullptr;
#pragma unroll
  for (int mi = 0; mi < C::MI; ++mi) {
    const int64_t r = (int64_t)it * C::BM + wm * WTM + mi * 8 + g;
#pragma unroll
    for (int ni = 0; ni < C::NI; ++ni) {
      const int64_t col = (int64_t)jt * C::BN + wn * WTN + ni * 8 + 2 * q;
      const double v0 = p.alpha * acc[mi][ni][0], v1 = p.alpha * acc[mi][ni][1];
      *reinterpret_cast<double2*>(c_base + r * p.ldc + col) = make_double2(v0, v1);
      if (ct_base) {
        ct_base[col * p.ldct + r] = v0;
        ct_base[(col + 1) * p.ldct + r] = v1;
      }
    }
  }
}

// packed lower triangle: element (i, k <= i)
__device__ __forceinline__ int tri(int i, int k) { return i * (i + 1) / 2 + k; }
constexpr int kTriDoubles = FB * (FB + 1) / 2;
constexpr int kCholSmemBytes = 2 * kTriDoubles * (int)sizeof(double) + 16;

// One CTA of 4 FB threads per output:  S = K22 - sum_z P_z (slices added in index order),  L22 = chol(S) (right-looking:
// the four threads 4 i .. 4 i + 3 own row i and split its trailing update by k mod 4),  W22 = L22^-1 (forward
// substitution in outer-product form, same ownership).  Writes the FB x FB block of W (lower triangle,
// zeros above) and of W^T, and the diagonal of L22.  A non-positive pivot sets info[o] = 1 + its global row index (first
// failure wins) and poisons the block with NaN, like a failed potrf.  All loads / stores of the block are spread over the
// 512 threads with independent addresses (a row-at-a-time loop is a chain of ~1 us global round trips).
constexpr int kCholThreads = 4 * FB;
static __global__ void __launch_bounds__(kCholThreads) chol_inv_block_kernel(
    const double* __restrict__ k22, int64_t ldk, int64_t k_batch, const double* __restrict__ part, int splits,
    int64_t part_split, int64_t part_batch, double* __restrict__ w, int64_t ldw, int64_t w_batch,
    double* __restrict__ wt, int64_t ldwt, int64_t wt_batch, double* __restrict__ diag, int64_t diag_batch, int row0,
    int32_t* __restrict__ info) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  double* lp = reinterpret_cast<double*>(smem_raw);  // L22, packed
  double* xp = lp + kTriDoubles;                      // W22, packed
  const int o = blockIdx.x, tid = threadIdx.x;
  const int t = tid >> 2, sub = tid & 3;  // row (Cholesky) / column (inverse) and the quarter of its sums
  k22 += o * k_batch;
  if (part) part += o * part_batch;
  // eight elements per thread at a time, so that eight loads are in flight per round trip (the slices are subtracted in
  // index order for every element: the result does not depend on the grouping)
  constexpr int kPerThread = FB * FB / kCholThreads, kGroup = 8;
#pragma unroll 1
  for (int base = 0; base < kPerThread; base += kGroup) {
    double v[kGroup];
#pragma unroll
    for (int u = 0; u < kGroup; ++u) {
      const int e = tid + kCholThreads * (base + u);
      v[u] = k22[(int64_t)(e / FB) * ldk + e % FB];
    }
    for (int zz = 0; zz < splits; ++zz) {
#pragma unroll
      for (int u = 0; u < kGroup; ++u) v[u] -= part[zz * part_split + tid + kCholThreads * (base + u)];
    }
#pragma unroll
    for (int u = 0; u < kGroup; ++u) {
      const int e = tid + kCholThreads * (base + u), r = e / FB, c = e % FB;
      if (c <= r) lp[tri(r, c)] = v[u];
    }
  }
  __syncthreads();
  // Right-looking Cholesky: column j is scaled by 1 / sqrt(pivot), then every row t > j subtracts its outer-product
  // share from the trailing triangle (independent FMAs, no dot-product chains).  The pivot stays in lp[tri(j, j)]; the
  // diagonal of L22 and its reciprocal go to dsm / rsm.  Two barriers per column.
  // sqrt and the reciprocal are software routines on the FP64 pipe: ONE thread evaluates them per column -- the owner of
  // the next pivot, right after its last update -- and publishes them through shared memory.
  __shared__ double dsm[FB], rsm[FB];
  auto publish_pivot = [&](int j, double piv) {
    if (!(piv > 0.0)) {
      atomicCAS(&info[o], 0, row0 + j + 1);
      piv = __longlong_as_double(0x7ff8000000000000ll);
    }
    const double sd = sqrt(piv);
    dsm[j] = sd, rsm[j] = 1.0 / sd;
  };
  if (tid == 0) publish_pivot(0, lp[tri(0, 0)]);
  __syncthreads();
  for (int j = 0; j < FB; ++j) {
    if (t > j && sub == 0) lp[tri(t, j)] *= rsm[j];
    __syncthreads();
    if (t > j) {
      const double ltj = lp[tri(t, j)];
      double* row = lp + tri(t, 0);
      for (int k = j + 1 + sub; k <= t; k += 4) row[k] = fma(-ltj, lp[tri(k, j)], row[k]);
      if (t == j + 1 && sub == 0) publish_pivot(j + 1, row[j + 1]);  // this thread just finished the next pivot
    }
    __syncthreads();
  }
  // W22 = L22^-1 by forward substitution in outer-product form on B = I: after step k - 1 row k of B is final and
  // X[k][:] = B[k][:] / L[k][k]; step k subtracts L[t][k] X[k][c] from every row t > k.  Rows are scaled when stored.
  for (int e = tid; e < kTriDoubles; e += kCholThreads) xp[e] = 0.0;
  __syncthreads();
  if (tid < FB) xp[tri(tid, tid)] = 1.0;
  __syncthreads();
  for (int k = 0; k < FB - 1; ++k) {
    if (t > k) {
      const double f = -lp[tri(t, k)] * rsm[k];
      const double* src = xp + tri(k, 0);
      double* row = xp + tri(t, 0);
      for (int c = sub; c <= k; c += 4) row[c] = fma(f, src[c], row[c]);
    }
    __syncthreads();
  }
  w += o * w_batch;
  wt += o * wt_batch;
  for (int e = tid; e < FB * FB; e += kCholThreads) {
    const int r = e / FB, c = e % FB;
    w[(int64_t)r * ldw + c] = c <= r ? xp[tri(r, c)] * rsm[r] : 0.0;
    wt[(int64_t)r * ldwt + c] = c >= r ? xp[tri(c, r)] * rsm[c] : 0.0;
  }
  if (tid < FB) diag[o * diag_batch + tid] = dsm[tid];
}

}  // namespace fac
}  // namespace se
