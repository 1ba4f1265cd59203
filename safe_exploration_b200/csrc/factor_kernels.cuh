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
  K_FULL = 0,       // k in [0, K)
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
  int splits;            // > 1: blockIdx.z takes an equal slice of every tile's K range and writes it to c + z c_split;
                         // the slices are added in index order by sum_slices_kernel / chol_inv_block_kernel
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
  if (p.splits > 1) {  // slice z of THIS tile's K range (slices past its end are empty and store zeros)
    const int per = ((k_hi - k_lo + p.splits - 1) / p.splits + C::BK - 1) / C::BK * C::BK;
    const int lo = min(k_hi, k_lo + z * per);
    k_hi = min(k_hi, lo + per), k_lo = lo;
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

// dst[o][r][c] = sum_z part[o][z][r][c] for a [rows x cols] block, slices added in index order (deterministic).
static __global__ void sum_slices_kernel(const double* __restrict__ part, int splits, int64_t slice_stride,
                                         int64_t part_batch, int64_t ld_part, int rows, int cols,
                                         double* __restrict__ dst, int64_t ld_dst, int64_t dst_batch) {
  const int64_t e = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  const int o = blockIdx.y;
  if (e >= (int64_t)rows * (cols / 2)) return;
  const int64_t r = e / (cols / 2), c = (e % (cols / 2)) * 2;
  const double* src = part + o * part_batch + r * ld_part + c;
  double2 acc = *reinterpret_cast<const double2*>(src);
  for (int z = 1; z < splits; ++z) {
    const double2 v = *reinterpret_cast<const double2*>(src + z * slice_stride);
    acc.x += v.x, acc.y += v.y;
  }
  *reinterpret_cast<double2*>(dst + o * dst_batch + r * ld_dst + c) = acc;
}

// packed lower triangle: element (i, k <= i)
__device__ __forceinline__ int tri(int i, int k) { return i * (i + 1) / 2 + k; }
constexpr int kTriDoubles = FB * (FB + 1) / 2;
constexpr int kCholSmemBytes = 2 * kTriDoubles * (int)sizeof(double) + 16;

// One CTA of 4 FB threads per output:  S = K22 - sum_z P_z (slices added in index order),  L22 = chol(S),  W22 = L22^-1.
// Writes the FB x FB block of W (lower triangle, zeros above) and of W^T, and the diagonal of L22.  A non-positive pivot
// sets info[o] = 1 + its global row index (first failure wins) and poisons the block with NaN, like a failed potrf.
//
// Measured with clock64 on the first version of this kernel (two barriers per column, one thread evaluating sqrt and the
// reciprocal, a barrier per substitution step): 247 k cycles for the Cholesky and 185 k for the inverse of ONE block --
// block-wide barriers and the single-thread pivot were the time, not the arithmetic.  Hence:
//  * Cholesky, right-looking on UNSCALED columns (an L D L^T sweep): with d_j the pivot, S[t][k] -= S[t][j] S[k][j] / d_j.
//    Every thread takes 1 / d_j itself after ONE barrier per column; L = S diag(1 / sqrt d) is formed at the end.
//    The four threads 4 t .. 4 t + 3 own row t and split its updates by k mod 4, four elements at a time with all loads
//    before the stores (row and column entries live in the same array: the compiler orders every load behind the
//    previous store otherwise).
//  * Inverse, column-parallel forward substitution in outer-product form: the four threads of column c (one warp holds
//    eight columns) carry x = e_c through k = c .. FB - 1: x_k is final, x_i -= L[i][k] x_k for i > k.  No block barrier.
constexpr int kCholThreads = 4 * FB;
static __global__ void __launch_bounds__(kCholThreads) chol_inv_block_kernel(
    const double* __restrict__ k22, int64_t ldk, int64_t k_batch, const double* __restrict__ part, int splits,
    int64_t part_split, int64_t part_batch, double* __restrict__ w, int64_t ldw, int64_t w_batch,
    double* __restrict__ wt, int64_t ldwt, int64_t wt_batch, double* __restrict__ diag, int64_t diag_batch, int row0,
    int32_t* __restrict__ info) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  double* lp = reinterpret_cast<double*>(smem_raw);  // S, then L22 (packed lower triangle)
  double* xp = lp + kTriDoubles;                      // W22 (packed)
  __shared__ double dsm[FB], rsm[FB];                 // diag(L22) and its reciprocal
  const int o = blockIdx.x, tid = threadIdx.x;
  const int t = tid >> 2, sub = tid & 3;  // row (Cholesky) / column (inverse) and the residue class of its updates
  k22 += o * k_batch;
  if (part) part += o * part_batch;
  // eight elements per thread and four slices at a time: 8 / 32 loads in flight per round trip (the slices are
  // subtracted in index order for every element, so the result does not depend on the grouping)
  constexpr int kPerThread = FB * FB / kCholThreads, kGroup = 8;
#pragma unroll 1
  for (int base = 0; base < kPerThread; base += kGroup) {
    double v[kGroup];
#pragma unroll
    for (int u = 0; u < kGroup; ++u) {
      const int e = tid + kCholThreads * (base + u);
      v[u] = k22[(int64_t)(e / FB) * ldk + e % FB];
    }
    int zz = 0;
    for (; zz + 4 <= splits; zz += 4) {
      double pz[4][kGroup];
#pragma unroll
      for (int q = 0; q < 4; ++q)
#pragma unroll
        for (int u = 0; u < kGroup; ++u) pz[q][u] = part[(zz + q) * part_split + tid + kCholThreads * (base + u)];
#pragma unroll
      for (int q = 0; q < 4; ++q)
#pragma unroll
        for (int u = 0; u < kGroup; ++u) v[u] -= pz[q][u];
    }
    for (; zz < splits; ++zz) {
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
  for (int j = 0; j < FB - 1; ++j) {
    const double piv = lp[tri(j, j)];
    if (t > j) {
      const double f = -lp[tri(t, j)] * __drcp_rn(piv);  // the reciprocal intrinsic: a third of the instructions of a division
      double* row = lp + tri(t, 0);
      int k = j + 1 + sub;
      for (; k + 12 <= t; k += 16) {
        const double c0 = lp[tri(k, j)], c1 = lp[tri(k + 4, j)], c2 = lp[tri(k + 8, j)], c3 = lp[tri(k + 12, j)];
        const double r0 = row[k], r1 = row[k + 4], r2 = row[k + 8], r3 = row[k + 12];
        row[k] = fma(f, c0, r0), row[k + 4] = fma(f, c1, r1);
        row[k + 8] = fma(f, c2, r2), row[k + 12] = fma(f, c3, r3);
      }
      for (; k <= t; k += 4) row[k] = fma(f, lp[tri(k, j)], row[k]);
    }
    __syncthreads();
  }
  if (tid < FB) {
    double piv = lp[tri(tid, tid)];
    if (!(piv > 0.0)) piv = __longlong_as_double(0x7ff8000000000000ll);
    const double r = rsqrt(piv);
    double sd = piv * r;
    sd = fma(fma(-sd, sd, piv), 0.5 * r, sd);
    dsm[tid] = sd, rsm[tid] = r;
  }
  __syncthreads();
  if (tid == 0) {  // first bad pivot in row order (a NaN pivot poisons everything after it)
    for (int j = 0; j < FB; ++j)
      if (!(dsm[j] == dsm[j])) {
        atomicCAS(&info[o], 0, row0 + j + 1);
        break;
      }
  }
  for (int e = tid; e < kTriDoubles; e += kCholThreads) xp[e] = 0.0;
  // L22 = S diag(1 / sqrt d): column j of the packed triangle times rsm[j]
  for (int e = tid; e < FB * FB; e += kCholThreads) {
    const int r = e / FB, c = e % FB;
    if (c < r) lp[tri(r, c)] *= rsm[c];
  }
  __syncthreads();
  // W22 = L22^-1, column t (warp-synchronous: the four threads of a column sit in one warp, which holds 8 columns and
  // starts at its first column so that __syncwarp is reached by all lanes)
  {
    const int c = t;
    double* xc = xp;  // element (i, c) at tri(i, c)
    for (int k = t & ~7; k < FB; ++k) {
      double xk = 0.0;
      if (k >= c) {
        const double b = k == c ? 1.0 : xc[tri(k, c)];
        xk = b * rsm[k];
      }
      __syncwarp();  // every lane has read x_k before lane `sub == 0` overwrites it with its final value
      if (k >= c) {
        if (sub == 0) xc[tri(k, c)] = xk;
        int i = k + 1 + sub;
        for (; i + 12 < FB; i += 16) {
          const double l0 = lp[tri(i, k)], l1 = lp[tri(i + 4, k)], l2 = lp[tri(i + 8, k)], l3 = lp[tri(i + 12, k)];
          const double x0 = xc[tri(i, c)], x1 = xc[tri(i + 4, c)], x2 = xc[tri(i + 8, c)], x3 = xc[tri(i + 12, c)];
          xc[tri(i, c)] = fma(-l0, xk, x0), xc[tri(i + 4, c)] = fma(-l1, xk, x1);
          xc[tri(i + 8, c)] = fma(-l2, xk, x2), xc[tri(i + 12, c)] = fma(-l3, xk, x3);
        }
        for (; i < FB; i += 4) xc[tri(i, c)] = fma(-lp[tri(i, k)], xk, xc[tri(i, c)]);
      }
      __syncwarp();  // the updates of step k are visible to the lane that reads x_{k+1}
    }
  }
  __syncthreads();
  w += o * w_batch;
  wt += o * wt_batch;
  for (int e = tid; e < FB * FB; e += kCholThreads) {
    const int r = e / FB, c = e % FB;
    w[(int64_t)r * ldw + c] = c <= r ? xp[tri(r, c)] : 0.0;
    wt[(int64_t)r * ldwt + c] = c >= r ? xp[tri(c, r)] : 0.0;
  }
  if (tid < FB) diag[o * diag_batch + tid] = dsm[tid];
}

}  // namespace fac
}  // namespace se
