// gp_kernels.cuh -- exact-GP posterior kernels (a23-a26 in SURVEY.md 8a).
//
//   kernel rows   k_o(z_c, x_i) = s_o exp(-1/2 sum_d ((z_cd - x_id)/l_od)^2)          gp_kstar_kernel
//   mean          sum_i alpha_oi k_oi            (fused into gp_kstar_kernel)
//   d mean / d z  -sum_i alpha_oi k_oi (z_d - x_id)/l_od^2   (fused into gp_kstar_kernel)
//   variance      s_o + noise_o - || L_o^-1 k_o ||^2   -> triangular GEMM with a fused row-sum-of-squares
//                 epilogue (gp_trisumsq_*), partial sums per column block reduced in fixed order (gp_var_kernel)
//
// Reference call sites restated here: safe_exploration/ssm_cem/gp_ssm_cem.py:45-57, 59-94 and
// safe_exploration/ssm_pytorch/utilities.py:54-85 (the arithmetic itself lives in gpytorch 0.3.2).
#pragma once
#include "common.cuh"

namespace se {

constexpr int kTile = 128;  // n is padded to a multiple of this; BN of every variance-GEMM variant divides it

// ------------------------------------------------------------------------------------------------------------------
// model packing (se_gp_create)
// ------------------------------------------------------------------------------------------------------------------
template <typename T>
__global__ void pack_xt_kernel(const T* __restrict__ x, int n, int d, int n_pad, T* __restrict__ xt) {
  int64_t idx = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (idx >= (int64_t)d * n_pad) return;
  int dd = idx / n_pad, i = idx % n_pad;
  xt[idx] = i < n ? x[(int64_t)i * d + dd] : T(0);
}

template <typename T>
__global__ void pack_rows_kernel(const T* __restrict__ src, int rows, int n, int n_pad, T* __restrict__ dst) {
  int64_t idx = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (idx >= (int64_t)rows * n_pad) return;
  int r = idx / n_pad, i = idx % n_pad;
  dst[idx] = i < n ? src[(int64_t)r * n + i] : T(0);
}

// [ns][n][n] -> [ns][n_pad][n_pad], strictly-upper part and padding forced to zero
template <typename T>
__global__ void pack_w_kernel(const T* __restrict__ src, int n, int n_pad, T* __restrict__ dst) {
  int j = blockIdx.x * blockDim.x + threadIdx.x;
  int i = blockIdx.y;
  int o = blockIdx.z;
  if (j >= n_pad) return;
  T v = T(0);
  if (i < n && j <= i) v = src[((int64_t)o * n + i) * n + j];
  dst[((int64_t)o * n_pad + i) * n_pad + j] = v;
}

// ------------------------------------------------------------------------------------------------------------------
// K1: kernel rows + mean + mean-Jacobian.  One warp owns R candidate rows; lanes stride over the training set so the
// K* stores and the X / alpha loads are fully coalesced (32 consecutive elements per warp instruction).
// ------------------------------------------------------------------------------------------------------------------
template <typename T, int NS, int D, bool JAC, int R>
__global__ void __launch_bounds__(256)
gp_kstar_kernel(GpDev<T> gp, const T* __restrict__ p, int64_t p_stride, const T* __restrict__ u, int64_t u_stride,
                int64_t row0, int nrows, int rows_pad, T* __restrict__ kst, T* __restrict__ mean,
                T* __restrict__ jac) {
  const int lane = threadIdx.x & 31;
  const int64_t warp = blockIdx.x * (int64_t)(blockDim.x >> 5) + (threadIdx.x >> 5);
  const int64_t r0 = warp * R;
  if (r0 >= rows_pad) return;
  const int n_pad = gp.n_pad;

  T z[R][D];
  bool rvalid[R];
#pragma unroll
  for (int r = 0; r < R; ++r) {
    rvalid[r] = (r0 + r) < nrows;
    const int64_t c = row0 + r0 + r;
#pragma unroll
    for (int dd = 0; dd < D; ++dd) {
      T v = T(0);
      if (rvalid[r]) v = dd < NS ? p[c * p_stride + dd] : u[c * u_stride + (dd - NS)];
      z[r][dd] = v;
    }
  }

  T m_acc[R][NS];
  T j_acc[R][NS][JAC ? D : 1];
#pragma unroll
  for (int r = 0; r < R; ++r)
#pragma unroll
    for (int o = 0; o < NS; ++o) {
      m_acc[r][o] = T(0);
#pragma unroll
      for (int dd = 0; dd < (JAC ? D : 1); ++dd) j_acc[r][o][dd] = T(0);
    }

  for (int i = lane; i < n_pad; i += 32) {
    T x[D], al[NS];
#pragma unroll
    for (int dd = 0; dd < D; ++dd) x[dd] = gp.xt[(int64_t)dd * n_pad + i];
#pragma unroll
    for (int o = 0; o < NS; ++o) al[o] = gp.alpha[(int64_t)o * n_pad + i];
    const bool ivalid = i < gp.n;
#pragma unroll
    for (int r = 0; r < R; ++r) {
      T diff[D];
#pragma unroll
      for (int dd = 0; dd < D; ++dd) diff[dd] = z[r][dd] - x[dd];
#pragma unroll
      for (int o = 0; o < NS; ++o) {
        T e = T(0);
#pragma unroll
        for (int dd = 0; dd < D; ++dd) e += diff[dd] * diff[dd] * gp.inv_l2[o][dd];
        T k = (ivalid && rvalid[r]) ? gp.outputscale[o] * Num<T>::exp_(T(-0.5) * e) : T(0);
        kst[((int64_t)o * rows_pad + r0 + r) * n_pad + i] = k;
        const T wk = al[o] * k;
        m_acc[r][o] += wk;
        if (JAC) {
#pragma unroll
          for (int dd = 0; dd < D; ++dd) j_acc[r][o][dd] -= wk * diff[dd] * gp.inv_l2[o][dd];
        }
      }
    }
  }

#pragma unroll
  for (int r = 0; r < R; ++r)
#pragma unroll
    for (int o = 0; o < NS; ++o) {
      T v = m_acc[r][o];
#pragma unroll
      for (int s = 16; s > 0; s >>= 1) v += __shfl_xor_sync(0xffffffffu, v, s);
      if (lane == 0 && rvalid[r]) mean[(row0 + r0 + r) * NS + o] = v;
      if (JAC) {
#pragma unroll
        for (int dd = 0; dd < D; ++dd) {
          T g = j_acc[r][o][dd];
#pragma unroll
          for (int s = 16; s > 0; s >>= 1) g += __shfl_xor_sync(0xffffffffu, g, s);
          if (lane == 0 && rvalid[r]) jac[((row0 + r0 + r) * NS + o) * D + dd] = g;
        }
      }
    }
}

// ------------------------------------------------------------------------------------------------------------------
// K2: partial[o][ib][row] = sum_{i in column block ib} ( sum_{j <= i} W_o[i][j] K*_o[row][j] )^2
//
// Work item = (output o, row tile rt, column block ib); its cost is proportional to ib + 1 (triangular depth), so
// items are handed out heaviest-first through an atomic counter (persistent CTAs, grid = multiple of the SM count).
// ------------------------------------------------------------------------------------------------------------------
struct TriItem {
  int o, rt, ib;
};
__device__ __forceinline__ TriItem decode_item(int idx, int nb, int rt_count, int ns) {
  TriItem it;
  const int per_ib = rt_count * ns;
  it.ib = nb - 1 - idx / per_ib;
  const int rem = idx % per_ib;
  it.rt = rem / ns;
  it.o = rem % ns;
  return it;
}

template <typename T>
__device__ __forceinline__ void load4(const T* __restrict__ src, T v[4]);
template <>
__device__ __forceinline__ void load4<float>(const float* __restrict__ src, float v[4]) {
  const float4 t = *reinterpret_cast<const float4*>(src);
  v[0] = t.x, v[1] = t.y, v[2] = t.z, v[3] = t.w;
}
template <>
__device__ __forceinline__ void load4<double>(const double* __restrict__ src, double v[4]) {
  const double2 t0 = *reinterpret_cast<const double2*>(src);
  const double2 t1 = *reinterpret_cast<const double2*>(src + 2);
  v[0] = t0.x, v[1] = t0.y, v[2] = t1.x, v[3] = t1.y;
}

// --- SIMT variant: 64x64 tile, BK = 16, 256 threads, 4x4 micro-tile (FFMA / DFMA).  Validation path and fallback. ---
template <typename T>
__global__ void __launch_bounds__(256)
gp_trisumsq_simt_kernel(const T* __restrict__ kst, const T* __restrict__ w, int n_pad, int rows_pad, int ns,
                        T* __restrict__ partial, int* __restrict__ counter) {
  constexpr int BM = 64, BN = 64, BK = 16, PADM = 4;
  __shared__ T As[BK][BM + PADM];
  __shared__ T Bs[BK][BN + PADM];
  __shared__ int s_item;

  const int tid = threadIdx.x;
  const int tx = tid & 15, ty = tid >> 4;
  const int lrow = tid >> 2, lk = (tid & 3) * 4;
  const int nb = n_pad / BN;
  const int rt_count = rows_pad / BM;
  const int total = nb * rt_count * ns;

  while (true) {
    if (tid == 0) s_item = atomicAdd(counter, 1);
    __syncthreads();
    const int idx = s_item;
    __syncthreads();
    if (idx >= total) break;
    const TriItem it = decode_item(idx, nb, rt_count, ns);
    const T* a_base = kst + ((int64_t)it.o * rows_pad + (int64_t)it.rt * BM) * n_pad;
    const T* b_base = w + ((int64_t)it.o * n_pad + (int64_t)it.ib * BN) * n_pad;
    const int kend = (it.ib + 1) * BN;

    T acc[4][4];
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
      for (int j = 0; j < 4; ++j) acc[i][j] = T(0);

    T ra[4], rb[4];
    load4<T>(a_base + (int64_t)lrow * n_pad + lk, ra);
    load4<T>(b_base + (int64_t)lrow * n_pad + lk, rb);
    for (int k0 = 0; k0 < kend; k0 += BK) {
#pragma unroll
      for (int e = 0; e < 4; ++e) {
        As[lk + e][lrow] = ra[e];
        Bs[lk + e][lrow] = rb[e];
      }
      __syncthreads();
      if (k0 + BK < kend) {
        load4<T>(a_base + (int64_t)lrow * n_pad + k0 + BK + lk, ra);
        load4<T>(b_base + (int64_t)lrow * n_pad + k0 + BK + lk, rb);
      }
#pragma unroll
      for (int kk = 0; kk < BK; ++kk) {
        T av[4], bv[4];
#pragma unroll
        for (int i = 0; i < 4; ++i) av[i] = As[kk][ty * 4 + i];
#pragma unroll
        for (int j = 0; j < 4; ++j) bv[j] = Bs[kk][tx * 4 + j];
#pragma unroll
        for (int i = 0; i < 4; ++i)
#pragma unroll
          for (int j = 0; j < 4; ++j) acc[i][j] += av[i] * bv[j];
      }
      __syncthreads();
    }

    // epilogue: row sums of squares over this column block; the 16 lanes of a half-warp share ty
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      T s = T(0);
#pragma unroll
      for (int j = 0; j < 4; ++j) s += acc[i][j] * acc[i][j];
#pragma unroll
      for (int sh = 8; sh > 0; sh >>= 1) s += __shfl_xor_sync(0xffffffffu, s, sh);
      if (tx == 0)
        partial[((int64_t)it.o * nb + it.ib) * rows_pad + (int64_t)it.rt * BM + ty * 4 + i] = s;
    }
  }
}

// --- FP64 tensor-core variant: DMMA m8n8k4, 128x128 tile, BK = 16, cp.async 3-stage pipeline --------------------
__device__ __forceinline__ void dmma_m8n8k4(double& c0, double& c1, double a, double b) {
  asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};\n"
               : "+d"(c0), "+d"(c1)
               : "d"(a), "d"(b));
}
__device__ __forceinline__ void cp_async16(void* smem_dst, const void* gmem_src) {
  const uint32_t s = static_cast<uint32_t>(__cvta_generic_to_shared(smem_dst));
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;\n" ::"r"(s), "l"(gmem_src));
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;\n" ::); }
template <int N>
__device__ __forceinline__ void cp_async_wait() {
  asm volatile("cp.async.wait_group %0;\n" ::"n"(N));
}

// Work-item order.  rc = row tiles per "L2 chunk": all items of rc consecutive row tiles are issued before the next
// chunk starts, so the K* rows of a chunk (rc * 128 * n_pad * ns elements) stay L2-resident while every column block
// re-reads them; inside a chunk items go heaviest (deepest column block) first.
struct TriOrder {
  int nb, rt_count, ns, rc;
};
__device__ __forceinline__ TriItem decode_item_chunked(int idx, const TriOrder& od) {
  const int per_chunk = od.rc * od.ns * od.nb;
  const int chunk = idx / per_chunk;
  const int rt0 = chunk * od.rc;
  const int rc_here = min(od.rc, od.rt_count - rt0);
  // the last chunk may hold fewer row tiles; its items follow all full chunks
  const int local = idx - chunk * per_chunk;
  const int per_ib = rc_here * od.ns;
  TriItem it;
  it.ib = od.nb - 1 - local / per_ib;
  const int rem = local % per_ib;
  it.rt = rt0 + rem / od.ns;
  it.o = rem % od.ns;
  return it;
}
__host__ __device__ inline int tri_total_items(int nb, int rt_count, int ns) { return nb * rt_count * ns; }

template <int WM_, int WN_, int BK_, int STAGES_, int BM_ = 128, int MINB_ = 1>
struct DmmaCfg {
  static constexpr int BM = BM_, BN = 128, BK = BK_, STAGES = STAGES_, WM = WM_, WN = WN_, MINB = MINB_;
  static constexpr int LDS = BK + 4;  // doubles per smem row: (BK + 4) mod 16 == 4 -> conflict-free LDS.64 fragments
  static constexpr int THREADS = WM * WN * 32;
  static constexpr int MI = BM / WM / 8, NI = BN / WN / 8;  // m8n8 tiles per warp
  static constexpr int STAGE_DOUBLES = (BM + BN) * LDS;
  static constexpr int SMEM_BYTES = STAGES * STAGE_DOUBLES * 8 + WN * BM * 8 + 16;
  static constexpr int CHUNKS_A = BM * (BK / 2) / THREADS;  // 16-byte cp.async chunks per thread, K* tile
  static constexpr int CHUNKS_B = BN * (BK / 2) / THREADS;  // ... and L^-1 tile
  static_assert(BM * (BK / 2) % THREADS == 0 && BN * (BK / 2) % THREADS == 0, "tile / thread mismatch");
};

template <typename C, bool SKIP_DIAG>
static __global__ void __launch_bounds__(C::THREADS, C::MINB)
gp_trisumsq_dmma_kernel(const double* __restrict__ kst, const double* __restrict__ w, int n_pad, int rows_pad,
                        TriOrder od, double* __restrict__ partial, int* __restrict__ counter) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  double* stages = reinterpret_cast<double*>(smem_raw);
  double* red = stages + C::STAGES * C::STAGE_DOUBLES;  // [WN][BM]
  int* s_item = reinterpret_cast<int*>(red + C::WN * C::BM);

  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int wm = warp / C::WN;
  // warp w runs on SM sub-partition w % 4 (each has its own DMMA unit): odd warp rows take the column groups in
  // reverse so that the triangular skip below removes the same amount of work from every sub-partition
  const int wn = (SKIP_DIAG && (wm & 1)) ? C::WN - 1 - warp % C::WN : warp % C::WN;
  const int g = lane >> 2, q = lane & 3;
  constexpr int WTM = C::BM / C::WM, WTN = C::BN / C::WN;  // warp tile
  const int total = tri_total_items(od.nb, od.rt_count, od.ns);

  while (true) {
    if (tid == 0) *s_item = atomicAdd(counter, 1);
    __syncthreads();
    const int idx = *s_item;
    __syncthreads();
    if (idx >= total) break;
    const TriItem it = decode_item_chunked(idx, od);
    const double* a_base = kst + ((int64_t)it.o * rows_pad + (int64_t)it.rt * C::BM) * n_pad;
    const double* b_base = w + ((int64_t)it.o * n_pad + (int64_t)it.ib * C::BN) * n_pad;
    const int ktiles = (it.ib + 1) * C::BN / C::BK;
    const int diag_kt = it.ib * C::BN / C::BK;  // first k-tile of the diagonal block (negative offsets never skip)

    auto issue_stage = [&](int kt) {
      double* sa = stages + (kt % C::STAGES) * C::STAGE_DOUBLES;
      double* sb = sa + C::BM * C::LDS;
      const int k0 = kt * C::BK;
#pragma unroll
      for (int c = 0; c < C::CHUNKS_A; ++c) {
        const int id = tid + C::THREADS * c;
        const int row = id / (C::BK / 2), kc = (id % (C::BK / 2)) * 2;
        cp_async16(sa + row * C::LDS + kc, a_base + (int64_t)row * n_pad + k0 + kc);
      }
#pragma unroll
      for (int c = 0; c < C::CHUNKS_B; ++c) {
        const int id = tid + C::THREADS * c;
        const int row = id / (C::BK / 2), kc = (id % (C::BK / 2)) * 2;
        cp_async16(sb + row * C::LDS + kc, b_base + (int64_t)row * n_pad + k0 + kc);
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
      const double* sb =
          stages + (kt % C::STAGES) * C::STAGE_DOUBLES + C::BM * C::LDS + (wn * WTN + g) * C::LDS + q;
      // Diagonal block: W[i][j] = 0 for j > i, so a warp whose 32 columns all have i < k0 has nothing to add.  The
      // predicate is warp-uniform and tile-granular (BK divides the warp-tile width); the warp -> column mapping above
      // pairs column groups s and WN-1-s on the same SM sub-partition, so the skipped DMMAs shorten the critical path.
      if (!SKIP_DIAG || (kt - diag_kt) * C::BK < (wn + 1) * WTN) {
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
    }
    cp_async_wait<0>();

    // epilogue: accumulator fragment (mi, ni) holds row g (+8 mi) and columns 2q, 2q+1 (+8 ni)
#pragma unroll
    for (int mi = 0; mi < C::MI; ++mi) {
      double s = 0.0;
#pragma unroll
      for (int ni = 0; ni < C::NI; ++ni) s += acc[mi][ni][0] * acc[mi][ni][0] + acc[mi][ni][1] * acc[mi][ni][1];
      s += __shfl_xor_sync(0xffffffffu, s, 1);
      s += __shfl_xor_sync(0xffffffffu, s, 2);
      if (q == 0) red[wn * C::BM + wm * WTM + mi * 8 + g] = s;
    }
    __syncthreads();
    if (tid < C::BM) {
      double s = 0.0;
#pragma unroll
      for (int x = 0; x < C::WN; ++x) s += red[x * C::BM + tid];
      partial[((int64_t)it.o * od.nb + it.ib) * rows_pad + (int64_t)it.rt * C::BM + tid] = s;
    }
    // the __syncthreads at the top of the loop orders these reads of `red` before the next item's writes
  }
}

// var[c][o] = s_o + noise_o - sum_ib partial[o][ib][row]  (fixed summation order => run-to-run reproducible).
// When the kernel-row kernel split the training set over CTAs (small batches), kpart [nseg][rows_pad][ns][1 + d] holds
// its per-segment mean / Jacobian sums; they are added here in segment order, exactly as the unsplit kernel adds them.
template <typename T>
__global__ void gp_var_kernel(GpDev<T> gp, const T* __restrict__ partial, int nb, int rows_pad, int64_t row0,
                              int nrows, T* __restrict__ var, const T* __restrict__ kpart, int nseg,
                              T* __restrict__ mean, T* __restrict__ jac) {
  const int64_t idx = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (idx >= (int64_t)nrows * gp.ns) return;
  const int o = idx % gp.ns;
  const int64_t r = idx / gp.ns;
  T s = T(0);
  for (int ib = 0; ib < nb; ++ib) s += partial[((int64_t)o * nb + ib) * rows_pad + r];
  T v = gp.outputscale[o] - s + gp.noise[o];
  if (nseg > 0) {
    const int w = 1 + gp.d;
    const T* src = kpart + (r * gp.ns + o) * w;
    const int64_t seg_stride = (int64_t)rows_pad * gp.ns * w;
    T m = T(0);
    for (int sg = 0; sg < nseg; ++sg) m += src[sg * seg_stride];
    mean[(row0 + r) * gp.ns + o] = m;
    if (jac)
      for (int dd = 0; dd < gp.d; ++dd) {
        T g = T(0);
        for (int sg = 0; sg < nseg; ++sg) g += src[sg * seg_stride + 1 + dd];
        jac[((row0 + r) * gp.ns + o) * gp.d + dd] = -g * gp.inv_l2[o][dd];
      }
  }
  // A non-finite input poisons the row's mean (the digit planes of the tensor-core path cannot carry a NaN); the
  // reference's GP returns NaN for mean AND variance there, which is what _fix_zeros_nans looks at
  // (gp_reachability_pytorch.py:234-243).
  const T m_final = mean[(row0 + r) * gp.ns + o];
  if (m_final != m_final) v = m_final;
  var[(row0 + r) * gp.ns + o] = v;
}

}  // namespace se
