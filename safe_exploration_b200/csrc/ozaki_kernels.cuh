// ozaki_kernels.cuh -- FP64-accurate variance GEMM on the 5th-generation tensor cores (tcgen05, kind::i8).
//
// The variance of the exact GP (gp_ssm_cem.py:93 via gpytorch) needs  q_c = || L^-1 k*_c ||^2  to ~1e-10 of the prior
// scale (the result is a difference of two numbers 10^3 apart), which rules out every fp32-accumulating tensor-core
// format.  tcgen05 has no f64 kind, but its s8 x s8 -> s32 path is EXACT, so the product is evaluated error-free from
// balanced radix-256 digits (Ozaki scheme I with full signed bytes; |digit| <= 128, top digit <= 127):
//
//     k*_cj   = 2^eA   sum_t dA_t[c][j] 2^(-7-8t)        (S digit planes, written by gp_kstar_slices)
//     W_ij    = 2^eB_i sum_u dB_u[i][j] 2^(-7-8u)        (S digit planes, packed once per model)
//     v_ci    = sum_j W_ij k*_cj = 2^(eA+eB_i-14) sum_l 2^(-8l) D_l[c][i],   D_l = sum_{t+u=l} dA_t dB_u^T   (int32, exact)
//
// Only the digit pairs with t + u <= S - 1 are formed (S (S + 1) / 2 int8 GEMMs); the dropped pairs are below
// 2^(-8S) of the row scales.  |D_l| <= (l + 1) 2^14 n stays inside int32 for n <= 16384 (measured: 2^21 at n = 2000).  All S levels D_l of a 128 x 64 output tile live in TMEM at once (S x 64 columns), so every
// digit tile that the bulk-copy engine brings into shared memory is used by up to S MMAs.  The epilogue recombines the
// levels in int64 (exact), converts once to double, squares and row-sums: partial[o][ib][c] like the DMMA kernel.
//
// Shared-memory operand images are the no-swizzle K-major canonical UMMA layout (8 rows x 16 bytes core matrices,
// LBO = 128 B between the two K halves of a K = 32 instruction, SBO = 256 B between 8-row groups); the global-memory
// images are stored in exactly that order, so one contiguous cp.async.bulk per operand fills a pipeline stage.
#pragma once
#include "common.cuh"
#include "gp_kernels.cuh"
#include "reach_kernels.cuh"

namespace se {
namespace oz {

constexpr int TM = 128;          // candidates per tile  (UMMA M)
constexpr int TN = 64;           // L^-1 rows per tile   (UMMA N)
constexpr int KB = 32;           // K elements per kind::i8 instruction
constexpr int A_BLK = TM * KB;   // bytes of one digit plane of one K step, candidates
constexpr int B_BLK = TN * KB;   // ... L^-1 rows
constexpr int MAX_S = 8;         // 8 levels x 64 columns = all 512 TMEM columns
constexpr int THREADS = 320;     // warp 0: bulk-copy producer, warp 1: MMA issuer, warps 2-9: epilogue
constexpr int EPI_THREADS = 256;
constexpr int TMEM_COLS = 512;
constexpr uint32_t LBO = 128, SBO = 256;

// byte offset of element (row r, k index c < 32) inside a digit-plane block of a K step
__host__ __device__ __forceinline__ int blk_off(int r, int c) {
  return (r >> 3) * 256 + (c >> 4) * 128 + (r & 7) * 16 + (c & 15);
}

template <int S>
struct Cfg {
  static constexpr int KS = 2;  // K steps per pipeline stage (one barrier round trip per KS * S (S + 1) / 2 MMAs)
  static constexpr int STEP_BYTES = S * (A_BLK + B_BLK);
  static constexpr int STAGE_BYTES = KS * STEP_BYTES;
  static constexpr int STAGES = (220 * 1024) / STAGE_BYTES;
  static constexpr int SMEM_BYTES = STAGES * STAGE_BYTES + 1024 /*alignment slack*/ + 256 /*barriers*/;
  static constexpr int PAIRS = S * (S + 1) / 2;
};

__device__ __forceinline__ uint64_t make_desc(uint32_t smem_addr, uint32_t lbo, uint32_t sbo);

// ---- PTX wrappers ---------------------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return static_cast<uint32_t>(__cvta_generic_to_shared(p)); }

__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;\n" ::"r"(bar), "r"(count));
}
__device__ __forceinline__ void mbar_fence_init() { asm volatile("fence.mbarrier_init.release.cluster;\n" ::: "memory"); }
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;\n" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];\n" ::"r"(bar) : "memory");
}
// Bounded wait: a protocol bug must end in a trap (launch failure), never in a hung GPU.
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
  uint32_t done = 0;
  const long long t0 = clock64();
  while (true) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}\n"
        : "=r"(done)
        : "r"(bar), "r"(parity)
        : "memory");
    if (done) break;
    if (clock64() - t0 > 4000000000ll) __trap();
  }
}
__device__ __forceinline__ void bulk_g2s(uint32_t dst, const void* src, uint32_t bytes, uint32_t bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];\n" ::"r"(dst),
               "l"(src), "r"(bytes), "r"(bar)
               : "memory");
}
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;\n" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;\n" ::: "memory"); }
__device__ __forceinline__ void tmem_alloc(uint32_t dst_smem, uint32_t cols) {  // whole warp
  asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;\n" ::"r"(dst_smem), "r"(cols) : "memory");
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;\n" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc(uint32_t taddr, uint32_t cols) {  // whole warp
  asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;\n" ::"r"(taddr), "r"(cols) : "memory");
}
__device__ __forceinline__ void tc_commit(uint32_t bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];\n" ::"r"(bar) : "memory");
}
// D[tmem] (+)= A[smem] * B[smem]^T, s8 x s8 -> s32, M = 128, N = 64, K = 32; issued by ONE thread.
// COLL: A-operand collector usage.  Consecutive MMAs that multiply the same A digit tile keep it in the tensor core's
// collector buffer (0 discard, 1 fill, 2 use, 3 lastuse), so shared memory is read for A once per digit, not per pair.
template <int COLL>
__device__ __forceinline__ void mma_i8(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc, uint32_t idesc, uint32_t accumulate) {
#define SE_MMA_I8(QUAL)                                                                                        \
  asm volatile(                                                                                                \
      "{\n\t.reg .pred p;\n\t"                                                                                 \
      "setp.ne.b32 p, %4, 0;\n\t"                                                                              \
      "tcgen05.mma.cta_group::1.kind::i8" QUAL " [%0], %1, %2, %3, p;\n\t}\n" ::"r"(d_tmem),                   \
      "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(accumulate)                                                    \
      : "memory")
  if constexpr (COLL == 1) SE_MMA_I8(".collector::a::fill");
  else if constexpr (COLL == 2) SE_MMA_I8(".collector::a::use");
  else if constexpr (COLL == 3) SE_MMA_I8(".collector::a::lastuse");
  else SE_MMA_I8("");
#undef SE_MMA_I8
}

// one lane of a converged warp (the async-issue idiom: the warp stays converged, so every operand stays uniform)
__device__ __forceinline__ bool elect_one() {
  uint32_t pred;
  asm volatile("{\n\t.reg .pred P;\n\telect.sync _|P, 0xffffffff;\n\tselp.u32 %0, 1, 0, P;\n\t}\n" : "=r"(pred));
  return pred != 0;
}
__device__ __forceinline__ uint64_t pack_desc(uint32_t lo, uint32_t hi) {
  uint64_t d;
  asm("mov.b64 %0, {%1, %2};\n" : "=l"(d) : "r"(lo), "r"(hi));
  return d;
}

// instruction descriptor: c = s32 (2 << 4), a = b = signed 8 bit (1 << 7, 1 << 10), both K-major, N >> 3, M >> 4
__host__ __device__ constexpr uint32_t make_idesc(int m, int n) {
  return (2u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(n >> 3) << 17) | ((uint32_t)(m >> 4) << 24);
}

// All digit pairs (t, u), t + u <= S - 1, of one K step: level t + u accumulates A_t B_u^T.
// a_lo / b_lo: low descriptor words of digit plane 0 (start address >> 4 | LBO); the other planes are a constant
// number of 16-byte units further on, so a descriptor costs one add.
// MODE 0 / 1: one N = 64 MMA per pair (1: the A digit tile stays in the collector across its pairs).  The issue rate
//   matters there: an N = 64 MMA occupies the tensor core for only 32 cycles.
// MODE 2: the digit planes of L^-1 lie back to back in the stage (64 rows = 8 row groups each, SBO apart) and the level
//   accumulators lie back to back in TMEM, so A_t times planes 0 .. S-1-t is ONE MMA of N = 64 (S - t) columns that
//   lands on levels t .. S-1 -- split only where N would exceed 256.  S = 6: 9 MMAs per K step instead of 21.
template <int S, int MODE>
__device__ __forceinline__ void issue_kstep(uint32_t tmem_base, uint32_t a_lo, uint32_t b_lo, uint32_t desc_hi,
                                            uint32_t idesc, uint32_t not_first) {
  if constexpr (MODE == 2) {
#pragma unroll
    for (int t = 0; t < S; ++t) {
      const uint64_t ad = pack_desc(a_lo + t * (A_BLK >> 4), desc_hi);
      const int planes = S - t, chunks = (planes + 3) / 4;
#pragma unroll
      for (int c = 0; c < chunks; ++c) {
        const int cnt = planes - 4 * c < 4 ? planes - 4 * c : 4;
        const uint64_t bd = pack_desc(b_lo + 4 * c * (B_BLK >> 4), desc_hi);
        const uint32_t acc = t > 0 ? 1u : not_first;  // level l is first written by digit 0 of K step 0
        const uint32_t dcol = tmem_base + (uint32_t)((t + 4 * c) * TN);
        const uint32_t id = make_idesc(TM, cnt * TN);
        if (chunks == 1) mma_i8<0>(dcol, ad, bd, id, acc);
        else if (c == 0) mma_i8<1>(dcol, ad, bd, id, acc);
        else if (c == chunks - 1) mma_i8<3>(dcol, ad, bd, id, acc);
        else mma_i8<2>(dcol, ad, bd, id, acc);
      }
    }
  } else {
#pragma unroll
    for (int t = 0; t < S; ++t) {
      const uint64_t ad = pack_desc(a_lo + t * (A_BLK >> 4), desc_hi);
#pragma unroll
      for (int uu = 0; uu + t < S; ++uu) {
        const uint64_t bd = pack_desc(b_lo + uu * (B_BLK >> 4), desc_hi);
        // level t + uu is first written by the pair (0, t + uu) of K step 0
        const uint32_t acc = t > 0 ? 1u : not_first;
        const uint32_t dcol = tmem_base + (uint32_t)((t + uu) * TN);
        const int uses = S - t;
        if (MODE == 0 || uses == 1) mma_i8<0>(dcol, ad, bd, idesc, acc);
        else if (uu == 0) mma_i8<1>(dcol, ad, bd, idesc, acc);
        else if (uu == uses - 1) mma_i8<3>(dcol, ad, bd, idesc, acc);
        else mma_i8<2>(dcol, ad, bd, idesc, acc);
      }
    }
  }
}
__device__ __forceinline__ uint32_t desc_lo(uint32_t smem_addr, uint32_t lbo) {
  return ((smem_addr >> 4) & 0x3FFFu) | (((lbo >> 4) & 0x3FFFu) << 16);
}
__device__ __forceinline__ uint32_t desc_hi(uint32_t sbo) { return ((sbo >> 4) & 0x3FFFu) | (1u << 14); }
__device__ __forceinline__ void tmem_ld8(uint32_t taddr, int32_t (&v)[8]) {
  asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];\n"
               : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7])
               : "r"(taddr));
}
__device__ __forceinline__ void tmem_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;\n" ::: "memory"); }

// K-major, no swizzle: start address, LBO (between the two 16-byte K halves), SBO (between 8-row groups), version 1
__device__ __forceinline__ uint64_t make_desc(uint32_t smem_addr, uint32_t lbo, uint32_t sbo) {
  return (uint64_t)((smem_addr >> 4) & 0x3FFFu) | ((uint64_t)((lbo >> 4) & 0x3FFFu) << 16) |
         ((uint64_t)((sbo >> 4) & 0x3FFFu) << 32) | (1ull << 46);
}

constexpr int kKstarUnroll = SE_KSTAR_UNROLL;
constexpr int RB = 8;  // radix bits
// exponent e with |x| / 2^e <= 0.99, so that the top digit of rint(x / 2^e * 2^(8 S - 1)) stays inside [-127, 127]
__host__ __device__ __forceinline__ int digit_exponent(double x) {
  int e = 0;
  if (x > 0.0 && frexp(x, &e) > 0.99) ++e;
  return e;
}
// Balanced radix-256 digits of four 63-bit fixed-point integers at once.  q = sum_t d_t 256^(7-t) with d_t in
// [-128, 127]  <=>  q + B = sum_t (d_t + 128) 256^(7-t) with B = 0x8080...80, i.e. the digits are the plain bytes of
// q + B with the top bit of every byte flipped: one 64-bit add and one xor instead of a carry chain.  A 4 x 4 byte
// transpose (PRMT) then gathers digit t of the four values into word[t] (byte e = value e), the unit the images store.
__device__ __forceinline__ void transpose4x4(uint32_t x0, uint32_t x1, uint32_t x2, uint32_t x3, uint32_t& b0,
                                             uint32_t& b1, uint32_t& b2, uint32_t& b3) {
  const uint32_t a = __byte_perm(x0, x1, 0x5140), b = __byte_perm(x2, x3, 0x5140);
  const uint32_t c = __byte_perm(x0, x1, 0x7362), d = __byte_perm(x2, x3, 0x7362);
  b0 = __byte_perm(a, b, 0x5410);
  b1 = __byte_perm(a, b, 0x7632);
  b2 = __byte_perm(c, d, 0x5410);
  b3 = __byte_perm(c, d, 0x7632);
}
__device__ __forceinline__ void digits4(const long long (&q)[4], uint32_t (&word)[MAX_S]) {
  static_assert(MAX_S == 8, "digits4 splits a 64-bit value into 8 byte digits");
  uint32_t lo[4], hi[4];
#pragma unroll
  for (int e = 0; e < 4; ++e) {
    const unsigned long long u = ((unsigned long long)q[e] + 0x8080808080808080ull) ^ 0x8080808080808080ull;
    lo[e] = (uint32_t)u;
    hi[e] = (uint32_t)(u >> 32);
  }
  // digit t = byte 7 - t of the 64-bit value: hi holds digits 3, 2, 1, 0 in its bytes 0 .. 3; lo holds 7, 6, 5, 4
  transpose4x4(hi[0], hi[1], hi[2], hi[3], word[3], word[2], word[1], word[0]);
  transpose4x4(lo[0], lo[1], lo[2], lo[3], word[7], word[6], word[5], word[4]);
}

// ---------------------------------------------------------------------------------------------------------------------
// model packing: per-row exponents of L^-1, scale factors of the epilogue, digit image of L^-1
// ---------------------------------------------------------------------------------------------------------------------
// scale_b[o][i] = 2^(eA_o + eB_oi - 12 - 21); exp_b[o][i] = eB_oi.  One warp per row.
static __global__ void oz_row_scale_kernel(const double* __restrict__ w, int n_pad, int ns, const int* __restrict__ exp_a,
                                    int* __restrict__ exp_b, double* __restrict__ scale_b) {
  const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (warp >= ns * n_pad) return;
  const int o = warp / n_pad, i = warp % n_pad;
  const double* row = w + ((int64_t)o * n_pad + i) * n_pad;
  double mx = 0.0;
  for (int j = lane; j <= i; j += 32) mx = fmax(mx, fabs(row[j]));
#pragma unroll
  for (int s = 16; s > 0; s >>= 1) mx = fmax(mx, __shfl_xor_sync(0xffffffffu, mx, s));
  if (lane == 0) {
    const int e = digit_exponent(mx);
    exp_b[warp] = e;
    scale_b[warp] = ldexp(1.0, exp_a[o] + e - 38);  // 2^(eA + eB - 14) from the digit weights, 2^-24 from `hi`
  }
}

// Digits are always taken from the 48-bit fixed-point value (7 digits); a model / workspace with S < 7 planes keeps the
// S most significant ones (the dropped balanced digits are at most half a unit of the last kept digit).
static __global__ void oz_pack_b_kernel(const double* __restrict__ w, int n_pad, const int* __restrict__ exp_b, int S,
                                 int8_t* __restrict__ b_img) {
  // one thread per 4 consecutive j of one row
  const int j0 = (blockIdx.x * blockDim.x + threadIdx.x) * 4;
  const int i = blockIdx.y, o = blockIdx.z;
  if (j0 >= n_pad) return;
  const double* row = w + ((int64_t)o * n_pad + i) * n_pad;
  const double sc = ldexp(1.0, RB * MAX_S - 1 - exp_b[o * n_pad + i]);
  uint32_t word[MAX_S];
  long long q[4];
#pragma unroll
  for (int e = 0; e < 4; ++e) q[e] = __double2ll_rn(row[j0 + e] * sc);
  digits4(q, word);
  const int nb = n_pad / TN, nkb = n_pad / KB;
  const int ib = i / TN, il = i % TN, kb = j0 / KB, c = j0 % KB;
  int8_t* dst = b_img + (((int64_t)o * nb + ib) * nkb + kb) * (int64_t)(S * B_BLK) + blk_off(il, c);
#pragma unroll
  for (int t = 0; t < MAX_S; ++t)
    if (t < S) *reinterpret_cast<uint32_t*>(dst + t * B_BLK) = word[t];
}

// ---------------------------------------------------------------------------------------------------------------------
// K1 (Ozaki flavour): kernel rows as S digit planes in the UMMA image layout + mean + mean-Jacobian.
// One warp owns R = 2 candidate rows (an even/odd pair shares every 32-byte sector of the image); a lane owns 4
// consecutive training points per 128-point block, so each digit plane is written with one 32-bit store per lane.
// ---------------------------------------------------------------------------------------------------------------------
template <int NS, int D, bool JAC, int R>
__global__ void __launch_bounds__(256, R == 1 ? 2 : 1)
gp_kstar_slices_kernel(GpDev<double> gp, const double* __restrict__ p, int64_t p_stride, const double* __restrict__ u,
                       int64_t u_stride, int64_t row0, int nrows, int rows_pad, const int* __restrict__ exp_a, int S,
                       int8_t* __restrict__ a_img, double* __restrict__ mean, double* __restrict__ jac) {
  const int lane = threadIdx.x & 31;
  const int64_t warp = blockIdx.x * (int64_t)(blockDim.x >> 5) + (threadIdx.x >> 5);
  const int64_t r0 = warp * R;
  if (r0 >= rows_pad) return;
  const int n_pad = gp.n_pad;
  const int nkb = n_pad / KB, rt_count = rows_pad / TM;

  double z[R][D];
  bool rvalid[R];
#pragma unroll
  for (int r = 0; r < R; ++r) {
    rvalid[r] = (r0 + r) < nrows;
    const int64_t c = row0 + r0 + r;
#pragma unroll
    for (int dd = 0; dd < D; ++dd) {
      double v = 0.0;
      if (rvalid[r]) v = dd < NS ? p[c * p_stride + dd] : u[c * u_stride + (dd - NS)];
      z[r][dd] = v;
    }
  }
  double qscale[NS], log_s[NS], nh_inv_l2[NS][D];
#pragma unroll
  for (int o = 0; o < NS; ++o) {
    log_s[o] = log(gp.outputscale[o]);
#pragma unroll
    for (int dd = 0; dd < D; ++dd) nh_inv_l2[o][dd] = -0.5 * gp.inv_l2[o][dd];
  }
#pragma unroll
  for (int o = 0; o < NS; ++o) qscale[o] = ldexp(1.0, RB * MAX_S - 1 - exp_a[o]);

  double m_acc[R][NS];
  double j_acc[R][NS][JAC ? D : 1];
#pragma unroll
  for (int r = 0; r < R; ++r)
#pragma unroll
    for (int o = 0; o < NS; ++o) {
      m_acc[r][o] = 0.0;
#pragma unroll
      for (int dd = 0; dd < (JAC ? D : 1); ++dd) j_acc[r][o][dd] = 0.0;
    }

  for (int jb = 0; jb < n_pad; jb += 128) {
    const int j0 = jb + 4 * lane;
    double x[D][4], al[NS][4];
#pragma unroll
    for (int dd = 0; dd < D; ++dd) {
      const double2 t0 = *reinterpret_cast<const double2*>(gp.xt + (int64_t)dd * n_pad + j0);
      const double2 t1 = *reinterpret_cast<const double2*>(gp.xt + (int64_t)dd * n_pad + j0 + 2);
      x[dd][0] = t0.x, x[dd][1] = t0.y, x[dd][2] = t1.x, x[dd][3] = t1.y;
    }
#pragma unroll
    for (int o = 0; o < NS; ++o) {
      const double2 t0 = *reinterpret_cast<const double2*>(gp.alpha + (int64_t)o * n_pad + j0);
      const double2 t1 = *reinterpret_cast<const double2*>(gp.alpha + (int64_t)o * n_pad + j0 + 2);
      al[o][0] = t0.x, al[o][1] = t0.y, al[o][2] = t1.x, al[o][3] = t1.y;
    }
    const int kb = j0 / KB, c = j0 % KB;
#pragma unroll
    for (int r = 0; r < R; ++r) {
      const int rc = (int)(r0 + r);  // row inside this chunk
      const int ct = rc / TM, rl = rc % TM;
      // Training points beyond n are zero-padded: their kernel values are finite garbage, but alpha and the columns
      // of L^-1 are exactly zero there, so neither the mean nor the contraction sees them (no predicate needed).
      long long q[NS][4];
#pragma unroll
      for (int e = 0; e < 4; ++e) {
        double diff[D], sq[D];
#pragma unroll
        for (int dd = 0; dd < D; ++dd) {
          diff[dd] = z[r][dd] - x[dd][e];
          sq[dd] = diff[dd] * diff[dd];
        }
#pragma unroll
        for (int o = 0; o < NS; ++o) {
          double arg = log_s[o];  // k = s exp(-1/2 sum_d diff_d^2 / l_d^2) = exp(ln s - sum_d diff_d^2 / (2 l_d^2))
#pragma unroll
          for (int dd = 0; dd < D; ++dd) arg = fma(sq[dd], nh_inv_l2[o][dd], arg);
          const double k = exp(arg);
          q[o][e] = __double2ll_rn(k * qscale[o]);
          const double wk = al[o][e] * k;
          m_acc[r][o] += wk;
          if (JAC) {
#pragma unroll
            for (int dd = 0; dd < D; ++dd) j_acc[r][o][dd] = fma(wk, diff[dd], j_acc[r][o][dd]);
          }
        }
      }
#pragma unroll
      for (int o = 0; o < NS; ++o) {
        uint32_t word[MAX_S];
        digits4(q[o], word);
        int8_t* dst = a_img + (((int64_t)o * rt_count + ct) * nkb + kb) * (int64_t)(S * A_BLK) + blk_off(rl, c);
#pragma unroll
        for (int t = 0; t < MAX_S; ++t)
          if (t < S) *reinterpret_cast<uint32_t*>(dst + t * A_BLK) = word[t];
      }
    }
  }

#pragma unroll
  for (int r = 0; r < R; ++r)
#pragma unroll
    for (int o = 0; o < NS; ++o) {
      double v = m_acc[r][o];
#pragma unroll
      for (int s = 16; s > 0; s >>= 1) v += __shfl_xor_sync(0xffffffffu, v, s);
      if (lane == 0 && rvalid[r]) mean[(row0 + r0 + r) * NS + o] = v;
      if (JAC) {
#pragma unroll
        for (int dd = 0; dd < D; ++dd) {
          double g = j_acc[r][o][dd];  // sum_i alpha_i k_i (z_d - x_id); d mean / d z_d = -that / l_d^2
#pragma unroll
          for (int s = 16; s > 0; s >>= 1) g += __shfl_xor_sync(0xffffffffu, g, s);
          if (lane == 0 && rvalid[r]) jac[((row0 + r0 + r) * NS + o) * D + dd] = -g * gp.inv_l2[o][dd];
        }
      }
    }
}

// ---------------------------------------------------------------------------------------------------------------------
// K1, row-group mapping (the default): one warp owns the 8 candidate rows of one core-matrix row group; lane = (row
// lane & 7, quad lane >> 3) owns 4 consecutive training points of every 16-point K half.  A digit-plane store of the
// warp is then one full 128-byte line (8 rows x 16 bytes, contiguous in the image), and a training point / alpha value
// is fetched once per 8 candidates instead of once per candidate (the one-row-per-warp kernel above pulls 80 KB through
// L1 per candidate at n = 2048 and sits at 84 % L1 throughput with the FP64 pipe at 41 %).
// ---------------------------------------------------------------------------------------------------------------------
// FUSE (small batches inside se_cem_solve): the kernel first advances its candidates' ellipsoids by the PREVIOUS
// horizon step (reach_step_body: variance from the GEMM's partial sums, ellipsoid step, CEM bookkeeping) and takes the
// new centres straight from registers -- one launch per horizon step less in a chain where every launch costs ~10 us.
// All four lanes of a candidate (and every training-set segment's CTA) run the step; one of them writes.  The step reads
// the old ellipsoids / segment sums and writes the new ones to different buffers (the host alternates them).
template <bool FUSE>
struct FuseArgs {};
template <>
struct FuseArgs<true> {
  ProbDev<double> pr;
  StepIO<double> io;
};

// OPT = outputs per thread.  OPT == NS: a thread carries all n_s outputs of its (row, points) cell (the (z - x)^2 terms
// are shared between the outputs).  OPT == 1: the CTA has NS times the warps and each warp group works on one output --
// the differences are recomputed per output, but a thread holds 1 / NS of the accumulators, so more warps fit an SM;
// the kernel waits on dependent FP64 issue (ncu: `wait` 2.5 per issue, 3.3 warps per scheduler at 128 registers), not
// on the pipe.  Same arithmetic per output either way: the results are bit-identical.
// SFIX > 0: the number of digit planes is a compile-time constant (the launcher passes S == SFIX): the eight predicated
// stores of a step, each with its own address arithmetic and compare against S, become SFIX plain stores off one base
// pointer -- 35 of the ~230 instructions a warp issues per 16-point step (the kernel is issue-bound: ncu shows 64 - 72 %
// of the issue slots busy, `not selected` and `math pipe throttle` the leading stall reasons).
template <int NS, int D, bool JAC, bool FUSE = false, int OPT = NS, int SFIX = 0>
// CTAs per SM: 8 (128 registers) is the measured optimum for n_s <= 3 (C2: 0.216 ms; 6 -> 0.279, 10 -> 0.226, 12 -> 0.240);
// with 4 and more outputs the accumulators no longer fit and 6 (168 registers) wins (C3: 3.10 vs 3.82 ms)
// (OPT == 1: 768 threads per SM -> 80 registers for inputs of up to 5 dimensions, 512 -> 128 registers beyond)
__global__ void __launch_bounds__(64 * (NS / OPT), OPT == NS ? (NS <= 3 ? 8 : 6)
                                                               : ((D <= 5 ? SE_KSTAR_TPS : 512) / (64 * NS) > 0 ? (D <= 5 ? SE_KSTAR_TPS : 512) / (64 * NS) : 1))
gp_kstar_slices8_kernel(GpDev<double> gp, const double* __restrict__ p, int64_t p_stride, const double* __restrict__ u,
                        int64_t u_stride, int64_t row0, int nrows, int rows_pad, const int* __restrict__ exp_a, int S_arg,
                        int8_t* __restrict__ a_img, double* __restrict__ mean, double* __restrict__ jac,
                        int seg_per_cta, double* __restrict__ kpart, FuseArgs<FUSE> fa) {
  static_assert(OPT == NS || OPT == 1, "a thread carries all outputs or one");
  const int S = SFIX > 0 ? SFIX : S_arg;
  constexpr int NT = 64 * (NS / OPT);  // threads per CTA
  const int lane = threadIdx.x & 31;
  const int o_base = OPT == NS ? 0 : (int)(threadIdx.x >> 6);  // this warp group's first output
  const int64_t warp = blockIdx.x * 2ll + ((threadIdx.x >> 5) & 1);
  const int64_t r0 = warp * 8;  // rows_pad is a multiple of 128 and the grid is rows_pad / 16 CTAs: every warp has rows
  const int r8 = lane & 7, quad = lane >> 3;
  const int64_t row = r0 + r8;
  const bool valid = row < nrows;
  const int n_pad = gp.n_pad;
  const int nkb = n_pad / KB, rt_count = rows_pad / TM;

  double z[D];
  if constexpr (FUSE) {
    double p1[NS];
#pragma unroll
    for (int dd = 0; dd < NS; ++dd) p1[dd] = 0.0;
    if (valid)
      reach_step_body<double, NS, D - NS>(fa.pr, fa.io, row0 + row, quad == 0 && blockIdx.y == 0 && o_base == 0, p1);
#pragma unroll
    for (int dd = 0; dd < D; ++dd) z[dd] = dd < NS ? p1[dd] : (valid ? u[(row0 + row) * u_stride + (dd - NS)] : 0.0);
  } else {
#pragma unroll
    for (int dd = 0; dd < D; ++dd) {
      double v = 0.0;
      if (valid) v = dd < NS ? p[(row0 + row) * p_stride + dd] : u[(row0 + row) * u_stride + (dd - NS)];
      z[dd] = v;
    }
  }
  // k = s exp(-1/2 sum_d ((z_d - x_d) / l_d)^2) is evaluated in scaled coordinates: with c_d = 1 / (sqrt 2 l_d),
  // e_d = z_d c_d - x_d c_d (one FMA) and arg = ln s - sum_d e_d^2 (one FMA each).  The fixed-point scale 2^shift of the
  // digit planes is folded into the exponential (exp_tab_scaled), so the sums below run on k 2^shift and are scaled back
  // by the exact power of two when a segment is folded -- bit-identical to summing k itself.
  double log_s[OPT], cs[OPT][D], zs[OPT][D], inv_q[OPT];
  int shift[OPT];
#pragma unroll
  for (int o = 0; o < OPT; ++o) {
    log_s[o] = log(gp.outputscale[o_base + o]);
    shift[o] = RB * MAX_S - 1 - exp_a[o_base + o];
    inv_q[o] = ldexp(1.0, -shift[o]);
#pragma unroll
    for (int dd = 0; dd < D; ++dd) {
      cs[o][dd] = sqrt(0.5 * gp.inv_l2[o_base + o][dd]);
      zs[o][dd] = z[dd] * cs[o][dd];
    }
  }
  // exp_nb() wants finite arguments: a non-finite input or hyper-parameter poisons the row's mean (and Jacobian)
  // instead, which is what the NaN check of the ellipsoid step looks at (gp_reachability_pytorch.py:234-243)
  bool poison = false;
#pragma unroll
  for (int dd = 0; dd < D; ++dd) poison |= !finite_f64(z[dd]);
#pragma unroll
  for (int o = 0; o < OPT; ++o) {
    poison |= !finite_f64(log_s[o]) && log_s[o] > 0.0;
    poison |= log_s[o] != log_s[o];
#pragma unroll
    for (int dd = 0; dd < D; ++dd) poison |= !finite_f64(cs[o][dd]);
  }
  // Mean and Jacobian sums are formed per 128-point segment (lane partials folded over the four quads) and the
  // segments are added in order -- by this CTA, or, when the training set is split over blockIdx.y because the batch
  // is too small to fill the machine, by gp_var_kernel from kpart.  Same additions in the same order either way, so
  // the result does not depend on how a batch was chunked or split.
  double m_acc[OPT], j_acc[OPT][JAC ? D : 1], m_tot[OPT], j_tot[OPT][JAC ? D : 1];
#pragma unroll
  for (int o = 0; o < OPT; ++o) {
    m_acc[o] = m_tot[o] = 0.0;
#pragma unroll
    for (int dd = 0; dd < (JAC ? D : 1); ++dd) j_acc[o][dd] = j_tot[o][dd] = 0.0;
  }
  // byte offset of (this lane's row, K index 4 quad) inside a digit-plane block; the K half adds 128
  const int ct = (int)(r0 / TM), rl = (int)(r0 % TM) + r8;
  const int64_t img_o = (int64_t)rt_count * nkb * (int64_t)(S * A_BLK);
  int8_t* const img_row = a_img + (int64_t)ct * nkb * (int64_t)(S * A_BLK) + blk_off(rl, 4 * quad) + o_base * img_o;

  // Training inputs and alpha go through shared memory: the CTA copies them in chunks of 128 points with cp.async
  // (no registers, double buffered, one chunk ahead), so a step reads its operands with short, fixed latency.  Straight
  // global loads leave ~40 % of the issue slots waiting on L2 (ncu: long scoreboard at the first use of x) because
  // ptxas sinks any register prefetch to the end of the loop body.
  constexpr int CH = 128;  // points per chunk: 1 KB per array
  __shared__ __align__(16) double stage[2][D + NS][CH];
  __shared__ double etab[64];  // exp_tab's table; published by the first barrier of the chunk loop
  if (threadIdx.x < 64) etab[threadIdx.x] = kExpTab[threadIdx.x];
  auto fill = [&](int chunk, int buf) {
    // 64 x 16 bytes = one array row of the chunk; the CTA's threads walk the (array, 16-byte piece) pairs
#pragma unroll
    for (int i = threadIdx.x; i < (D + NS) * 64; i += NT) {
      const int a = i >> 6, piece = i & 63;
      const double* src = (a < D ? gp.xt + (int64_t)a * n_pad : gp.alpha + (int64_t)(a - D) * n_pad) + chunk * CH +
                          2 * piece;
      const uint32_t dst = smem_u32(&stage[buf][a][2 * piece]);
      asm volatile("cp.async.cg.shared.global [%0], [%1], 16;\n" ::"r"(dst), "l"(src) : "memory");
    }
    asm volatile("cp.async.commit_group;\n" ::: "memory");
  };
  const int ch0 = blockIdx.y * seg_per_cta;
  const int chunks = min(n_pad / CH, ch0 + seg_per_cta);
  const bool split = gridDim.y > 1;
  fill(ch0, 0);
#pragma unroll 1
  for (int ch = ch0; ch < chunks; ++ch) {
    const int buf = (ch - ch0) & 1;
    if (ch + 1 < chunks) {
      fill(ch + 1, buf ^ 1);
      asm volatile("cp.async.wait_group 1;\n" ::: "memory");
    } else {
      asm volatile("cp.async.wait_group 0;\n" ::: "memory");
    }
    __syncthreads();
#pragma unroll kKstarUnroll
  for (int jl = 0; jl < CH; jl += 16) {
    const int j16 = ch * CH + jl;
    double x[D][4], al[OPT][4];
#pragma unroll
    for (int dd = 0; dd < D; ++dd) {
      const double2 t0 = *reinterpret_cast<const double2*>(&stage[buf][dd][jl + 4 * quad]);
      const double2 t1 = *reinterpret_cast<const double2*>(&stage[buf][dd][jl + 4 * quad + 2]);
      x[dd][0] = t0.x, x[dd][1] = t0.y, x[dd][2] = t1.x, x[dd][3] = t1.y;
    }
#pragma unroll
    for (int o = 0; o < OPT; ++o) {
      const double2 t0 = *reinterpret_cast<const double2*>(&stage[buf][D + o_base + o][jl + 4 * quad]);
      const double2 t1 = *reinterpret_cast<const double2*>(&stage[buf][D + o_base + o][jl + 4 * quad + 2]);
      al[o][0] = t0.x, al[o][1] = t0.y, al[o][2] = t1.x, al[o][3] = t1.y;
    }
    // zero-padded training points: finite garbage kernel values that meet alpha = 0 and zero columns of L^-1
    long long q[OPT][4];
#pragma unroll
    for (int e = 0; e < 4; ++e) {
#pragma unroll
      for (int o = 0; o < OPT; ++o) {
        double diff[D];
        double arg = log_s[o];
#pragma unroll
        for (int dd = 0; dd < D; ++dd) {
          diff[dd] = fma(-x[dd][e], cs[o][dd], zs[o][dd]);
          arg = fma(-diff[dd], diff[dd], arg);
        }
        const double kq = exp_tab_scaled(arg, etab, shift[o]);  // k 2^shift
        q[o][e] = __double2ll_rn(kq);
        const double wk = al[o][e] * kq;
        m_acc[o] += wk;
        if (JAC) {
#pragma unroll
          for (int dd = 0; dd < D; ++dd) j_acc[o][dd] = fma(wk, diff[dd], j_acc[o][dd]);
        }
      }
    }
    int8_t* const dst0 = img_row + (int64_t)(j16 >> 5) * (S * A_BLK) + ((j16 & 16) ? 128 : 0);
#pragma unroll
    for (int o = 0; o < OPT; ++o) {
      uint32_t word[MAX_S];
      digits4(q[o], word);
      int8_t* dst = dst0 + o * img_o;
#pragma unroll
      for (int t = 0; t < MAX_S; ++t)
        if (SFIX > 0 ? t < SFIX : t < S) *reinterpret_cast<uint32_t*>(dst + t * A_BLK) = word[t];
    }
  }
    // segment done: fold the quads (lanes r8, r8 + 8, r8 + 16, r8 + 24 hold disjoint training points), then add the
    // segment to the running total or hand it to gp_var_kernel
#pragma unroll
    for (int o = 0; o < OPT; ++o) {
      double v = m_acc[o];
      v += __shfl_xor_sync(0xffffffffu, v, 8);
      v += __shfl_xor_sync(0xffffffffu, v, 16);
      v *= inv_q[o];  // exact: a power of two
      m_acc[o] = 0.0;
      if (poison) v = __longlong_as_double(0x7ff8000000000000ll);
      if (split) {
        if (quad == 0) kpart[(((int64_t)ch * rows_pad + row) * NS + o_base + o) * (1 + D)] = v;
      } else {
        m_tot[o] += v;
      }
      if (JAC) {
#pragma unroll
        for (int dd = 0; dd < D; ++dd) {
          double g = j_acc[o][dd];
          g += __shfl_xor_sync(0xffffffffu, g, 8);
          g += __shfl_xor_sync(0xffffffffu, g, 16);
          g = g * inv_q[o] / cs[o][dd];  // back to sum_i alpha_i k_i (z_d - x_id)
          j_acc[o][dd] = 0.0;
          if (poison) g = __longlong_as_double(0x7ff8000000000000ll);
          if (split) {
            if (quad == 0) kpart[(((int64_t)ch * rows_pad + row) * NS + o_base + o) * (1 + D) + 1 + dd] = g;
          } else {
            j_tot[o][dd] += g;
          }
        }
      }
    }
    __syncthreads();  // everyone is done with this buffer before the chunk after next lands in it
  }
  if (split) return;  // gp_var_kernel adds the segments

#pragma unroll
  for (int o = 0; o < OPT; ++o) {
    double v = m_tot[o];
    if (poison) v = __longlong_as_double(0x7ff8000000000000ll);
    if (quad == 0 && valid) mean[(row0 + row) * NS + o_base + o] = v;
    if (JAC) {
#pragma unroll
      for (int dd = 0; dd < D; ++dd) {
        double g = j_tot[o][dd];  // sum_i alpha_i k_i (z_d - x_id); d mean / d z_d = -that / l_d^2
        if (poison) g = __longlong_as_double(0x7ff8000000000000ll);
        if (quad == 0 && valid) jac[((row0 + row) * NS + o_base + o) * D + dd] = -g * gp.inv_l2[o_base + o][dd];
      }
    }
  }
}

// ---------------------------------------------------------------------------------------------------------------------
// K2 (Ozaki flavour): partial[o][ib][c] = sum_{i in 64-row block ib} ( sum_{j <= i} W_o[i][j] K*_o[c][j] )^2
// Persistent, warp-specialised: bulk-copy producer / single-thread tcgen05.mma issuer / 4 epilogue warps.
// ---------------------------------------------------------------------------------------------------------------------
// K steps and MMA width of row tile ib when only n of the n_pad rows / columns are real (the rest are zero padding)
__device__ __forceinline__ int item_ksteps(int ib, int n) { return (min((ib + 1) * TN, n) + KB - 1) / KB; }
__device__ __forceinline__ int item_cols(int ib, int n) { return min(TN, (n - ib * TN + 15) & ~15); }

// item sequence of CTA b: round r -> position r * G + (r odd ? G - 1 - b : b)  (snake, so heavy/light items alternate)
__device__ __forceinline__ bool next_item(int round, const TriOrder& od, int total, TriItem* it) {
  const int G = gridDim.x, b = blockIdx.x;
  const int idx = round * G + ((round & 1) ? G - 1 - b : b);
  if (idx >= total) return false;
  *it = decode_item_chunked(idx, od);
  return true;
}

template <int S, int MODE>
__global__ void __launch_bounds__(THREADS, 1)
gp_trisumsq_ozaki_kernel(const int8_t* __restrict__ a_img, const int8_t* __restrict__ b_img,
                         const double* __restrict__ scale_b, int n, int n_pad, int rows_pad, TriOrder od,
                         double* __restrict__ partial, int dbg) {
  using C = Cfg<S>;
  extern __shared__ unsigned char smem_raw[];
  const uint32_t raw = smem_u32(smem_raw);
  const uint32_t base = (raw + 1023u) & ~1023u;                    // stage 0
  const uint32_t bars = base + C::STAGES * C::STAGE_BYTES;         // full[STAGES], empty[STAGES], tmem_full, tmem_empty
  const uint32_t tmem_slot = bars + (2 * C::STAGES + 2) * 8;
  auto full_bar = [&](int s) { return bars + 8u * s; };
  auto empty_bar = [&](int s) { return bars + 8u * (C::STAGES + s); };
  const uint32_t tmem_full = bars + 8u * (2 * C::STAGES), tmem_empty = tmem_full + 8u;

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int total = tri_total_items(od.nb, od.rt_count, od.ns);
  const int rounds = (total + gridDim.x - 1) / gridDim.x;
  const int nkb = n_pad / KB, nb_img = n_pad / TN;

  if (threadIdx.x == 0) {
    for (int s = 0; s < C::STAGES; ++s) {
      mbar_init(full_bar(s), 1);
      mbar_init(empty_bar(s), 1);
    }
    mbar_init(tmem_full, 1);
    mbar_init(tmem_empty, EPI_THREADS);
    mbar_fence_init();
  }
  if (warp == 1) tmem_alloc(tmem_slot, TMEM_COLS);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  uint32_t tmem_base;
  asm volatile("ld.shared.u32 %0, [%1];\n" : "=r"(tmem_base) : "r"(tmem_slot));

  if (warp == 0) {
    // ===== producer: one elected lane streams (A digits | B digits) of every K step into the stage ring =====
    if (lane == 0) {
      int stage = 0;
      uint32_t phase = 0;
      for (int r = 0; r < rounds; ++r) {
        TriItem it;
        if (!next_item(r, od, total, &it)) continue;
        const int8_t* a_src = a_img + ((int64_t)it.o * od.rt_count + it.rt) * nkb * (int64_t)(S * A_BLK);
        const int8_t* b_src = b_img + ((int64_t)it.o * nb_img + it.ib) * nkb * (int64_t)(S * B_BLK);
        const int nk = item_ksteps(it.ib, n);
        for (int kb = 0; kb < nk; kb += C::KS) {
          const int ks = min(C::KS, nk - kb);  // K steps in this stage: stage = [A digits of ks steps | B digits]
          mbar_wait(empty_bar(stage), phase ^ 1u);
          if (dbg & 1) {  // measurement only: no operand traffic
            mbar_arrive(full_bar(stage));
          } else if (dbg & 12) {  // measurement only: one operand's traffic (4: candidates' digits skipped, 8: L^-1's)
            const uint32_t dst = base + stage * C::STAGE_BYTES;
            if (dbg & 4) {
              mbar_expect_tx(full_bar(stage), ks * S * B_BLK);
              bulk_g2s(dst + C::KS * S * A_BLK, b_src + (int64_t)kb * (S * B_BLK), ks * S * B_BLK, full_bar(stage));
            } else {
              mbar_expect_tx(full_bar(stage), ks * S * A_BLK);
              bulk_g2s(dst, a_src + (int64_t)kb * (S * A_BLK), ks * S * A_BLK, full_bar(stage));
            }
          } else {
            mbar_expect_tx(full_bar(stage), ks * C::STEP_BYTES);
            const uint32_t dst = base + stage * C::STAGE_BYTES;
            bulk_g2s(dst, a_src + (int64_t)kb * (S * A_BLK), ks * S * A_BLK, full_bar(stage));
            bulk_g2s(dst + C::KS * S * A_BLK, b_src + (int64_t)kb * (S * B_BLK), ks * S * B_BLK, full_bar(stage));
          }
          if (++stage == C::STAGES) stage = 0, phase ^= 1u;
        }
      }
    }
  } else if (warp == 1) {
    // ===== MMA issuer: S (S + 1) / 2 int8 MMAs per K step into S level accumulators (64 TMEM columns each) =====
    // The whole warp walks the loop (so stage / phase / descriptors stay warp-uniform and live in uniform registers);
    // one elected lane issues the tcgen05 instructions.
    const uint32_t tb = __shfl_sync(0xffffffffu, tmem_base, 0);
    const uint32_t a_lo0 = desc_lo(__shfl_sync(0xffffffffu, base, 0), LBO), dhi = desc_hi(SBO);
    int stage = 0;
    uint32_t phase = 0, tphase = 0;
    for (int r = 0; r < rounds; ++r) {
      TriItem it;
      if (!next_item(r, od, total, &it)) continue;
      const int nk = item_ksteps(it.ib, n);
      // the last row tile may hold fewer than 64 training points: a narrower MMA (N is a multiple of 16) skips the padding
      const uint32_t idesc = make_idesc(TM, item_cols(it.ib, n));
      const bool full_width = item_cols(it.ib, n) == TN;
      mbar_wait(tmem_empty, tphase ^ 1u);  // the epilogue has drained the previous item's accumulators
      tc_fence_after();
      for (int kb = 0; kb < nk; kb += C::KS) {
        const int ks = min(C::KS, nk - kb);
        mbar_wait(full_bar(stage), phase);
        tc_fence_after();
        const uint32_t a_lo = a_lo0 + stage * (C::STAGE_BYTES >> 4);
        const uint32_t b_lo = a_lo + (C::KS * S * A_BLK >> 4);
        if (elect_one()) {
          if (MODE == 2 && !full_width) {  // ragged last row tile: per-pair MMAs of the narrower width
#pragma unroll
            for (int k = 0; k < C::KS; ++k)
              if (k < ks)
                issue_kstep<S, 1>(tb, a_lo + k * (S * A_BLK >> 4), b_lo + k * (S * B_BLK >> 4), dhi, idesc,
                                  kb + k > 0 ? 1u : 0u);
          } else {
#pragma unroll
            for (int k = 0; k < C::KS; ++k)
              if (k < ks)
                issue_kstep<S, MODE>(tb, a_lo + k * (S * A_BLK >> 4), b_lo + k * (S * B_BLK >> 4), dhi, idesc,
                                     kb + k > 0 ? 1u : 0u);
          }
          tc_commit(empty_bar(stage));  // frees the stage once these MMAs have read it
          if (kb + ks == nk) tc_commit(tmem_full);
        }
        __syncwarp();
        if (++stage == C::STAGES) stage = 0, phase ^= 1u;
      }
      tphase ^= 1u;
    }
  } else {
    // ===== epilogue: exact int64 recombination of the levels, one double conversion, square, row sum =====
    const int quad = warp & 3;  // TMEM lanes a warp may touch: 32 (warp % 4) ..
    const int half = (warp - 2) >> 2;  // two warps per lane quadrant, 32 of the 64 columns each
    const int row = quad * 32 + lane;
    const double lo_scale = S > 4 ? ldexp(1.0, 3 * RB - RB * (S - 1)) : 0.0;
    uint32_t tphase = 0;
    for (int r = 0; r < rounds; ++r) {
      TriItem it;
      if (!next_item(r, od, total, &it)) continue;
      mbar_wait(tmem_full, tphase);
      tc_fence_after();
      const double* sbp = scale_b + (int64_t)it.o * n_pad + (int64_t)it.ib * TN;
      const uint32_t tlane = tmem_base + ((uint32_t)(quad * 32) << 16);
      const int c_end = (dbg & 2) ? 0 : min((half + 1) * (TN / 2), item_cols(it.ib, n));
      // Phase 1 (the tensor core waits for it): drain this warp's 32 columns of every level and fold the levels into
      // two exact 64-bit integers per element (integer pipe only), then hand the accumulators back to the MMA warp.
      long long hi[TN / 2], lo[TN / 2];
#pragma unroll
      for (int b = 0; b < TN / 16; ++b) {
        const int c8 = half * (TN / 2) + b * 8;
        if (c8 < c_end) {  // warp-uniform: c_end is a multiple of 16
          int32_t d[S][8];
#pragma unroll
          for (int l = 0; l < S; ++l) tmem_ld8(tlane + (uint32_t)(l * TN + c8), d[l]);
          tmem_ld_wait();
#pragma unroll
          for (int j = 0; j < 8; ++j) {
            long long h = 0, w = 0;
#pragma unroll
            for (int l = 0; l < S && l < 4; ++l) h += (long long)d[l][j] << (RB * (3 - l));
#pragma unroll
            for (int l = 4; l < S; ++l) w += (long long)d[l][j] << (RB * (S - 1 - l));
            hi[b * 8 + j] = h, lo[b * 8 + j] = w;
          }
        } else {
#pragma unroll
          for (int j = 0; j < 8; ++j) hi[b * 8 + j] = 0, lo[b * 8 + j] = 0;
        }
      }
      tc_fence_before();
      mbar_arrive(tmem_empty);
      // Phase 2 (overlaps the next item's MMAs): one conversion to double per half, row scale, square, row sum.
      double sum = 0.0;
#pragma unroll
      for (int e = 0; e < TN / 2; ++e) {
        const double v = ((double)hi[e] + (double)lo[e] * lo_scale) * __ldg(sbp + half * (TN / 2) + e);
        sum += v * v;
      }
      // the two column halves are separate partial blocks (gp_var_kernel sums 2 nb of them in a fixed order)
      partial[((int64_t)it.o * (2 * od.nb) + 2 * it.ib + half) * rows_pad + (int64_t)it.rt * TM + row] = sum;
      tphase ^= 1u;
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem_base, TMEM_COLS);
}

// ---------------------------------------------------------------------------------------------------------------------
// K2, cluster flavour: a thread-block cluster of CA x CB CTAs works on CA neighbouring row tiles of L^-1 times CB
// neighbouring candidate tiles (CTA rank = rb CA + ra takes row tile CA ibg + ra and candidate tile CB rtg + rb).
// Digit planes are fetched from L2 once per GROUP instead of once per CTA: the CB CTAs that share a row tile each issue one
// cp.async.bulk ... .multicast::cluster for 1 / CB of the stage's L^-1 bytes, the CA CTAs that share a candidate tile each
// 1 / CA of its kernel-row bytes; a multicast lands at the same shared-memory offset in every CTA of its mask and signals
// the full barrier there.  A stage may only be refilled when every CTA that receives the copy has read it, so a stage's
// empty barrier counts CA + CB - 1 arrivals and the MMA warps commit to all CTAs that write into them
// (tcgen05.commit ... .multicast::cluster).  All CTAs of a cluster run the same item sequence with the K extent of the
// group's deepest row tile -- the extra K steps of the shallower tiles multiply the exact zeros that the image holds
// above the diagonal, so the level sums (integers) are unchanged -- and the rings stay in lock step by construction.
// L2 -> SM traffic per K step per SM at S = 6: 24 / CA + 12 / CB KB instead of 36.  Same partial-sum layout.
// ---------------------------------------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t cluster_ctarank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;\n" : "=r"(r));
  return r;
}
__device__ __forceinline__ void cluster_sync_all() {
  asm volatile("barrier.cluster.arrive.release.aligned;\n" ::: "memory");
  asm volatile("barrier.cluster.wait.acquire.aligned;\n" ::: "memory");
}
__device__ __forceinline__ void bulk_g2s_multicast(uint32_t dst, const void* src, uint32_t bytes, uint32_t bar,
                                                   uint16_t cta_mask) {
  asm volatile(
      "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes.multicast::cluster [%0], [%1], %2, [%3], %4;\n" ::"r"(dst),
      "l"(src), "r"(bytes), "r"(bar), "h"(cta_mask)
      : "memory");
}
__device__ __forceinline__ void tc_commit_multicast(uint32_t bar, uint16_t cta_mask) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;\n" ::"r"(bar),
               "h"(cta_mask)
               : "memory");
}

// item sequence of cluster b (of G clusters): as next_item, over (output, candidate-tile group, row-tile group)
template <int CA, int CB>
__device__ __forceinline__ bool next_group_item(int round, const TriOrder& od_grp, int total, int ra, int rb, TriItem* it) {
  const int G = gridDim.x / (CA * CB), b = blockIdx.x / (CA * CB);
  const int idx = round * G + ((round & 1) ? G - 1 - b : b);
  if (idx >= total) return false;
  *it = decode_item_chunked(idx, od_grp);
  it->rt = CB * it->rt + rb;
  it->ib = CA * it->ib + ra;
  return true;
}

template <int S, int CA, int CB>
__global__ void __launch_bounds__(THREADS, 1)
gp_trisumsq_ozaki_cluster_kernel(const int8_t* __restrict__ a_img, const int8_t* __restrict__ b_img,
                                 const double* __restrict__ scale_b, int n, int n_pad, int rows_pad, TriOrder od_grp,
                                 int rt_count, int nb, double* __restrict__ partial) {
  using C = Cfg<S>;
  static_assert((S * A_BLK) % (16 * CA) == 0 && (S * B_BLK) % (16 * CB) == 0, "a multicast piece is a multiple of 16 bytes");
  static_assert(CA <= 2, "the image of L^-1 holds an even number of row tiles (n_pad is a multiple of 128), not a multiple of 4");
  extern __shared__ unsigned char smem_raw[];
  const uint32_t raw = smem_u32(smem_raw);
  const uint32_t base = (raw + 1023u) & ~1023u;
  const uint32_t bars = base + C::STAGES * C::STAGE_BYTES;
  const uint32_t tmem_slot = bars + (2 * C::STAGES + 2) * 8;
  auto full_bar = [&](int s) { return bars + 8u * s; };
  auto empty_bar = [&](int s) { return bars + 8u * (C::STAGES + s); };
  const uint32_t tmem_full = bars + 8u * (2 * C::STAGES), tmem_empty = tmem_full + 8u;

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int crank = (int)cluster_ctarank(), ra = crank % CA, rb = crank / CA;
  // CTAs that share this CTA's candidate tile (same rb) / its row tile (same ra); every CTA that writes into this one
  uint16_t mask_a = 0, mask_b = 0;
#pragma unroll
  for (int a = 0; a < CA; ++a) mask_a |= (uint16_t)(1u << (rb * CA + a));
#pragma unroll
  for (int b = 0; b < CB; ++b) mask_b |= (uint16_t)(1u << (b * CA + ra));
  const uint16_t mask_all = mask_a | mask_b;
  const int total = tri_total_items(od_grp.nb, od_grp.rt_count, od_grp.ns);
  const int G = gridDim.x / (CA * CB);
  const int rounds = (total + G - 1) / G;
  const int nkb = n_pad / KB, nb_img = n_pad / TN;
  // K extent of a group = that of its deepest row tile (rows of the image beyond nb are zero padding)
  auto group_ksteps = [&](int ib) { return item_ksteps(min(ib - ib % CA + CA - 1, nb_img - 1), n); };

  if (threadIdx.x == 0) {
    for (int s = 0; s < C::STAGES; ++s) {
      mbar_init(full_bar(s), 1);
      mbar_init(empty_bar(s), CA + CB - 1);  // every CTA this one receives copies from (itself included) releases a stage
    }
    mbar_init(tmem_full, 1);
    mbar_init(tmem_empty, EPI_THREADS);
    mbar_fence_init();
  }
  if (warp == 1) tmem_alloc(tmem_slot, TMEM_COLS);
  tc_fence_before();
  cluster_sync_all();  // the peers' barriers are initialised before anything is multicast or committed to them
  tc_fence_after();
  uint32_t tmem_base;
  asm volatile("ld.shared.u32 %0, [%1];\n" : "=r"(tmem_base) : "r"(tmem_slot));

  if (warp == 0) {
    if (lane == 0) {
      int stage = 0;
      uint32_t phase = 0;
      for (int r = 0; r < rounds; ++r) {
        TriItem it;
        if (!next_group_item<CA, CB>(r, od_grp, total, ra, rb, &it)) continue;
        const int8_t* a_src = a_img + ((int64_t)it.o * rt_count + it.rt) * nkb * (int64_t)(S * A_BLK);
        const int8_t* b_src = b_img + ((int64_t)it.o * nb_img + it.ib) * nkb * (int64_t)(S * B_BLK);
        const int nk = group_ksteps(it.ib);
        for (int kb = 0; kb < nk; kb += C::KS) {
          const int ks = min(C::KS, nk - kb);
          mbar_wait(empty_bar(stage), phase ^ 1u);
          mbar_expect_tx(full_bar(stage), ks * C::STEP_BYTES);
          const uint32_t dst = base + stage * C::STAGE_BYTES;
          // this CTA's piece of the stage's kernel-row bytes -> the CTAs on its candidate tile; of its L^-1 bytes -> the
          // CTAs on its row tile
          const uint32_t a_piece = ks * S * A_BLK / CA, b_piece = ks * S * B_BLK / CB;
          if (CA == 1)
            bulk_g2s(dst, a_src + (int64_t)kb * (S * A_BLK), a_piece, full_bar(stage));
          else
            bulk_g2s_multicast(dst + ra * a_piece, a_src + (int64_t)kb * (S * A_BLK) + ra * a_piece, a_piece,
                               full_bar(stage), mask_a);
          if (CB == 1)
            bulk_g2s(dst + C::KS * S * A_BLK, b_src + (int64_t)kb * (S * B_BLK), b_piece, full_bar(stage));
          else
            bulk_g2s_multicast(dst + C::KS * S * A_BLK + rb * b_piece, b_src + (int64_t)kb * (S * B_BLK) + rb * b_piece,
                               b_piece, full_bar(stage), mask_b);
          if (++stage == C::STAGES) stage = 0, phase ^= 1u;
        }
      }
    }
  } else if (warp == 1) {
    const uint32_t tb = __shfl_sync(0xffffffffu, tmem_base, 0);
    const uint32_t a_lo0 = desc_lo(__shfl_sync(0xffffffffu, base, 0), LBO), dhi = desc_hi(SBO);
    int stage = 0;
    uint32_t phase = 0, tphase = 0;
    for (int r = 0; r < rounds; ++r) {
      TriItem it;
      if (!next_group_item<CA, CB>(r, od_grp, total, ra, rb, &it)) continue;
      const int nk = group_ksteps(it.ib);
      // a row tile beyond nb is padding (zero rows): the narrowest MMA keeps the protocol going at next to no cost
      const int cols = it.ib < nb ? item_cols(it.ib, n) : 16;
      const uint32_t idesc = make_idesc(TM, cols);
      const bool full_width = cols == TN;
      mbar_wait(tmem_empty, tphase ^ 1u);
      tc_fence_after();
      for (int kb = 0; kb < nk; kb += C::KS) {
        const int ks = min(C::KS, nk - kb);
        mbar_wait(full_bar(stage), phase);
        tc_fence_after();
        const uint32_t a_lo = a_lo0 + stage * (C::STAGE_BYTES >> 4);
        const uint32_t b_lo = a_lo + (C::KS * S * A_BLK >> 4);
        if (elect_one()) {
          if (!full_width) {
#pragma unroll
            for (int k = 0; k < C::KS; ++k)
              if (k < ks)
                issue_kstep<S, 1>(tb, a_lo + k * (S * A_BLK >> 4), b_lo + k * (S * B_BLK >> 4), dhi, idesc,
                                  kb + k > 0 ? 1u : 0u);
          } else {
#pragma unroll
            for (int k = 0; k < C::KS; ++k)
              if (k < ks)
                issue_kstep<S, 2>(tb, a_lo + k * (S * A_BLK >> 4), b_lo + k * (S * B_BLK >> 4), dhi, idesc,
                                  kb + k > 0 ? 1u : 0u);
          }
          tc_commit_multicast(empty_bar(stage), mask_all);  // this CTA is done with the stage: tell every writer
          if (kb + ks == nk) tc_commit(tmem_full);
        }
        __syncwarp();
        if (++stage == C::STAGES) stage = 0, phase ^= 1u;
      }
      tphase ^= 1u;
    }
  } else {
    const int quad = warp & 3;
    const int half = (warp - 2) >> 2;
    const int row = quad * 32 + lane;
    const double lo_scale = S > 4 ? ldexp(1.0, 3 * RB - RB * (S - 1)) : 0.0;
    uint32_t tphase = 0;
    for (int r = 0; r < rounds; ++r) {
      TriItem it;
      if (!next_group_item<CA, CB>(r, od_grp, total, ra, rb, &it)) continue;
      mbar_wait(tmem_full, tphase);
      tc_fence_after();
      const bool real = it.ib < nb;
      const double* sbp = scale_b + (int64_t)it.o * n_pad + (int64_t)it.ib * TN;
      const uint32_t tlane = tmem_base + ((uint32_t)(quad * 32) << 16);
      const int c_end = real ? min((half + 1) * (TN / 2), item_cols(it.ib, n)) : 0;
      long long hi[TN / 2], lo[TN / 2];
#pragma unroll
      for (int b = 0; b < TN / 16; ++b) {
        const int c8 = half * (TN / 2) + b * 8;
        if (c8 < c_end) {
          int32_t d[S][8];
#pragma unroll
          for (int l = 0; l < S; ++l) tmem_ld8(tlane + (uint32_t)(l * TN + c8), d[l]);
          tmem_ld_wait();
#pragma unroll
          for (int j = 0; j < 8; ++j) {
            long long h = 0, w = 0;
#pragma unroll
            for (int l = 0; l < S && l < 4; ++l) h += (long long)d[l][j] << (RB * (3 - l));
#pragma unroll
            for (int l = 4; l < S; ++l) w += (long long)d[l][j] << (RB * (S - 1 - l));
            hi[b * 8 + j] = h, lo[b * 8 + j] = w;
          }
        } else {
#pragma unroll
          for (int j = 0; j < 8; ++j) hi[b * 8 + j] = 0, lo[b * 8 + j] = 0;
        }
      }
      tc_fence_before();
      mbar_arrive(tmem_empty);
      if (real) {
        double sum = 0.0;
#pragma unroll
        for (int e = 0; e < TN / 2; ++e) {
          const double v = ((double)hi[e] + (double)lo[e] * lo_scale) * __ldg(sbp + half * (TN / 2) + e);
          sum += v * v;
        }
        partial[((int64_t)it.o * (2 * nb) + 2 * it.ib + half) * rows_pad + (int64_t)it.rt * TM + row] = sum;
      }
      tphase ^= 1u;
    }
  }
  tc_fence_before();
  cluster_sync_all();  // peers may still be arriving on this CTA's barriers: leave together
  if (warp == 1) tmem_dealloc(tmem_base, TMEM_COLS);
}

// ---------------------------------------------------------------------------------------------------------------------
// Probe (tests): one CTA, nk K steps of a single 128 x 64 tile; dumps every level accumulator.  out[S][128][64] int32.
// lbo / sbo are parameters so the descriptor convention is pinned by a test and not by a comment.
// ---------------------------------------------------------------------------------------------------------------------
template <int S, int MODE>
__global__ void __launch_bounds__(128, 1)
oz_probe_kernel(const int8_t* __restrict__ a_img, const int8_t* __restrict__ b_img, int nk, uint32_t lbo, uint32_t sbo,
                int32_t* __restrict__ out) {
  extern __shared__ unsigned char smem_raw[];
  const uint32_t raw = smem_u32(smem_raw);
  const uint32_t base = (raw + 1023u) & ~1023u;
  const uint32_t bars = base + S * (A_BLK + B_BLK);
  const uint32_t full = bars, done = bars + 8, tmem_slot = bars + 16;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (threadIdx.x == 0) {
    mbar_init(full, 1);
    mbar_init(done, 1);
    mbar_fence_init();
  }
  if (warp == 0) tmem_alloc(tmem_slot, TMEM_COLS);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  uint32_t tmem_base;
  asm volatile("ld.shared.u32 %0, [%1];\n" : "=r"(tmem_base) : "r"(tmem_slot));
  constexpr uint32_t idesc = make_idesc(TM, TN);
  for (int kb = 0; kb < nk; ++kb) {
    if (threadIdx.x == 0) {
      mbar_expect_tx(full, S * (A_BLK + B_BLK));
      bulk_g2s(base, a_img + (int64_t)kb * S * A_BLK, S * A_BLK, full);
      bulk_g2s(base + S * A_BLK, b_img + (int64_t)kb * S * B_BLK, S * B_BLK, full);
      mbar_wait(full, kb & 1);
      tc_fence_after();
      issue_kstep<S, MODE>(tmem_base, desc_lo(base, lbo), desc_lo(base + S * A_BLK, lbo), desc_hi(sbo), idesc, kb > 0 ? 1u : 0u);
      tc_commit(done);
      mbar_wait(done, kb & 1);  // single stage: the next K step overwrites the operands
    }
    __syncthreads();
  }
  tc_fence_after();
  const int row = warp * 32 + lane;
  const uint32_t tlane = tmem_base + ((uint32_t)(warp * 32) << 16);
  for (int l = 0; l < S; ++l)
    for (int c8 = 0; c8 < TN; c8 += 8) {
      int32_t v[8];
      tmem_ld8(tlane + (uint32_t)(l * TN + c8), v);
      tmem_ld_wait();
#pragma unroll
      for (int j = 0; j < 8; ++j) out[((int64_t)l * TM + row) * TN + c8 + j] = v[j];
    }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) tmem_dealloc(tmem_base, TMEM_COLS);
}

// ---------------------------------------------------------------------------------------------------------------------
// Pipe peak (se_measure_peak kind 3): back-to-back 128 x 256 x 32 s8 MMAs on every SM, operands resident in shared
// memory, no loads and no epilogue -- the denominator bench.py quotes for the tcgen05 variance GEMM.
// ---------------------------------------------------------------------------------------------------------------------
static __global__ void __launch_bounds__(128, 1) oz_peak_kernel(int iters, int n_mma, int issuers, int32_t* __restrict__ sink) {
  extern __shared__ unsigned char smem_raw[];
  const uint32_t raw = smem_u32(smem_raw);
  const uint32_t base = (raw + 1023u) & ~1023u;
  const uint32_t bar = base + 12 * 1024, tmem_slot = bar + 16;
  const int warp = threadIdx.x >> 5;
  // pseudo-random digit bytes: the power drawn by the tensor pipe (and so the sustained clock) depends on the data
  for (int i = threadIdx.x; i < 12 * 1024 / 4; i += blockDim.x) {
    uint32_t h = (uint32_t)i * 2654435761u + blockIdx.x * 40503u;
    h ^= h >> 15, h *= 2246822519u, h ^= h >> 13, h *= 3266489917u, h ^= h >> 16;
    reinterpret_cast<uint32_t*>(smem_raw + (base - raw))[i] = h;
  }
  if (threadIdx.x == 0) {
    mbar_init(bar, 1);
    mbar_init(bar + 8, 1);
    mbar_fence_init();
  }
  if (warp == 0) tmem_alloc(tmem_slot, TMEM_COLS);
  asm volatile("fence.proxy.async.shared::cta;\n" ::: "memory");  // generic-proxy smem writes -> async proxy (UMMA)
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  uint32_t tmem_base;
  asm volatile("ld.shared.u32 %0, [%1];\n" : "=r"(tmem_base) : "r"(tmem_slot));
  // issuers = 2 (n_mma = 64 shape only): two warps issue half of the MMAs each, into their own 256 TMEM columns -- does
  // the time one thread needs between two tcgen05.mma limit the pipe, or the pipe itself?
  if (warp < issuers) {
    const uint32_t tb = __shfl_sync(0xffffffffu, tmem_base, 0) + (uint32_t)(warp * 256);
    const uint32_t alo = desc_lo(__shfl_sync(0xffffffffu, base, 0), LBO), dhi = desc_hi(SBO);
    const uint32_t idesc = make_idesc(128, n_mma);
    const int my_iters = iters / issuers;
    if (elect_one()) {
      const uint64_t ad = pack_desc(alo, dhi), bd = pack_desc(alo + (4096 >> 4), dhi);
      if (n_mma == 256) {
        for (int i = 0; i < my_iters; i += 2) {
          mma_i8<0>(tb, ad, bd, idesc, 1u);
          mma_i8<0>(issuers == 1 ? tb + 256 : tb, ad, bd, idesc, 1u);
        }
      } else {  // the variance GEMM's own shape: one A tile kept in the collector across several N = 64 MMAs
        const uint64_t bd1 = pack_desc(alo + (6144 >> 4), dhi), bd2 = pack_desc(alo + (8192 >> 4), dhi);
        const uint64_t bd3 = pack_desc(alo + (10240 >> 4), dhi);
        const uint32_t span = issuers == 2 ? 256u : 512u;  // TMEM columns this issuer owns
        const uint32_t c1 = tb + (uint32_t)n_mma % span, c2 = tb + (uint32_t)(2 * n_mma) % span;
        const uint32_t c3 = tb + (uint32_t)(3 * n_mma) % span;
        for (int i = 0; i < my_iters; i += 4) {
          mma_i8<1>(tb, ad, bd, idesc, 1u);
          mma_i8<2>(c1, ad, bd1, idesc, 1u);
          mma_i8<2>(c2, ad, bd2, idesc, 1u);
          mma_i8<3>(c3, ad, bd3, idesc, 1u);
        }
      }
      tc_commit(bar + 8 * warp);
    }
    __syncwarp();
  }
  mbar_wait(bar, 0);
  if (issuers == 2) mbar_wait(bar + 8, 0);
  tc_fence_after();
  if (sink == reinterpret_cast<int32_t*>(1)) {  // never true: keeps the accumulators observable
    int32_t v[8];
    tmem_ld8(tmem_base, v);
    tmem_ld_wait();
    sink[0] = v[0];
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) tmem_dealloc(tmem_base, TMEM_COLS);
}

}  // namespace oz
}  // namespace se
