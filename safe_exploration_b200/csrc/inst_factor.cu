// inst_factor.cu -- C ABI of the on-device factor refresh (factor_kernels.cuh): L^-1 block row by block row, and
// (K + noise I)^-1 = W^T W for the marginal-likelihood gradient.  No cuSOLVER / cuBLAS.
#include <cstdlib>

#include "factor_kernels.cuh"

namespace se {

static int launch_gemm_nt(const fac::GemmNT& p, int batch, cudaStream_t st) {
  if (p.m_tiles <= 0 || p.n_tiles <= 0 || batch <= 0) return 0;
  SE_TRY(set_max_dyn_smem(fac::dgemm_nt_kernel, fac::Cfg::SMEM_BYTES));
  const dim3 grid((unsigned)(p.m_tiles * p.n_tiles), (unsigned)batch, (unsigned)(p.splits > 1 ? p.splits : 1));
  SE_LAUNCH(fac::dgemm_nt_kernel, grid, fac::Cfg::THREADS, fac::Cfg::SMEM_BYTES, st, p);
  return 0;
}

// The Schur complement + small Cholesky of a block row (a few CTAs, latency-bound) run on a side stream beside the
// T^T product (throughput-bound) -- both only need L21.  One side stream per device, created on first use; the two events
// that order it against the caller's stream belong to the call (concurrent calls on one device must not share them).
static int factor_side_stream(cudaStream_t* out) {
  static std::mutex mu;
  static std::map<int, cudaStream_t> per_device;
  int dev = 0;
  SE_CUDA(cudaGetDevice(&dev));
  std::lock_guard<std::mutex> lock(mu);
  cudaStream_t& s = per_device[dev];
  if (!s) SE_CUDA(cudaStreamCreateWithFlags(&s, cudaStreamNonBlocking));
  *out = s;
  return 0;
}
struct FactorEvents {
  cudaEvent_t ev_l21 = nullptr, ev_chol = nullptr;
  int create() {
    SE_CUDA(cudaEventCreateWithFlags(&ev_l21, cudaEventDisableTiming));
    SE_CUDA(cudaEventCreateWithFlags(&ev_chol, cudaEventDisableTiming));
    return 0;
  }
  ~FactorEvents() {  // destroying a recorded event is allowed: its resources go when the work has passed it
    if (ev_l21) cudaEventDestroy(ev_l21);
    if (ev_chol) cudaEventDestroy(ev_chol);
  }
};

constexpr int kMaxProductSplits = 8;  // K slices of the two long products of a block row

// C = sum of `splits` K slices of A B^T: the product writes its slices to `slices`, sum_slices_kernel adds them into g.c.
// Few tiles (early block rows, small models) and tiles of very different depth (the K extent of a tile follows the
// triangle) both leave SMs idle with one CTA per tile; equal slices of every tile give ~4 units per CTA slot.
static int launch_gemm_nt_sliced(fac::GemmNT g, int batch, int rows, int cols, double* slices, int sm_count,
                                 cudaStream_t st) {
  static const int env = getenv("SE_FACTOR_SPLITS") ? atoi(getenv("SE_FACTOR_SPLITS")) : -1;  // -1: automatic, 1: off
  const int64_t tiles = (int64_t)g.m_tiles * g.n_tiles * batch, slots = 2ll * sm_count;
  int splits = env > 0 ? env : (int)((4 * slots + tiles - 1) / tiles);
  if (splits > kMaxProductSplits) splits = kMaxProductSplits;
  if (splits > g.k / 128) splits = g.k / 128 > 0 ? g.k / 128 : 1;  // at least 128 deep
  if (splits <= 1) return launch_gemm_nt(g, batch, st);
  fac::GemmNT s = g;
  const int64_t elems = (int64_t)rows * g.ldc;  // one slice of one batch entry, laid out like the destination
  s.c = slices, s.c_batch = (int64_t)kMaxProductSplits * elems, s.c_split = elems, s.ct = nullptr, s.splits = splits;
  s.alpha = g.alpha;
  SE_TRY(launch_gemm_nt(s, batch, st));
  SE_REQUIRE(g.ct == nullptr, "sliced products do not write a transposed copy");
  const int64_t work = (int64_t)rows * (cols / 2);
  SE_LAUNCH(fac::sum_slices_kernel, dim3((unsigned)ceil_div(work, (int64_t)256), (unsigned)batch), 256, 0, st, slices,
            splits, elems, (int64_t)kMaxProductSplits * elems, g.ldc, rows, cols, g.c, g.ldc, g.c_batch);
  return 0;
}

}  // namespace se

using namespace se;

// doubles: L21 [n_s][FB][n_pad] | split-K slices of the Schur complement [n_s][kMaxSplits][FB * FB] | T^T [n_s][n_pad][FB]
// | K slices of the two long products [n_s][kMaxProductSplits][FB * n_pad]
extern "C" int64_t se_gp_factor_workspace(int n_s, int64_t n_pad) {
  return (int64_t)sizeof(double) * n_s *
         ((2 + kMaxProductSplits) * fac::FB * n_pad + (int64_t)fac::kMaxSplits * fac::FB * fac::FB);
}

extern "C" int se_gp_factor(int n_s, int64_t n_pad, int first_block, const double* dev_k, double* dev_w,
                            double* dev_wt, double* dev_diag, double* dev_workspace, int32_t* dev_info,
                            void* stream) {
  SE_REQUIRE(n_s >= 1 && n_s <= SE_MAX_NS, "se_gp_factor: n_s = %d outside 1..%d", n_s, SE_MAX_NS);
  SE_REQUIRE(n_pad >= fac::FB && n_pad % fac::FB == 0 && n_pad <= (1 << 20), "se_gp_factor: n_pad = %lld must be a "
             "positive multiple of %d", (long long)n_pad, fac::FB);
  const int blocks = (int)(n_pad / fac::FB);
  SE_REQUIRE(first_block >= 0 && first_block <= blocks, "se_gp_factor: first_block = %d outside 0..%d", first_block,
             blocks);
  SE_REQUIRE(dev_k && dev_w && dev_wt && dev_diag && dev_workspace && dev_info, "se_gp_factor: NULL device pointer");
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const int64_t ld = n_pad, mat = n_pad * n_pad;
  double* l21 = dev_workspace;                                   // [n_s][FB][n_pad]
  double* part = l21 + (int64_t)n_s * fac::FB * n_pad;           // [n_s][kMaxSplits][FB * FB]
  double* tt = part + (int64_t)n_s * fac::kMaxSplits * fac::FB * fac::FB;  // [n_s][n_pad][FB]
  double* slices = tt + (int64_t)n_s * n_pad * fac::FB;  // [n_s][kMaxProductSplits][FB * n_pad]
  int dev = 0, sm_count = 148;
  SE_CUDA(cudaGetDevice(&dev));
  SE_CUDA(cudaDeviceGetAttribute(&sm_count, cudaDevAttrMultiProcessorCount, dev));
  SE_TRY(set_max_dyn_smem(fac::chol_inv_block_kernel, fac::kCholSmemBytes));
  cudaStream_t side = nullptr;
  FactorEvents ev;
  SE_TRY(factor_side_stream(&side));
  SE_TRY(ev.create());
  static const int overlap = getenv("SE_FACTOR_OVERLAP") ? atoi(getenv("SE_FACTOR_OVERLAP")) : 1;
  for (int b = first_block; b < blocks; ++b) {
    const int n1 = b * fac::FB;
    int splits = 0;
    cudaStream_t sc = st;  // stream of the Schur complement + Cholesky
    if (n1 > 0) {
      fac::GemmNT g{};
      // L21 = K21 W11^T
      g.a = dev_k + (int64_t)n1 * ld, g.lda = ld, g.a_batch = mat;
      g.b = dev_w, g.ldb = ld, g.b_batch = mat;
      g.c = l21, g.ldc = ld, g.c_batch = (int64_t)fac::FB * n_pad, g.c_split = 0;
      g.alpha = 1.0, g.m_tiles = fac::FB / fac::Cfg::BM, g.n_tiles = b, g.k = n1, g.kmode = fac::K_B_LOWER, g.splits = 1;
      SE_TRY(launch_gemm_nt_sliced(g, n_s, fac::FB, n1, slices, sm_count, st));
      if (overlap) {
        SE_CUDA(cudaEventRecord(ev.ev_l21, st));
        SE_CUDA(cudaStreamWaitEvent(side, ev.ev_l21, 0));
        sc = side;
      }
      // split-K slices of L21 L21^T
      splits = n1 / 256 < 1 ? 1 : (n1 / 256 > fac::kMaxSplits ? fac::kMaxSplits : n1 / 256);
      fac::GemmNT s{};
      s.a = l21, s.lda = ld, s.a_batch = (int64_t)fac::FB * n_pad;
      s.b = l21, s.ldb = ld, s.b_batch = (int64_t)fac::FB * n_pad;
      s.c = part, s.ldc = fac::FB, s.c_batch = (int64_t)fac::kMaxSplits * fac::FB * fac::FB, s.c_split = fac::FB * fac::FB;
      s.alpha = 1.0, s.m_tiles = fac::FB / fac::Cfg::BM, s.n_tiles = 1, s.k = n1, s.kmode = fac::K_FULL, s.splits = splits;
      // with splits == 1 the kernel takes the whole K range and writes slice 0
      SE_TRY(launch_gemm_nt(s, n_s, sc));
    }
    // L22 = chol(K22 - L21 L21^T), W22 = L22^-1 -> diagonal blocks of W and W^T
    SE_LAUNCH(fac::chol_inv_block_kernel, n_s, fac::kCholThreads, fac::kCholSmemBytes, sc, dev_k + (int64_t)n1 * ld + n1, ld, mat,
              n1 > 0 ? part : nullptr, splits, (int64_t)fac::FB * fac::FB, (int64_t)fac::kMaxSplits * fac::FB * fac::FB,
              dev_w + (int64_t)n1 * ld + n1, ld, mat, dev_wt + (int64_t)n1 * ld + n1, ld, mat, dev_diag + n1, n_pad, n1,
              dev_info);
    if (n1 > 0) {
      if (sc != st) SE_CUDA(cudaEventRecord(ev.ev_chol, sc));
      // T^T = W11^T L21^T
      fac::GemmNT t{};
      t.a = dev_wt, t.lda = ld, t.a_batch = mat;
      t.b = l21, t.ldb = ld, t.b_batch = (int64_t)fac::FB * n_pad;
      t.c = tt, t.ldc = fac::FB, t.c_batch = n_pad * fac::FB, t.c_split = 0;
      t.alpha = 1.0, t.m_tiles = n1 / fac::Cfg::BM, t.n_tiles = 1, t.k = n1, t.kmode = fac::K_A_UPPER, t.splits = 1;
      SE_TRY(launch_gemm_nt_sliced(t, n_s, n1, fac::FB, slices, sm_count, st));
      if (sc != st) SE_CUDA(cudaStreamWaitEvent(st, ev.ev_chol, 0));
      // W21^T = -T^T W22^T -> new columns of W^T, new rows of W
      fac::GemmNT u{};
      u.a = tt, u.lda = fac::FB, u.a_batch = n_pad * fac::FB;
      u.b = dev_w + (int64_t)n1 * ld + n1, u.ldb = ld, u.b_batch = mat;
      u.c = dev_wt + n1, u.ldc = ld, u.c_batch = mat, u.c_split = 0;
      u.ct = dev_w + (int64_t)n1 * ld, u.ldct = ld, u.ct_batch = mat;
      u.alpha = -1.0, u.m_tiles = n1 / fac::Cfg::BM, u.n_tiles = 1, u.k = fac::FB, u.kmode = fac::K_FULL, u.splits = 1;
      SE_TRY(launch_gemm_nt(u, n_s, st));
    }
  }
  return 0;
}

extern "C" int se_gp_kinv(int n_s, int64_t n_pad, const double* dev_wt, double* dev_kinv, void* stream) {
  SE_REQUIRE(n_s >= 1 && n_s <= SE_MAX_NS, "se_gp_kinv: n_s = %d outside 1..%d", n_s, SE_MAX_NS);
  SE_REQUIRE(n_pad >= fac::FB && n_pad % fac::FB == 0, "se_gp_kinv: n_pad = %lld must be a positive multiple of %d",
             (long long)n_pad, fac::FB);
  SE_REQUIRE(dev_wt && dev_kinv, "se_gp_kinv: NULL device pointer");
  fac::GemmNT g{};
  g.a = dev_wt, g.lda = n_pad, g.a_batch = n_pad * n_pad;
  g.b = dev_wt, g.ldb = n_pad, g.b_batch = n_pad * n_pad;
  g.c = dev_kinv, g.ldc = n_pad, g.c_batch = n_pad * n_pad, g.c_split = 0;
  g.alpha = 1.0, g.m_tiles = (int)(n_pad / fac::Cfg::BM), g.n_tiles = (int)(n_pad / fac::Cfg::BN), g.k = (int)n_pad;
  g.kmode = fac::K_SYM_LOWER, g.splits = 1;
  return launch_gemm_nt(g, n_s, static_cast<cudaStream_t>(stream));
}
