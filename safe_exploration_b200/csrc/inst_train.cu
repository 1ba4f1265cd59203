// inst_train.cu -- C ABI of the model-refresh kernels (covariance assembly, marginal-log-likelihood gradient)
#include "train_kernels.cuh"

namespace se {

static int make_hyp(int n_s, int d, const double* ls, const double* os, const double* nz, TrainHyp* h) {
  SE_REQUIRE(n_s >= 1 && n_s <= SE_MAX_NS && d >= 1 && d <= SE_MAX_D, "n_s = %d / d = %d outside 1..%d / 1..%d", n_s,
             d, SE_MAX_NS, SE_MAX_D);
  SE_REQUIRE(ls && os && nz, "hyper-parameter pointers must not be NULL");
  memset(h, 0, sizeof(*h));
  for (int o = 0; o < n_s; ++o) {
    for (int c = 0; c < d; ++c) {
      const double l = ls[o * d + c];
      SE_REQUIRE(l > 0.0, "lengthscale[%d][%d] = %g is not positive", o, c, l);
      h->neg_half_inv_l2[o][c] = -0.5 / (l * l);
      h->inv_l3[o][c] = 1.0 / (l * l * l);
      h->lengthscale[o][c] = l;
    }
    h->outputscale[o] = os[o];
    h->noise[o] = nz[o];
  }
  return 0;
}

template <int D>
static int launch_kernel_matrix(const double* x, int n, int n_s, const TrainHyp& h, double* k, cudaStream_t st) {
  const int tiles = (int)ceil_div(n, kTrainTile);
  SE_LAUNCH((gp_kernel_matrix_kernel<D>), dim3(tiles, tiles, n_s), kTrainThreads, 0, st, x, n, h, k);
  return 0;
}

template <int D>
static int launch_mll_grad(const double* x, const double* alpha, const double* kinv, int n, int n_s, const TrainHyp& h,
                           double* part, int tri, cudaStream_t st) {
  SE_LAUNCH((gp_mll_grad_tiles_kernel<D>), dim3(tri, n_s), kTrainThreads, 0, st, x, alpha, kinv, n, h, part);
  return 0;
}

#define SE_FOR_D(X) X(1) X(2) X(3) X(4) X(5) X(6) X(7) X(8) X(9)

}  // namespace se

using namespace se;

extern "C" int se_gp_kernel_matrix(int n_s, int d, int64_t n, const double* dev_x, const double* host_lengthscale,
                                   const double* host_outputscale, const double* host_noise, double* dev_k,
                                   void* stream) {
  TrainHyp h;
  SE_TRY(make_hyp(n_s, d, host_lengthscale, host_outputscale, host_noise, &h));
  SE_REQUIRE(n >= 0 && n <= 65535 * (int64_t)kTrainTile, "se_gp_kernel_matrix: n = %lld out of range", (long long)n);
  if (n == 0) return 0;
  SE_REQUIRE(dev_x && dev_k, "se_gp_kernel_matrix: NULL device pointer");
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  switch (d) {
#define X(D) \
  case D:    \
    return launch_kernel_matrix<D>(dev_x, (int)n, n_s, h, dev_k, st);
    SE_FOR_D(X)
#undef X
  }
  return fail("se_gp_kernel_matrix: unsupported d = %d", d);
}

extern "C" int64_t se_gp_mll_grad_workspace(int n_s, int d, int64_t n) {
  const int64_t tiles = ceil_div(n, kTrainTile);
  return (int64_t)n_s * (tiles * (tiles + 1) / 2) * (d + 2) * (int64_t)sizeof(double);
}

extern "C" int se_gp_mll_grad(int n_s, int d, int64_t n, const double* dev_x, const double* dev_alpha,
                              const double* dev_kinv, const double* host_lengthscale, const double* host_outputscale,
                              const double* host_noise, double* dev_workspace, double* dev_grad, void* stream) {
  TrainHyp h;
  SE_TRY(make_hyp(n_s, d, host_lengthscale, host_outputscale, host_noise, &h));
  SE_REQUIRE(n >= 1 && n <= 46340 * (int64_t)kTrainTile / 2, "se_gp_mll_grad: n = %lld out of range", (long long)n);
  SE_REQUIRE(dev_x && dev_alpha && dev_kinv && dev_workspace && dev_grad, "se_gp_mll_grad: NULL device pointer");
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const int tiles = (int)ceil_div(n, kTrainTile);
  const int tri = tiles * (tiles + 1) / 2;
  switch (d) {
#define X(D)                                                                                          \
  case D:                                                                                             \
    SE_TRY(launch_mll_grad<D>(dev_x, dev_alpha, dev_kinv, (int)n, n_s, h, dev_workspace, tri, st)); \
    break;
    SE_FOR_D(X)
#undef X
    default:
      return fail("se_gp_mll_grad: unsupported d = %d", d);
  }
  SE_LAUNCH(gp_mll_grad_reduce_kernel, n_s, 256, 0, st, dev_workspace, tri, d, h, dev_grad);
  return 0;
}

// ---------------------------------------------------------------------------------------------------------------
// se_gp_mll_grad_static: the same gradient with the per-tile partial sums in library-owned static device memory, for
// callers that may not allocate (CUDA graph capture, serving loops).  No counterpart in the reference.
// ---------------------------------------------------------------------------------------------------------------

// The scratch array.  File scope, so every GPU that loads libsecem.so holds these 4 MiB until the process exits.
// In HOST code the name `g_mll_scratch` is only a stand-in for the symbol, never a device address: the address comes
// from cudaGetSymbolAddress, once per device, with that device current (mll_scratch_ptr below).
constexpr int64_t kMllScratchDoubles = SE_MLL_GRAD_STATIC_MAX_BYTES / (int64_t)sizeof(double);
static __device__ __align__(256) double g_mll_scratch[kMllScratchDoubles];

namespace se {

constexpr int kMaxDevices = 64;

// Device address of g_mll_scratch on the CURRENT device (resolved on first use per device, then cached).
static int mll_scratch_ptr(double** out) {
  static std::mutex mu;
  static double* ptrs[kMaxDevices] = {};
  int dev = 0;
  SE_CUDA(cudaGetDevice(&dev));
  SE_REQUIRE(dev >= 0 && dev < kMaxDevices, "se_gp_mll_grad_static: device ordinal %d outside 0..%d", dev,
             kMaxDevices - 1);
  std::lock_guard<std::mutex> lock(mu);
  if (!ptrs[dev]) {
    double* p = nullptr;
    const cudaError_t e = cudaGetSymbolAddress(reinterpret_cast<void**>(&p), g_mll_scratch);
    if (e != cudaSuccess || !p)
      return fail("se_gp_mll_grad_static: cudaGetSymbolAddress(g_mll_scratch) on device %d -> %s", dev,
                  cudaGetErrorString(e));
    ptrs[dev] = p;
  }
  *out = ptrs[dev];
  return 0;
}

}  // namespace se

extern "C" int64_t se_gp_mll_grad_static_max_bytes(void) { return SE_MLL_GRAD_STATIC_MAX_BYTES; }

extern "C" int se_gp_mll_grad_static(int n_s, int d, int64_t n, const double* dev_x, const double* dev_alpha,
                                     const double* dev_kinv, const double* host_lengthscale,
                                     const double* host_outputscale, const double* host_noise, double* dev_grad,
                                     void* stream) {
  SE_REQUIRE(n_s >= 1 && n_s <= SE_MAX_NS && d >= 1 && d <= SE_MAX_D, "n_s = %d / d = %d outside 1..%d / 1..%d", n_s,
             d, SE_MAX_NS, SE_MAX_D);
  SE_REQUIRE(n >= 1 && n <= 46340 * (int64_t)kTrainTile / 2, "se_gp_mll_grad_static: n = %lld out of range",
             (long long)n);
  const int64_t need = se_gp_mll_grad_workspace(n_s, d, n);
  SE_REQUIRE(need <= SE_MLL_GRAD_STATIC_MAX_BYTES,
             "se_gp_mll_grad_static: n_s = %d, d = %d, n = %lld needs %lld bytes of per-tile partial sums, the static "
             "scratch array holds %lld (SE_MLL_GRAD_STATIC_MAX_BYTES); use se_gp_mll_grad with a caller workspace",
             n_s, d, (long long)n, (long long)need, (long long)SE_MLL_GRAD_STATIC_MAX_BYTES);
  double* scratch = nullptr;
  SE_TRY(mll_scratch_ptr(&scratch));
  // Every slot the reduction reads is written by the tile kernel of the same call (all o, tiles, d + 2 sums), so the
  // previous call's contents never need clearing.  The two kernels below are the only work: no allocation anywhere.
  return se_gp_mll_grad(n_s, d, n, dev_x, dev_alpha, dev_kinv, host_lengthscale, host_outputscale, host_noise, scratch,
                        dev_grad, stream);
}
