// secem.cu -- C ABI of libsecem.so (include/secem.h): handle management and kernel orchestration.
#include <dlfcn.h>

#include <cmath>
#include <cstdlib>
#include <mutex>
#include <unordered_map>
#include <vector>

#include "cem_kernels.cuh"
#include "gp_kernels.cuh"
#include "launchers.cuh"
#include "reach_kernels.cuh"
#include "rollout_kernels.cuh"

namespace se {
thread_local char g_last_error[1024] = "";
std::atomic<uint64_t> g_launches{0};
}  // namespace se

using namespace se;

// -------------------------------------------------------------------------------------------------------------------
// handles
// -------------------------------------------------------------------------------------------------------------------
constexpr int64_t kSplitRows = 4096;  // below this a kernel-row launch splits the training set over CTAs

struct se_gp_s {
  uint64_t generation = 0;  // unique per se_gp_create (never reused, unlike the handle's address)
  se_dtype dtype = SE_F64;
  int ns = 0, nu = 0, d = 0, n = 0, n_pad = 0;
  size_t esz = 8;
  void* xt = nullptr;
  void* alpha = nullptr;
  void* w = nullptr;
  double inv_l2[SE_MAX_NS][SE_MAX_D];
  double outputscale[SE_MAX_NS];
  double noise[SE_MAX_NS];
  // workspace: materialised kernel rows K*[ns][rows_chunk][n_pad] + partial sums + posterior buffers
  uint64_t ws_limit = 6ull << 30;
  void* kst = nullptr;
  size_t kst_bytes = 0;
  void* partial = nullptr;
  size_t partial_bytes = 0;
  void* kpart = nullptr;  // per-segment mean / Jacobian sums of a split kernel-row launch (small batches, tcgen05 path)
  size_t kpart_bytes = 0;
  void* post_mean = nullptr;  // [cap, ns]
  void* post_var = nullptr;
  void* post_jac = nullptr;  // [cap, ns, d]
  int64_t post_cap = 0;
  int* counter = nullptr;
  se_gemm_impl impl = SE_GEMM_AUTO;
  bool impl_explicit = false;  // se_gp_set_gemm_impl named an engine: the solve then keeps the four-kernel step
  int sm_count = 148;
  // tcgen05 digit-sliced variance GEMM: digit image of L^-1, per-row scales (built on first use)
  int tc_digits_req = 0;  // 0 = chosen from the prior / noise ratio (tc_auto_digits)
  int tc_digits = 0;      // resolved value (tc_prepare)
  int tc_ready_digits = 0;
  int8_t* b_img = nullptr;
  int* exp_ab = nullptr;  // [ns] exponents of the output scales, then [ns][n_pad] row exponents of L^-1
  double* scale_b = nullptr;
  // fused small-model rollout (n_pad == 128): L^-T, built on first use
  void* wt = nullptr;
  // second stream + events of the two-part pipeline in se_cem_solve (created on first use)
  cudaStream_t side = nullptr;
  cudaEvent_t ev_fork = nullptr, ev_join = nullptr, ev_k[2] = {nullptr, nullptr}, ev_g[2] = {nullptr, nullptr};
  // timing
  bool timing = false;
  std::vector<cudaEvent_t> ev_gemm, ev_kstar;  // (start, stop) pairs
};

// Live GP handles and their generation.  A CEM handle keeps a raw pointer to its GP; the allocator readily hands the
// address of a destroyed se_gp_s to the next se_gp_create (update_model destroys and re-creates the model), so a
// pointer comparison cannot tell "same model" from "new model at the old address".  se_cem_solve looks the pointer up
// here and compares generations instead, and the generation is part of the CUDA-graph key.
static std::mutex g_gp_mutex;
static std::unordered_map<const se_gp_s*, uint64_t> g_live_gp;
static uint64_t g_gp_generation = 0;

static uint64_t gp_live_generation(const se_gp_s* g) {  // 0 = not a live handle
  std::lock_guard<std::mutex> lock(g_gp_mutex);
  auto it = g_live_gp.find(g);
  return it == g_live_gp.end() ? 0 : it->second;
}

template <typename T>
static GpDev<T> gp_dev(const se_gp_s* g) {
  GpDev<T> v;
  v.n = g->n, v.n_pad = g->n_pad, v.ns = g->ns, v.nu = g->nu, v.d = g->d;
  v.xt = static_cast<const T*>(g->xt);
  v.alpha = static_cast<const T*>(g->alpha);
  v.w = static_cast<const T*>(g->w);
  for (int o = 0; o < SE_MAX_NS; ++o) {
    for (int dd = 0; dd < SE_MAX_D; ++dd) v.inv_l2[o][dd] = (T)g->inv_l2[o][dd];
    v.outputscale[o] = (T)g->outputscale[o];
    v.noise[o] = (T)g->noise[o];
  }
  return v;
}

static int grow(void** ptr, size_t* have, size_t want) {
  if (*have >= want) return 0;
  if (*ptr) SE_CUDA(cudaFree(*ptr));
  *ptr = nullptr;
  *have = 0;
  SE_CUDA(cudaMalloc(ptr, want));
  *have = want;
  return 0;
}

extern "C" const char* se_last_error(void) { return g_last_error; }
extern "C" int se_version(void) { return 100; }
extern "C" uint64_t se_launch_count(void) { return g_launches.load(); }
extern "C" void se_launch_count_reset(void) { g_launches.store(0); }
extern "C" int se_measure_peak(int kind, double* flops_out) {
  SE_REQUIRE(kind >= 0 && kind <= 10 && flops_out, "se_measure_peak: bad arguments");
  if (kind == 7) return measure_peak_i8(0, 64, 2, flops_out);    // N = 64 shape, two issuing warps (burst)
  if (kind == 8) return measure_peak_i8(0, 256, 2, flops_out);   // N = 256, two issuing warps (burst)
  if (kind == 9) return measure_peak_i8(0, 128, 1, flops_out);   // N = 128 shape (burst)
  if (kind == 10) return measure_peak_i8(0, 128, 2, flops_out);  // N = 128 shape, two issuing warps (burst)
  if (kind >= 3) return measure_peak_i8(kind == 4 || kind == 6, kind >= 5 ? 64 : 256, 1, flops_out);
  return measure_peak(kind, flops_out);
}

// SE_GEMM_AUTO: float64 models of 256 .. 16384 training points run the variance GEMM on the tcgen05 tensor cores
// (2.9x the FP64 DMMA path at n = 2000); smaller ones are launch-bound and stay on the exact DMMA kernel; float32
// models run FFMA.  SE_GEMM_AUTO_IMPL=mma|simt|tc overrides (benchmarks).
static se_gemm_impl resolve_impl(const se_gp_s* g, se_gemm_impl impl) {
  if (impl != SE_GEMM_AUTO) return impl;
  if (const char* e = getenv("SE_GEMM_AUTO_IMPL")) {
    if (!strcmp(e, "mma")) return SE_GEMM_MMA;
    if (!strcmp(e, "simt")) return SE_GEMM_SIMT;
    if (!strcmp(e, "tc") && g->dtype == SE_F64) return SE_GEMM_TC;
  }
  if (g->dtype == SE_F64 && g->n_pad >= 256 && g->n_pad <= 16384) return SE_GEMM_TC;
  return g->dtype == SE_F64 ? SE_GEMM_MMA : SE_GEMM_SIMT;
}

extern "C" int se_gp_create(se_gp_t* out, int n_s, int n_u, int n, se_dtype dtype, const void* dev_x,
                            const void* dev_alpha, const void* dev_linv, const double* host_lengthscale,
                            const double* host_outputscale, const double* host_noise, void* stream) {
  SE_REQUIRE(out, "se_gp_create: out is NULL");
  SE_REQUIRE(n_s >= 1 && n_s <= SE_MAX_NS, "se_gp_create: n_s = %d outside [1, %d]", n_s, SE_MAX_NS);
  SE_REQUIRE(n_u >= 1 && n_u <= SE_MAX_NU, "se_gp_create: n_u = %d outside [1, %d]", n_u, SE_MAX_NU);
  SE_REQUIRE(n >= 1, "se_gp_create: empty training set");
  SE_REQUIRE(dtype == SE_F32 || dtype == SE_F64, "se_gp_create: bad dtype");
  SE_REQUIRE(dev_x && dev_alpha && dev_linv && host_lengthscale && host_outputscale && host_noise,
             "se_gp_create: NULL argument");
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  se_gp_s* g = new se_gp_s();
  g->dtype = dtype;
  g->esz = dtype == SE_F64 ? 8 : 4;
  g->ns = n_s, g->nu = n_u, g->d = n_s + n_u, g->n = n;
  g->n_pad = (int)round_up(n, kTile);
  memset(g->inv_l2, 0, sizeof(g->inv_l2));
  memset(g->outputscale, 0, sizeof(g->outputscale));
  memset(g->noise, 0, sizeof(g->noise));
  for (int o = 0; o < n_s; ++o) {
    for (int dd = 0; dd < g->d; ++dd) {
      const double l = host_lengthscale[o * g->d + dd];
      if (!(l > 0.0)) {
        delete g;
        return fail("se_gp_create: lengthscale[%d][%d] = %g must be > 0", o, dd, l);
      }
      g->inv_l2[o][dd] = 1.0 / (l * l);
    }
    g->outputscale[o] = host_outputscale[o];
    g->noise[o] = host_noise[o];
  }
  int dev = 0;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&g->sm_count, cudaDevAttrMultiProcessorCount, dev);
  auto cleanup_fail = [&]() {
    se_gp_destroy(g);
    return -1;
  };
  const size_t np = (size_t)g->n_pad;
  if (cudaMalloc(&g->xt, g->d * np * g->esz) != cudaSuccess || cudaMalloc(&g->alpha, n_s * np * g->esz) != cudaSuccess ||
      cudaMalloc(&g->w, (size_t)n_s * np * np * g->esz) != cudaSuccess ||
      cudaMalloc(&g->counter, 64 * sizeof(int)) != cudaSuccess) {
    fail("se_gp_create: cudaMalloc failed (n_pad = %d): %s", g->n_pad, cudaGetErrorString(cudaGetLastError()));
    return cleanup_fail();
  }
  int rc;
  if (dtype == SE_F64)
    rc = launch_pack_model<double>((const double*)dev_x, (const double*)dev_alpha, (const double*)dev_linv, n, g->d,
                                   n_s, g->n_pad, (double*)g->xt, (double*)g->alpha, (double*)g->w, st);
  else
    rc = launch_pack_model<float>((const float*)dev_x, (const float*)dev_alpha, (const float*)dev_linv, n, g->d, n_s,
                                  g->n_pad, (float*)g->xt, (float*)g->alpha, (float*)g->w, st);
  if (rc != 0) return cleanup_fail();
  g->impl = resolve_impl(g, SE_GEMM_AUTO);
  {
    std::lock_guard<std::mutex> lock(g_gp_mutex);
    g->generation = ++g_gp_generation;
    g_live_gp[g] = g->generation;
  }
  *out = g;
  return 0;
}

extern "C" int se_gp_destroy(se_gp_t g) {
  if (!g) return 0;
  {
    std::lock_guard<std::mutex> lock(g_gp_mutex);
    g_live_gp.erase(g);
  }
  for (void* p : {g->xt, g->alpha, g->w, g->kst, g->partial, g->kpart, g->post_mean, g->post_var, g->post_jac, (void*)g->counter,
                  (void*)g->b_img, (void*)g->exp_ab, (void*)g->scale_b, g->wt})
    if (p) cudaFree(p);
  for (auto e : g->ev_gemm) cudaEventDestroy(e);
  for (auto e : g->ev_kstar) cudaEventDestroy(e);
  for (cudaEvent_t e : {g->ev_fork, g->ev_join, g->ev_k[0], g->ev_k[1], g->ev_g[0], g->ev_g[1]})
    if (e) cudaEventDestroy(e);
  if (g->side) cudaStreamDestroy(g->side);
  delete g;
  return 0;
}

extern "C" int se_gp_set_workspace_limit(se_gp_t g, uint64_t bytes) {
  SE_REQUIRE(g, "NULL handle");
  g->ws_limit = bytes;
  return 0;
}
extern "C" int se_gp_set_gemm_impl(se_gp_t g, se_gemm_impl impl) {
  SE_REQUIRE(g, "NULL handle");
  SE_REQUIRE(impl != SE_GEMM_TC || g->dtype == SE_F64, "SE_GEMM_TC evaluates float64 models");
  g->impl = resolve_impl(g, impl);
  g->impl_explicit = impl != SE_GEMM_AUTO;
  return 0;
}
extern "C" int se_gp_set_tc_digits(se_gp_t g, int digits) {
  SE_REQUIRE(g, "NULL handle");
  SE_REQUIRE(digits == 0 || (digits >= 3 && digits <= 8), "se_gp_set_tc_digits: %d is neither 0 (auto) nor in [3, 8]",
             digits);
  g->tc_digits_req = digits;
  return 0;
}
extern "C" int se_gp_get_gemm_info(se_gp_t g, int* impl_out, int* tc_digits_out) {
  SE_REQUIRE(g, "NULL handle");
  if (impl_out) *impl_out = (int)g->impl;
  if (tc_digits_out) *tc_digits_out = g->tc_digits;
  return 0;
}

// Digit planes for a target relative variance error of 1e-8.  The variance s + noise - ||L^-1 k||^2 cancels down to
// ~noise, so an error eps relative to the prior scale s shows up as eps * s / noise; measured on the bench workloads
// (tools/tc_check.py) the truncation after S planes is eps ~ 480 * 2^(-8 S) ... i.e. 1.7e-9 relative at S = 6 and
// s / noise = 10^3.
static int tc_auto_digits(const se_gp_s* g) {
  double ratio = 1.0;
  for (int o = 0; o < g->ns; ++o)
    if (g->noise[o] > 0.0) ratio = std::fmax(ratio, g->outputscale[o] / g->noise[o]);
  for (int S = 6; S < 8; ++S)
    if (480.0 * ratio * std::ldexp(1.0, -8 * S) <= 1e-8) return S;
  return 8;
}

// digit image of L^-1 for the tcgen05 path; rebuilt when the number of digit planes changes
static int tc_prepare(se_gp_s* g, cudaStream_t st) {
  int want = g->tc_digits_req ? g->tc_digits_req : tc_auto_digits(g);
  if (want == 8 && g->n_pad > 8192 && !g->tc_digits_req) want = 7;
  if (g->tc_ready_digits == want) return 0;
  g->tc_digits = want;
  SE_REQUIRE(g->dtype == SE_F64, "the tcgen05 variance GEMM evaluates float64 models");
  SE_REQUIRE(g->n_pad <= (g->tc_digits == 8 ? 8192 : 16384),
             "the tcgen05 variance GEMM holds int32 level sums: n_train must be <= 16384 (8192 with 8 digit planes)");
  SE_CUDA(cudaStreamSynchronize(st));  // a previous image may still be in use
  if (g->b_img) SE_CUDA(cudaFree(g->b_img));
  g->b_img = nullptr;
  const size_t np = (size_t)g->n_pad;
  SE_CUDA(cudaMalloc(&g->b_img, (size_t)g->ns * np * np * g->tc_digits));
  if (!g->exp_ab) SE_CUDA(cudaMalloc(&g->exp_ab, (size_t)(g->ns + g->ns * np) * sizeof(int)));
  if (!g->scale_b) SE_CUDA(cudaMalloc(&g->scale_b, (size_t)g->ns * np * sizeof(double)));
  SE_TRY(launch_oz_pack_model((const double*)g->w, g->n_pad, g->ns, g->outputscale, g->tc_digits, g->exp_ab,
                              g->exp_ab + g->ns, g->scale_b, g->b_img, st));
  g->tc_ready_digits = g->tc_digits;
  return 0;
}

extern "C" int se_debug_tc_probe(const int8_t* dev_a_img, const int8_t* dev_b_img, int nk, int digits, int lbo, int sbo,
                                 int32_t* dev_out, void* stream) {
  SE_REQUIRE(dev_a_img && dev_b_img && dev_out && nk >= 1, "se_debug_tc_probe: bad arguments");
  return launch_oz_probe(dev_a_img, dev_b_img, nk, digits, lbo, sbo, dev_out, static_cast<cudaStream_t>(stream));
}

extern "C" int se_debug_eig_max(const double* host_m, int n, double* host_out) {
  SE_REQUIRE(host_m && host_out && n >= 1 && n <= SE_MAX_NS, "se_debug_eig_max: bad arguments");
  double a[SE_MAX_NS][SE_MAX_NS] = {};
  for (int i = 0; i < n; ++i)
    for (int j = 0; j < n; ++j) a[i][j] = host_m[i * n + j];
  *host_out = max_real_eig_general<double, SE_MAX_NS>(a, n);
  return 0;
}

extern "C" int se_debug_exp(const double* dev_x, int64_t n, double* dev_out, void* stream) {
  SE_REQUIRE(n >= 0 && (n == 0 || (dev_x && dev_out)), "se_debug_exp: bad arguments");
  return launch_debug_exp(dev_x, n, dev_out, 0, static_cast<cudaStream_t>(stream));
}
extern "C" int se_debug_exp_table(const double* dev_x, int64_t n, double* dev_out, void* stream) {
  SE_REQUIRE(n >= 0 && (n == 0 || (dev_x && dev_out)), "se_debug_exp_table: bad arguments");
  return launch_debug_exp(dev_x, n, dev_out, 1, static_cast<cudaStream_t>(stream));
}

extern "C" int se_gp_timing_enable(se_gp_t g, int enable) {
  SE_REQUIRE(g, "NULL handle");
  g->timing = enable != 0;
  for (auto e : g->ev_gemm) cudaEventDestroy(e);
  for (auto e : g->ev_kstar) cudaEventDestroy(e);
  g->ev_gemm.clear();
  g->ev_kstar.clear();
  return 0;
}
extern "C" int se_gp_timing_read(se_gp_t g, double* gemm_ms, int64_t* gemm_launches, double* kstar_ms,
                                 int64_t* kstar_launches, void* stream) {
  SE_REQUIRE(g, "NULL handle");
  SE_CUDA(cudaStreamSynchronize(static_cast<cudaStream_t>(stream)));
  auto total = [](std::vector<cudaEvent_t>& ev, double* ms, int64_t* cnt) {
    double s = 0;
    for (size_t i = 0; i + 1 < ev.size(); i += 2) {
      float t = 0.f;
      cudaEventElapsedTime(&t, ev[i], ev[i + 1]);
      s += t;
    }
    if (ms) *ms = s;
    if (cnt) *cnt = (int64_t)ev.size() / 2;
    for (auto e : ev) cudaEventDestroy(e);
    ev.clear();
  };
  total(g->ev_gemm, gemm_ms, gemm_launches);
  total(g->ev_kstar, kstar_ms, kstar_launches);
  return 0;
}

// rows of K* per chunk so that ns * rows * n_pad * esz <= ws_limit (multiple of the row tile)
static int64_t chunk_rows(const se_gp_s* g, int64_t N) {
  const int64_t n_rows_pad = round_up(N, kTile);
  int64_t by_limit = (int64_t)(g->ws_limit / ((uint64_t)g->ns * g->n_pad * g->esz));
  by_limit = by_limit / kTile * kTile;
  if (by_limit < kTile) by_limit = kTile;
  return n_rows_pad < by_limit ? n_rows_pad : by_limit;
}

static size_t kpart_set_bytes(const se_gp_s* g, int64_t split_rows) {
  return (size_t)(g->n_pad / 128) * split_rows * g->ns * (1 + g->d) * sizeof(double);
}
// Ensure workspace for a batch of N candidates.  Allocates (cudaMalloc) only when the batch grows, so a caller that
// wants to capture a CUDA graph warms the handle up with one eager call first.
static int gp_reserve(se_gp_s* g, int64_t N) {
  const int64_t rows = chunk_rows(g, N);
  SE_TRY(grow(&g->kst, &g->kst_bytes, (size_t)g->ns * rows * g->n_pad * g->esz));
  SE_TRY(grow(&g->partial, &g->partial_bytes, (size_t)g->ns * (g->n_pad / 32) * rows * g->esz));
  if (g->impl == SE_GEMM_TC) {
    const int64_t split_rows = rows < kSplitRows ? rows : kSplitRows;
    // two sets for small batches: the fused ellipsoid step reads one step's segment sums while the kernel rows of the
    // next are written; one set for the whole chunk when large batches cut the training set to balance the tail
    const size_t two_small = 2 * kpart_set_bytes(g, split_rows), one_large = kpart_set_bytes(g, rows);
    SE_TRY(grow(&g->kpart, &g->kpart_bytes, two_small > one_large ? two_small : one_large));
  }
  return 0;
}
// Resident CTAs per SM of the kernel-row kernel of the tensor-core path (its launch bounds, ozaki_kernels.cuh): one
// thread per cell carrying all outputs (n_s = 1) -> 64-thread CTAs, 8 (n_s <= 3) or 6 per SM; one output per warp
// group (n_s >= 2, SE_KSTAR_OPT != 0) -> 64 n_s threads, SE_KSTAR_TPS (d <= 5) or 512 threads per SM.
static int kstar_ctas_per_sm(const se_gp_s* g) {
  static const int opt_env = getenv("SE_KSTAR_OPT") ? atoi(getenv("SE_KSTAR_OPT")) : 1;
  if (g->ns < 2 || opt_env != 1) return g->ns <= 3 ? 8 : 6;
  const int per_sm = (g->d <= 5 ? SE_KSTAR_TPS : 512) / (64 * g->ns);
  return per_sm > 0 ? per_sm : 1;
}
// How a kernel-row launch of rows_pad candidates cuts the training set over blockIdx.y (runs of 128-point segments);
// returns segments per CTA, 0 = one CTA walks all of it.  The segment sums are added in a fixed order by the consumer,
// so the result does not depend on the cut.  Two reasons to cut:
//  - fewer than kSplitRows candidates do not fill the machine with one CTA per 16 rows;
//  - more than one wave of CTAs leaves a tail: C2's 1024 CTAs on 148 x 6 = 888 slots run as one full wave and one that
//    is 15 % full (ncu: achieved occupancy 28.8 % of a theoretical 37.5 %).  Shorter CTAs shrink the tail to a
//    fraction of a wave; the cut with the lowest modelled time ceil(CTAs / slots) x (length + fixed cost) is taken.
static int kstar_segments_per_cta(const se_gp_s* g, int rows_pad, int64_t cap_rows) {
  static const int env = getenv("SE_KSTAR_SPLIT") ? atoi(getenv("SE_KSTAR_SPLIT")) : 1;
  static const int tail_env = getenv("SE_KSTAR_TAIL") ? atoi(getenv("SE_KSTAR_TAIL")) : 1;  // > 1: force that many ways
  const int segs = g->n_pad / 128;
  if (!env || g->impl != SE_GEMM_TC || segs < 2 || rows_pad > cap_rows) return 0;
  const int row_ctas = rows_pad / 16;
  if (rows_pad < kSplitRows) {
    int ways = (2 * g->sm_count + row_ctas - 1) / row_ctas;  // ~2 CTAs per SM in total
    if (ways > segs) ways = segs;
    if (ways < 2) return 0;
    const int spc = (segs + ways - 1) / ways;
    return spc < segs ? spc : 0;
  }
  if (!tail_env) return 0;
  const int64_t slots = (int64_t)g->sm_count * kstar_ctas_per_sm(g);
  auto cost = [&](int spc) {  // in units of one full-length CTA; 0.5 % of it per CTA for the prologue
    const int ny = (segs + spc - 1) / spc;
    const int64_t waves = ((int64_t)row_ctas * ny + slots - 1) / slots;
    return (double)waves * ((double)spc / segs + 0.005);
  };
  int best = segs;
  if (tail_env > 1) {
    best = (segs + tail_env - 1) / tail_env;
  } else {
    for (int spc = segs - 1; spc >= 1; --spc)
      if (cost(spc) < 0.97 * cost(best)) best = spc;  // a finer cut has to pay for its extra segment traffic
  }
  return best < segs ? best : 0;
}
static int gp_reserve_posterior(se_gp_s* g, int64_t N) {
  if (g->post_cap >= N) return 0;
  for (void** p : {&g->post_mean, &g->post_var, &g->post_jac})
    if (*p) {
      cudaFree(*p);
      *p = nullptr;
    }
  g->post_cap = 0;
  SE_CUDA(cudaMalloc(&g->post_mean, (size_t)N * g->ns * g->esz));
  SE_CUDA(cudaMalloc(&g->post_var, (size_t)N * g->ns * g->esz));
  SE_CUDA(cudaMalloc(&g->post_jac, (size_t)N * g->ns * g->d * g->esz));
  g->post_cap = N;
  return 0;
}

static int record_event(std::vector<cudaEvent_t>& ev, cudaStream_t st) {
  cudaEvent_t e;
  SE_CUDA(cudaEventCreate(&e));
  SE_CUDA(cudaEventRecord(e, st));
  ev.push_back(e);
  return 0;
}

// ---- GP posterior in three stages, for rows [row0, row0 + nrows) of a candidate batch -------------------------------
// The stages of one part use the workspace slice `slot` (kernel rows / digit image and partial sums), so two parts can
// be in flight on two streams (cem_solve_t pipelines the population in halves).
struct PostPart {
  int64_t row0;
  int nrows, rows_pad, nb;
  void* kst;      // this part's kernel rows (plain) or digit image (tcgen05 path)
  void* partial;  // this part's partial sums
  void* kpart;    // this part's per-segment mean / Jacobian sums (split kernel-row launch)
  int seg_per_cta, nseg;  // 0 / 0: the kernel-row kernel wrote mean and Jacobian itself
};

template <typename T>
static PostPart post_part(se_gp_s* g, int64_t row0, int nrows, int64_t slot_row0) {
  PostPart pp;
  pp.row0 = row0, pp.nrows = nrows, pp.rows_pad = (int)round_up(nrows, kTile), pp.nb = 0;
  // slot_row0: first (padded) row of this part inside the workspace
  pp.kst = (char*)g->kst + (size_t)g->ns * slot_row0 * g->n_pad * g->esz;
  pp.partial = (char*)g->partial + (size_t)g->ns * (g->n_pad / 32) * slot_row0 * g->esz;
  // rows of segment sums the workspace holds beyond this part's slot (small batches: half of it, see gp_reserve)
  const int64_t set_rows = g->kpart ? (int64_t)(g->kpart_bytes / kpart_set_bytes(g, 1)) : 0;
  const int64_t cap_rows = pp.rows_pad < kSplitRows ? (set_rows / 2 < kSplitRows ? set_rows / 2 : kSplitRows) - slot_row0
                                                    : set_rows - slot_row0;
  pp.seg_per_cta = kstar_segments_per_cta(g, pp.rows_pad, cap_rows);
  pp.nseg = pp.seg_per_cta > 0 ? g->n_pad / 128 : 0;
  pp.kpart = g->kpart ? (char*)g->kpart + (size_t)(g->n_pad / 128) * slot_row0 * g->ns * (1 + g->d) * sizeof(double)
                      : nullptr;
  return pp;
}

template <typename T>
static int post_kstar(se_gp_s* g, const GpDev<T>& gd, const PostPart& pp, const T* p, int64_t p_stride, const T* u,
                      int64_t u_stride, T* mean, T* jac, cudaStream_t st, const ProbDev<double>* fuse_pr = nullptr,
                      const StepIO<double>* fuse_io = nullptr) {
  const bool tc = g->impl == SE_GEMM_TC;
  if (g->timing) SE_TRY(record_event(g->ev_kstar, st));
  if constexpr (sizeof(T) == 8) {
    if (tc)  // kernel rows as int8 digit planes (<= 8 bytes per element, like the float64 rows they replace)
      SE_TRY(launch_kstar_slices(gd, p, p_stride, u, u_stride, pp.row0, pp.nrows, pp.rows_pad, g->exp_ab,
                                 g->tc_digits, (int8_t*)pp.kst, mean, jac, pp.seg_per_cta, (double*)pp.kpart, st,
                                 fuse_pr, fuse_io));
  }
  if (!tc)
    SE_TRY(launch_kstar<T>(gd, p, p_stride, u, u_stride, pp.row0, pp.nrows, pp.rows_pad, (T*)pp.kst, mean, jac, st));
  if (g->timing) SE_TRY(record_event(g->ev_kstar, st));
  return 0;
}

template <typename T>
static int post_gemm(se_gp_s* g, const GpDev<T>& gd, PostPart& pp, cudaStream_t st) {
  const bool tc = g->impl == SE_GEMM_TC;
  if (g->timing) SE_TRY(record_event(g->ev_gemm, st));
  if constexpr (sizeof(T) == 8) {
    if (tc)
      SE_TRY(launch_trisumsq_ozaki((const int8_t*)pp.kst, g->b_img, g->scale_b, g->n, g->n_pad, pp.rows_pad, g->ns,
                                   g->tc_digits, (double*)pp.partial, g->sm_count, &pp.nb, st));
  }
  if (!tc)
    SE_TRY(launch_trisumsq<T>(g->impl, (const T*)pp.kst, gd.w, g->n_pad, pp.rows_pad, g->ns, (T*)pp.partial,
                              g->counter, g->sm_count, &pp.nb, st));
  if (g->timing) SE_TRY(record_event(g->ev_gemm, st));
  return 0;
}

template <typename T>
static int post_var(se_gp_s* g, const GpDev<T>& gd, const PostPart& pp, T* var, T* mean, T* jac, cudaStream_t st) {
  return launch_var<T>(gd, (const T*)pp.partial, pp.nb, pp.rows_pad, pp.row0, pp.nrows, var, (const T*)pp.kpart,
                       pp.nseg, mean, jac, st);
}

// GP posterior for N candidates whose inputs are z_c = [p[c*p_stride : +ns] | u[c*u_stride : +nu]]
template <typename T>
static int gp_posterior(se_gp_s* g, const T* p, int64_t p_stride, const T* u, int64_t u_stride, int64_t N, T* mean,
                        T* var, T* jac, cudaStream_t st) {
  if (N <= 0) return 0;
  SE_TRY(gp_reserve(g, N));
  if (g->impl == SE_GEMM_TC) SE_TRY(tc_prepare(g, st));
  const GpDev<T> gd = gp_dev<T>(g);
  const int64_t rows_chunk = chunk_rows(g, N);
  for (int64_t row0 = 0; row0 < N; row0 += rows_chunk) {
    const int nrows = (int)((N - row0) < rows_chunk ? (N - row0) : rows_chunk);
    PostPart pp = post_part<T>(g, row0, nrows, 0);
    SE_TRY(post_kstar<T>(g, gd, pp, p, p_stride, u, u_stride, mean, jac, st));
    SE_TRY(post_gemm<T>(g, gd, pp, st));
    SE_TRY(post_var<T>(g, gd, pp, var, mean, jac, st));
  }
  return 0;
}

extern "C" int se_gp_predict(se_gp_t g, const void* dev_z, int64_t N, void* dev_mean, void* dev_var, void* dev_jac,
                             void* stream) {
  SE_REQUIRE(g && dev_z && dev_mean && dev_var, "se_gp_predict: NULL argument");
  SE_REQUIRE(N >= 0, "se_gp_predict: N < 0");
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (g->dtype == SE_F64) {
    const double* z = (const double*)dev_z;
    return gp_posterior<double>(g, z, g->d, z + g->ns, g->d, N, (double*)dev_mean, (double*)dev_var,
                                (double*)dev_jac, st);
  }
  const float* z = (const float*)dev_z;
  return gp_posterior<float>(g, z, g->d, z + g->ns, g->d, N, (float*)dev_mean, (float*)dev_var, (float*)dev_jac, st);
}

// -------------------------------------------------------------------------------------------------------------------
// problem constants
// -------------------------------------------------------------------------------------------------------------------
template <typename T>
static int make_prob(const se_problem* pr, const double* k_fb_override, ProbDev<T>* out) {
  SE_REQUIRE(pr, "se_problem is NULL");
  const int ns = pr->n_s, nu = pr->n_u;
  SE_REQUIRE(ns >= 1 && ns <= SE_MAX_NS && nu >= 1 && nu <= SE_MAX_NU, "se_problem: bad (n_s, n_u) = (%d, %d)", ns, nu);
  SE_REQUIRE(pr->m_safe >= 0 && pr->m_safe <= SE_MAX_FACES, "se_problem: m_safe = %d outside [0, %d]", pr->m_safe,
             SE_MAX_FACES);
  SE_REQUIRE(pr->m_act >= 0 && pr->m_act <= SE_MAX_ACT_FACES, "se_problem: m_act = %d outside [0, %d]", pr->m_act,
             SE_MAX_ACT_FACES);
  SE_REQUIRE(pr->l_mu && pr->l_sigma, "se_problem: l_mu / l_sigma is NULL");
  SE_REQUIRE(pr->m_safe == 0 || (pr->h_mat && pr->h_vec), "se_problem: h_mat / h_vec is NULL");
  SE_REQUIRE(pr->m_act == 0 || (pr->act_a && pr->act_b), "se_problem: act_a / act_b is NULL");
  SE_REQUIRE(pr->m_obs >= 0 && pr->m_obs <= SE_MAX_OBS_FACES, "se_problem: m_obs = %d outside [0, %d]", pr->m_obs,
             SE_MAX_OBS_FACES);
  SE_REQUIRE(pr->m_obs == 0 || (pr->h_mat_obs && pr->h_vec_obs), "se_problem: h_mat_obs / h_vec_obs is NULL");
  ProbDev<T>& o = *out;
  memset(&o, 0, sizeof(o));
  o.ns = ns, o.nu = nu, o.m_safe = pr->m_safe, o.m_act = pr->m_act;
  o.objective = pr->objective, o.all_states = pr->all_states_constraint;
  o.clamp_nonpos = pr->clamp_nonpositive;
  const double* kf = k_fb_override ? k_fb_override : pr->k_fb;
  double kfb[SE_MAX_NU][SE_MAX_NS] = {};
  for (int k = 0; k < nu; ++k)
    for (int j = 0; j < ns; ++j) {
      kfb[k][j] = kf ? kf[k * ns + j] : 0.0;
      o.k_fb[k][j] = (T)kfb[k][j];
    }
  for (int i = 0; i < ns; ++i) {
    for (int j = 0; j < ns; ++j) o.a[i][j] = (T)(pr->a ? pr->a[i * ns + j] : (i == j ? 1.0 : 0.0));  // :60-62
    for (int k = 0; k < nu; ++k) o.b[i][k] = (T)(pr->b ? pr->b[i * nu + k] : 0.0);
    o.l_mu[i] = (T)pr->l_mu[i];
    o.l_sigma[i] = (T)pr->l_sigma[i];
  }
  // B = S S^T = I + k_fb^T k_fb (utils.py:175-177) and its lower Cholesky factor
  double bm[SE_MAX_NS][SE_MAX_NS] = {}, ch[SE_MAX_NS][SE_MAX_NS] = {};
  for (int i = 0; i < ns; ++i)
    for (int j = 0; j < ns; ++j) {
      double s = i == j ? 1.0 : 0.0;
      for (int k = 0; k < nu; ++k) s += kfb[k][i] * kfb[k][j];
      bm[i][j] = s;
    }
  for (int j = 0; j < ns; ++j) {
    double dsum = bm[j][j];
    for (int k = 0; k < j; ++k) dsum -= ch[j][k] * ch[j][k];
    ch[j][j] = std::sqrt(dsum);
    for (int i = j + 1; i < ns; ++i) {
      double s = bm[i][j];
      for (int k = 0; k < j; ++k) s -= ch[i][k] * ch[j][k];
      ch[i][j] = s / ch[j][j];
    }
  }
  for (int i = 0; i < ns; ++i)
    for (int j = 0; j < ns; ++j) {
      o.bmat[i][j] = (T)bm[i][j];
      o.bchol[i][j] = (T)ch[i][j];
    }
  o.c_safety = (T)pr->c_safety;
  for (int f = 0; f < pr->m_safe; ++f) {
    for (int j = 0; j < ns; ++j) o.h_mat[f][j] = (T)pr->h_mat[f * ns + j];
    o.h_vec[f] = (T)pr->h_vec[f];
  }
  o.m_obs = pr->m_obs;
  for (int f = 0; f < pr->m_obs; ++f) {
    for (int j = 0; j < ns; ++j) o.hobs_mat[f][j] = (T)pr->h_mat_obs[f * ns + j];
    o.hobs_vec[f] = (T)pr->h_vec_obs[f];
  }
  for (int f = 0; f < pr->m_act; ++f) {
    for (int k = 0; k < nu; ++k) o.act_a[f][k] = (T)pr->act_a[f * nu + k];
    o.act_b[f] = (T)pr->act_b[f];
  }
  o.act_penalty = (T)pr->act_penalty;
  o.state_penalty = (T)pr->state_penalty;
  o.objective_theta = (T)pr->objective_theta;
  return 0;
}

// -------------------------------------------------------------------------------------------------------------------
// reachability
// -------------------------------------------------------------------------------------------------------------------
template <typename T>
static int reach_onestep_t(se_gp_s* g, const se_problem* prob, const T* p, const T* q, const T* u, int64_t N, T* p1,
                           T* q1, T* sigma, int32_t* flags, cudaStream_t st) {
  ProbDev<T> pr;
  SE_TRY(make_prob<T>(prob, nullptr, &pr));
  SE_REQUIRE(pr.ns == g->ns && pr.nu == g->nu, "se_reach_onestep: problem / GP dimension mismatch");
  SE_REQUIRE(q == nullptr || prob->k_fb != nullptr, "se_reach_onestep: k_fb is required for ellipsoid states");
  SE_TRY(gp_reserve_posterior(g, N));
  T* mean = (T*)g->post_mean;
  T* var = (T*)g->post_var;
  T* jac = q ? (T*)g->post_jac : nullptr;
  SE_TRY(gp_posterior<T>(g, p, g->ns, u, g->nu, N, mean, var, jac, st));
  StepIO<T> io;
  memset(&io, 0, sizeof(io));
  io.p_in = p, io.q_in = q, io.p_out = p1, io.q_out = q1;
  io.u = u, io.u_stride = g->nu;
  io.mean = mean, io.var = var, io.jac = jac;
  io.sigma_out = sigma;
  io.point = q ? 0 : 1;
  io.first_step = 1, io.last_step = 1, io.T_len = 1;
  io.flags = flags;
  if (flags) {  // batch-wide rule of _fix_zeros_nans: does this batch hold an exact-zero variance?
    SE_TRY(launch_var_zero_scan<T>(var, N * g->ns, flags, st));
    io.batch_zero_scan = 1;
  }
  return launch_reach_step<T>(pr, io, N, st);
}

template <typename T>
static int reach_step_t(const se_problem* prob, const T* p, const T* q, const T* u, const T* mean, const T* var,
                        const T* jac, int64_t N, T* p1, T* q1, T* sigma, int32_t* flags, cudaStream_t st) {
  ProbDev<T> pr;
  SE_TRY(make_prob<T>(prob, nullptr, &pr));
  SE_REQUIRE(q == nullptr || (prob->k_fb != nullptr && jac != nullptr),
             "se_reach_step: k_fb and the mean Jacobian are required for ellipsoid states");
  StepIO<T> io;
  memset(&io, 0, sizeof(io));
  io.p_in = p, io.q_in = q, io.p_out = p1, io.q_out = q1;
  io.u = u, io.u_stride = pr.nu;
  io.mean = mean, io.var = var, io.jac = q ? jac : nullptr;
  io.sigma_out = sigma;
  io.point = q ? 0 : 1;
  io.first_step = 1, io.last_step = 1, io.T_len = 1;
  io.flags = flags;
  if (flags) {
    SE_TRY(launch_var_zero_scan<T>(var, N * pr.ns, flags, st));
    io.batch_zero_scan = 1;
  }
  return launch_reach_step<T>(pr, io, N, st);
}

extern "C" int se_reach_step(se_dtype dtype, const se_problem* prob, const void* dev_p, const void* dev_q,
                             const void* dev_u, const void* dev_mean, const void* dev_var, const void* dev_jac,
                             int64_t N, void* dev_p1, void* dev_q1, void* dev_sigma, int32_t* dev_flags,
                             void* stream) {
  SE_REQUIRE(prob && dev_p && dev_u && dev_mean && dev_var && dev_p1 && dev_q1, "se_reach_step: NULL argument");
  SE_REQUIRE(dtype == SE_F32 || dtype == SE_F64, "se_reach_step: bad dtype");
  SE_REQUIRE(N >= 0, "se_reach_step: N < 0");
  if (N == 0) return 0;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (dtype == SE_F64)
    return reach_step_t<double>(prob, (const double*)dev_p, (const double*)dev_q, (const double*)dev_u,
                                (const double*)dev_mean, (const double*)dev_var, (const double*)dev_jac, N,
                                (double*)dev_p1, (double*)dev_q1, (double*)dev_sigma, dev_flags, st);
  return reach_step_t<float>(prob, (const float*)dev_p, (const float*)dev_q, (const float*)dev_u,
                             (const float*)dev_mean, (const float*)dev_var, (const float*)dev_jac, N, (float*)dev_p1,
                             (float*)dev_q1, (float*)dev_sigma, dev_flags, st);
}

extern "C" int se_reach_onestep(se_gp_t g, const se_problem* prob, const void* dev_p, const void* dev_q,
                                const void* dev_u, int64_t N, void* dev_p1, void* dev_q1, void* dev_sigma,
                                int32_t* dev_flags, void* stream) {
  SE_REQUIRE(g && prob && dev_p && dev_u && dev_p1 && dev_q1, "se_reach_onestep: NULL argument");
  SE_REQUIRE(N >= 0, "se_reach_onestep: N < 0");
  if (N == 0) return 0;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (g->dtype == SE_F64)
    return reach_onestep_t<double>(g, prob, (const double*)dev_p, (const double*)dev_q, (const double*)dev_u, N,
                                   (double*)dev_p1, (double*)dev_q1, (double*)dev_sigma, dev_flags, st);
  return reach_onestep_t<float>(g, prob, (const float*)dev_p, (const float*)dev_q, (const float*)dev_u, N,
                                (float*)dev_p1, (float*)dev_q1, (float*)dev_sigma, dev_flags, st);
}

template <typename T>
static int reach_multistep_t(se_gp_s* g, const se_problem* prob, const T* p0, const T* q0, const T* u, int64_t N,
                             int T_len, const double* k_fb_steps, T* p_all, T* q_all, T* sigma_all, T* pT, T* qT,
                             int32_t* flags, cudaStream_t st) {
  SE_REQUIRE(T_len >= 1 && T_len <= SE_MAX_T, "se_reach_multistep: T = %d outside [1, %d]", T_len, SE_MAX_T);
  SE_REQUIRE(prob->k_fb || k_fb_steps || (T_len == 1 && !q0), "se_reach_multistep: k_fb is required");
  SE_TRY(gp_reserve_posterior(g, N));
  T* mean = (T*)g->post_mean;
  T* var = (T*)g->post_var;
  for (int t = 0; t < T_len; ++t) {
    ProbDev<T> pr;
    // step 0 uses prob->k_fb (k_fb_init of gp_reachability.py:204-205); step i >= 1 uses k_fb[i-1] (:213-216)
    const double* kf = (k_fb_steps && t >= 1) ? k_fb_steps + (size_t)(t - 1) * g->nu * g->ns : nullptr;
    SE_TRY(make_prob<T>(prob, kf, &pr));
    SE_REQUIRE(pr.ns == g->ns && pr.nu == g->nu, "se_reach_multistep: problem / GP dimension mismatch");
    const bool point = (t == 0 && q0 == nullptr);
    const T* p_in = t == 0 ? p0 : pT;
    const T* q_in = t == 0 ? q0 : qT;
    T* jac = point ? nullptr : (T*)g->post_jac;
    SE_TRY(gp_posterior<T>(g, p_in, g->ns, u + (int64_t)t * g->nu, (int64_t)T_len * g->nu, N, mean, var, jac, st));
    StepIO<T> io;
    memset(&io, 0, sizeof(io));
    io.p_in = p_in, io.q_in = q_in, io.p_out = pT, io.q_out = qT;
    io.u = u + (int64_t)t * g->nu, io.u_stride = (int64_t)T_len * g->nu;
    io.mean = mean, io.var = var, io.jac = jac;
    io.point = point ? 1 : 0;
    io.first_step = t == 0, io.last_step = t == T_len - 1, io.T_len = T_len, io.t_index = t;
    io.p_all = p_all, io.q_all = q_all, io.sigma_all = sigma_all;
    io.flags = flags;
    if (flags) {
      SE_TRY(launch_var_zero_scan<T>(var, N * g->ns, flags, st));
      io.batch_zero_scan = 1;
    }
    SE_TRY(launch_reach_step<T>(pr, io, N, st));
  }
  return 0;
}

extern "C" int se_reach_multistep(se_gp_t g, const se_problem* prob, const void* dev_p0, const void* dev_q0,
                                  const void* dev_u, int64_t N, int T, const double* host_k_fb_steps, void* dev_p_all,
                                  void* dev_q_all, void* dev_sigma_all, void* dev_pT, void* dev_qT, int32_t* dev_flags,
                                  void* stream) {
  SE_REQUIRE(g && prob && dev_p0 && dev_u && dev_pT && dev_qT, "se_reach_multistep: NULL argument");
  SE_REQUIRE(N >= 0, "se_reach_multistep: N < 0");
  if (N == 0) return 0;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (g->dtype == SE_F64)
    return reach_multistep_t<double>(g, prob, (const double*)dev_p0, (const double*)dev_q0, (const double*)dev_u, N,
                                     T, host_k_fb_steps, (double*)dev_p_all, (double*)dev_q_all,
                                     (double*)dev_sigma_all, (double*)dev_pT, (double*)dev_qT, dev_flags, st);
  return reach_multistep_t<float>(g, prob, (const float*)dev_p0, (const float*)dev_q0, (const float*)dev_u, N, T,
                                  host_k_fb_steps, (float*)dev_p_all, (float*)dev_q_all, (float*)dev_sigma_all,
                                  (float*)dev_pT, (float*)dev_qT, dev_flags, st);
}

extern "C" int se_safety_distance(se_dtype dtype, int n_s, int m, const double* host_h_mat, const double* host_h_vec,
                                  double c_safety, const void* dev_p, const void* dev_q, int64_t N, void* dev_d,
                                  uint8_t* dev_inside, void* stream) {
  SE_REQUIRE(host_h_mat && host_h_vec && dev_p && dev_q, "se_safety_distance: NULL argument");
  SE_REQUIRE(N >= 0, "se_safety_distance: N < 0");
  if (N == 0) return 0;
  se_problem pr;
  memset(&pr, 0, sizeof(pr));
  const double zeros[SE_MAX_NS] = {};
  pr.n_s = n_s, pr.n_u = 1, pr.l_mu = zeros, pr.l_sigma = zeros;
  pr.c_safety = c_safety, pr.m_safe = m, pr.h_mat = host_h_mat, pr.h_vec = host_h_vec;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (dtype == SE_F64) {
    ProbDev<double> pd;
    SE_TRY(make_prob<double>(&pr, nullptr, &pd));
    return launch_safety_distance<double>(pd, (const double*)dev_p, (const double*)dev_q, N, (double*)dev_d,
                                          dev_inside, st);
  }
  ProbDev<float> pf;
  SE_TRY(make_prob<float>(&pr, nullptr, &pf));
  return launch_safety_distance<float>(pf, (const float*)dev_p, (const float*)dev_q, N, (float*)dev_d, dev_inside, st);
}

// -------------------------------------------------------------------------------------------------------------------
// NCCL, resolved at run time from the already-loaded libnccl.so.2 (torch ships it); single-GPU use needs no NCCL
// -------------------------------------------------------------------------------------------------------------------
namespace {
struct NcclId {
  char bytes[128];
};
struct NcclApi {
  int (*GetUniqueId)(NcclId*) = nullptr;
  int (*CommInitRank)(void**, int, NcclId, int) = nullptr;
  int (*CommDestroy)(void*) = nullptr;
  int (*AllGather)(const void*, void*, size_t, int, void*, cudaStream_t) = nullptr;
  const char* (*GetErrorString)(int) = nullptr;
  bool ok = false;
};
NcclApi g_nccl;

int nccl_load() {
  if (g_nccl.ok) return 0;
  void* h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
  if (!h) h = dlopen("libnccl.so", RTLD_NOW | RTLD_GLOBAL);
  SE_REQUIRE(h, "NCCL not found: %s", dlerror());
  g_nccl.GetUniqueId = reinterpret_cast<decltype(g_nccl.GetUniqueId)>(dlsym(h, "ncclGetUniqueId"));
  g_nccl.CommInitRank = reinterpret_cast<decltype(g_nccl.CommInitRank)>(dlsym(h, "ncclCommInitRank"));
  g_nccl.CommDestroy = reinterpret_cast<decltype(g_nccl.CommDestroy)>(dlsym(h, "ncclCommDestroy"));
  g_nccl.AllGather = reinterpret_cast<decltype(g_nccl.AllGather)>(dlsym(h, "ncclAllGather"));
  g_nccl.GetErrorString = reinterpret_cast<decltype(g_nccl.GetErrorString)>(dlsym(h, "ncclGetErrorString"));
  SE_REQUIRE(g_nccl.GetUniqueId && g_nccl.CommInitRank && g_nccl.CommDestroy && g_nccl.AllGather,
             "NCCL symbols missing");
  g_nccl.ok = true;
  return 0;
}
}  // namespace

extern "C" int se_nccl_unique_id(char* out128) {
  SE_TRY(nccl_load());
  NcclId id;
  const int r = g_nccl.GetUniqueId(&id);
  SE_REQUIRE(r == 0, "ncclGetUniqueId failed: %d", r);
  memcpy(out128, id.bytes, 128);
  return 0;
}
extern "C" int se_nccl_comm_init(void** comm, int world_size, const char* id128, int rank) {
  SE_TRY(nccl_load());
  NcclId id;
  memcpy(id.bytes, id128, 128);
  const int r = g_nccl.CommInitRank(comm, world_size, id, rank);
  SE_REQUIRE(r == 0, "ncclCommInitRank failed: %s", g_nccl.GetErrorString ? g_nccl.GetErrorString(r) : "?");
  return 0;
}
extern "C" int se_nccl_comm_destroy(void* comm) {
  if (!comm) return 0;
  SE_TRY(nccl_load());
  g_nccl.CommDestroy(comm);
  return 0;
}

// -------------------------------------------------------------------------------------------------------------------
// CEM
// -------------------------------------------------------------------------------------------------------------------
struct GraphKey {
  const void* ptr[32];
  int64_t val[8];
};
struct GraphEntry {
  GraphKey key;
  cudaGraphExec_t exec = nullptr;  // null: buffer set seen once, not captured yet
  uint64_t kernels = 0;            // kernel launches inside the graph (for se_launch_count)
  uint64_t stamp = 0;
};

struct se_cem_s {
  se_gp_s* gp = nullptr;
  uint64_t gp_generation = 0;  // generation of *gp when this handle was created
  se_cem_config cfg{};
  ProbDev<float> pf;
  ProbDev<double> pd;
  void* comm = nullptr;
  int rank = 0, world = 1;
  // device buffers
  void *actions = nullptr, *p = nullptr, *q = nullptr, *objective = nullptr, *constraint = nullptr;
  void *p2 = nullptr, *q2 = nullptr;  // second ellipsoid buffers: the fused step of small batches alternates them
  void *pack_local = nullptr, *pack_all = nullptr;
  double *mean = nullptr, *stdv = nullptr;  // [I + 1][M][T * nu]
  void* act_out = nullptr;                   // scratch outputs when the caller passes NULL
  int32_t *found = nullptr, *best = nullptr;
  // CUDA-graph replay of the enqueue phase (se_cem_solve)
  uint64_t* seed_dev = nullptr;
  cudaStream_t gstream = nullptr;
  std::vector<GraphEntry> graphs;
  uint64_t graph_clock = 0;
  bool graph_failed = false;
};

extern "C" int se_cem_destroy(se_cem_t c) {
  if (!c) return 0;
  for (void* p : {c->actions, c->p, c->q, c->p2, c->q2, c->objective, c->constraint, c->pack_local, c->pack_all, (void*)c->mean,
                  (void*)c->stdv, c->act_out, (void*)c->found, (void*)c->best})
    if (p) cudaFree(p);
  if (c->seed_dev) cudaFree(c->seed_dev);
  for (auto& e : c->graphs)
    if (e.exec) cudaGraphExecDestroy(e.exec);
  if (c->gstream) cudaStreamDestroy(c->gstream);
  delete c;
  return 0;
}

extern "C" int se_cem_create(se_cem_t* out, se_gp_t gp, const se_problem* prob, const se_cem_config* cfg) {
  SE_REQUIRE(out && gp && prob && cfg, "se_cem_create: NULL argument");
  SE_REQUIRE(cfg->num_instances >= 1, "se_cem_create: num_instances < 1");
  SE_REQUIRE(cfg->time_horizon >= 1 && cfg->time_horizon <= SE_MAX_T, "se_cem_create: time_horizon = %d outside [1, %d]",
             cfg->time_horizon, SE_MAX_T);
  SE_REQUIRE(cfg->num_iterations >= 1, "se_cem_create: num_iterations < 1");
  SE_REQUIRE(cfg->num_elites >= 1 && cfg->num_elites <= kMaxElites && cfg->num_elites <= cfg->num_rollouts,
             "se_cem_create: num_elites = %d must be in [1, min(%d, num_rollouts = %d)]", cfg->num_elites, kMaxElites,
             cfg->num_rollouts);
  SE_REQUIRE(cfg->local_rollouts >= 1 && cfg->first_rollout >= 0 &&
                 cfg->first_rollout + cfg->local_rollouts <= cfg->num_rollouts,
             "se_cem_create: bad shard [%d, +%d) of %d", cfg->first_rollout, cfg->local_rollouts, cfg->num_rollouts);
  SE_REQUIRE(cfg->first_instance >= 0 &&
                 ((int64_t)cfg->first_instance + cfg->num_instances) * cfg->num_rollouts < (1ll << 31),
             "se_cem_create: population too large");
  SE_REQUIRE(prob->n_s == gp->ns && prob->n_u == gp->nu, "se_cem_create: problem / GP dimension mismatch");
  SE_REQUIRE(prob->k_fb && prob->a && prob->b, "se_cem_create: k_fb, a, b are required");
  SE_REQUIRE(gp_live_generation(gp) != 0, "se_cem_create: the GP handle is not live (destroyed?)");
  se_cem_s* c = new se_cem_s();
  c->gp = gp;
  c->gp_generation = gp->generation;
  c->cfg = *cfg;
  if (make_prob<float>(prob, nullptr, &c->pf) != 0 || make_prob<double>(prob, nullptr, &c->pd) != 0) {
    delete c;
    return -1;
  }
  const size_t esz = gp->esz;
  const size_t M = cfg->num_instances, Nl = cfg->local_rollouts, T = cfg->time_horizon, ns = gp->ns, nu = gp->nu;
  const size_t cand = M * Nl, tnu = T * nu;
  bool ok = true;
  auto alloc = [&](void** p, size_t bytes) { ok = ok && cudaMalloc(p, bytes) == cudaSuccess; };
  alloc(&c->actions, cand * tnu * esz);
  alloc(&c->p, cand * ns * esz);
  alloc(&c->q, cand * ns * ns * esz);
  alloc(&c->p2, cand * ns * esz);
  alloc(&c->q2, cand * ns * ns * esz);
  alloc(&c->objective, cand * esz);
  alloc(&c->constraint, cand * esz);
  alloc(&c->pack_local, 2 * cand * esz);
  alloc((void**)&c->mean, (cfg->num_iterations + 1) * M * tnu * sizeof(double));
  alloc((void**)&c->stdv, (cfg->num_iterations + 1) * M * tnu * sizeof(double));
  alloc(&c->act_out, M * tnu * esz);
  alloc((void**)&c->found, M * sizeof(int32_t));
  alloc((void**)&c->best, M * sizeof(int32_t));
  if (!ok) {
    fail("se_cem_create: cudaMalloc failed: %s", cudaGetErrorString(cudaGetLastError()));
    se_cem_destroy(c);
    return -1;
  }
  if (gp_reserve(gp, (int64_t)cand) != 0 || gp_reserve_posterior(gp, (int64_t)cand) != 0) {
    se_cem_destroy(c);
    return -1;
  }
  *out = c;
  return 0;
}

extern "C" int se_cem_set_comm(se_cem_t c, void* nccl_comm, int rank, int world_size) {
  SE_REQUIRE(c, "NULL handle");
  SE_REQUIRE(world_size >= 1 && rank >= 0 && rank < world_size, "se_cem_set_comm: bad rank %d / %d", rank, world_size);
  if (world_size > 1) {
    SE_REQUIRE(nccl_comm, "se_cem_set_comm: NULL communicator");
    SE_REQUIRE(c->cfg.local_rollouts * world_size == c->cfg.num_rollouts &&
                   c->cfg.first_rollout == rank * c->cfg.local_rollouts,
               "se_cem_set_comm: population must be split evenly: rank r owns [r N/W, (r+1) N/W)");
    SE_TRY(nccl_load());
    if (c->pack_all) cudaFree(c->pack_all);
    const size_t cand = (size_t)c->cfg.num_instances * c->cfg.local_rollouts;
    SE_CUDA(cudaMalloc(&c->pack_all, (size_t)world_size * 2 * cand * c->gp->esz));
  }
  c->comm = nccl_comm, c->rank = rank, c->world = world_size;
  return 0;
}

template <typename T>
static int cem_solve_t(se_cem_s* c, const ProbDev<T>& pr, const T* p0, const T* noise, uint64_t seed,
                       const uint64_t* seed_dev, const se_cem_outputs* out, cudaStream_t st, int phase) {
  // phase bit 0: prepare (allocations, digit image, L^-T -- everything that may allocate or synchronise);
  // phase bit 1: enqueue the solve on `st` (pure stream work: this part is what se_cem_solve captures into a graph)
  se_gp_s* g = c->gp;
  const se_cem_config& cf = c->cfg;
  const int M = cf.num_instances, Nl = cf.local_rollouts, Ng = cf.num_rollouts, T_len = cf.time_horizon;
  const int ns = g->ns, nu = g->nu, tnu = T_len * nu, S = ns + ns * ns;
  const int64_t cand = (int64_t)M * Nl;
  SE_REQUIRE(c->world == 1 ? (Nl == Ng && cf.first_rollout == 0) : true,
             "se_cem_solve: a sharded population needs se_cem_set_comm");
  T* actions = (T*)c->actions;
  T* p = (T*)c->p;
  T* q = (T*)c->q;
  T* objective = (T*)c->objective;
  T* constraint = (T*)c->constraint;
  T* pack_local = (T*)c->pack_local;
  T* pack_all = c->world > 1 ? (T*)c->pack_all : pack_local;
  T* mean = (T*)g->post_mean;
  T* var = (T*)g->post_var;
  T* jac = (T*)g->post_jac;
  const size_t dist = (size_t)M * tnu;

  // pipeline the population in two halves when it is large enough and fits the workspace in one piece
  // (off by default: on the 1 kW B200 both the int8 GEMM and the FP64 kernel-row kernel run into the power cap, so
  // overlapping them does not shorten the solve -- measured 132 ms pipelined vs 124 ms sequential at C2)
  const int pipe_env = getenv("SE_PIPELINE") ? atoi(getenv("SE_PIPELINE")) : 0;
  const int64_t half0_rows = round_up((cand + 1) / 2, kTile);
  const int64_t two_part_rows = half0_rows + round_up(cand - half0_rows > 0 ? cand - half0_rows : 1, kTile);
  const bool pipelined = pipe_env != 0 && cand >= 4 * kTile && chunk_rows(g, two_part_rows) >= two_part_rows;
  // Models of at most 128 training points: the fused rollout kernel (SE_ROLLOUT_FUSED=0 keeps the four-kernel step).
  const int fused_env = getenv("SE_ROLLOUT_FUSED") ? atoi(getenv("SE_ROLLOUT_FUSED")) : 1;
  const bool fused = fused_env != 0 && g->n_pad == kSmallN && !g->timing && !g->impl_explicit;
  // Small batches on the tensor-core path: the ellipsoid step rides on the next kernel-row launch (SE_STEP_FUSED=0 keeps
  // the three-launch step)
  const int step_env = getenv("SE_STEP_FUSED") ? atoi(getenv("SE_STEP_FUSED")) : 1;
  const bool step_fused = step_env != 0 && !fused && !pipelined && sizeof(T) == 8 && g->impl == SE_GEMM_TC &&
                          !g->timing && cand < kSplitRows && chunk_rows(g, cand) >= cand;
  if (phase & 1) {
    if (fused && !g->wt) {
      SE_CUDA(cudaMalloc(&g->wt, (size_t)ns * g->n_pad * g->n_pad * g->esz));
      SE_TRY(launch_transpose_w<T>((const T*)g->w, g->n_pad, ns, (T*)g->wt, st));
    }
    if (!fused) SE_TRY(gp_reserve(g, pipelined ? two_part_rows : cand));
    if (!fused) SE_TRY(gp_reserve_posterior(g, cand));
    if (!fused && g->impl == SE_GEMM_TC) SE_TRY(tc_prepare(g, st));
    if (pipelined && !g->side) {
      SE_CUDA(cudaStreamCreateWithFlags(&g->side, cudaStreamNonBlocking));
      for (cudaEvent_t* e : {&g->ev_fork, &g->ev_join, &g->ev_k[0], &g->ev_k[1], &g->ev_g[0], &g->ev_g[1]})
        SE_CUDA(cudaEventCreateWithFlags(e, cudaEventDisableTiming));
    }
  }
  if (!(phase & 2)) return 0;
  // the posterior buffers may have been re-allocated by the prepare phase
  mean = (T*)g->post_mean, var = (T*)g->post_var, jac = (T*)g->post_jac;

  if (out->flags) SE_CUDA(cudaMemsetAsync(out->flags, 0, 4 * sizeof(int32_t), st));
  SE_TRY(launch_cem_init_dist(c->mean, c->stdv, (int)dist, st));

  ActionSrc<T> src;
  memset(&src, 0, sizeof(src));
  src.seed = seed, src.seed_dev = seed_dev, src.n_global = Ng, src.tnu = tnu;
  src.inst_offset = cf.first_instance;
  CostView<T> costs{pack_all, M, Nl};

  for (int it = 0; it < cf.num_iterations; ++it) {
    src.mode = noise ? 1 : 0;
    src.iteration = it;
    src.ext = noise ? noise + (size_t)it * M * Ng * tnu : nullptr;
    src.mean = c->mean + (size_t)it * dist;
    src.stdv = c->stdv + (size_t)it * dist;
    T* rec_act = out->rollout_actions ? (T*)out->rollout_actions + (size_t)it * cand * tnu : nullptr;
    T* rec_traj = out->trajectories ? (T*)out->trajectories + (size_t)it * cand * (T_len + 1) * S : nullptr;
    if (!fused) SE_TRY(launch_cem_sample<T>(src, M, cf.first_rollout, Nl, ns, p0, actions, p, rec_act, st));

    // RolloutFunction._sample_trajectory (constrained_cem_mpc.py:175-183): T sequential ellipsoid steps
    auto make_io = [&](int t, int64_t first) {
      StepIO<T> io;
      memset(&io, 0, sizeof(io));
      io.p_in = p, io.q_in = q, io.p_out = p, io.q_out = q;
      io.u = actions + (int64_t)t * nu, io.u_stride = tnu;
      io.mean = mean, io.var = var, io.jac = jac;
      io.point = t == 0 ? 1 : 0;  // get_action starts from a point state: q = None (safempc_cem.py:234-235, quirk Q9)
      io.objective = objective, io.constraint = constraint;
      io.first_step = t == 0, io.last_step = t == T_len - 1;
      io.act_all = actions, io.T_len = T_len, io.t_index = t;
      io.traj = rec_traj;
      io.flags = out->flags;
      io.first = first;
      return io;
    };
    if (fused) {
      // small model: one warp carries a candidate through all T steps (rollout_kernels.cuh), one launch per iteration
      RolloutIO<T> rio;
      memset(&rio, 0, sizeof(rio));
      rio.src = src, rio.first = cf.first_rollout, rio.n_local = Nl, rio.p0 = p0, rio.T_len = T_len;
      rio.pack = pack_local, rio.rec_obj = out->objective ? (T*)out->objective + (size_t)it * cand : nullptr;
      rio.rec_con = out->constraint ? (T*)out->constraint + (size_t)it * cand : nullptr;
      rio.rec_actions = rec_act, rio.traj = rec_traj, rio.flags = out->flags, rio.N = cand;
      SE_TRY(launch_rollout_small<T>(gp_dev<T>(g), (const T*)g->wt, pr, rio, st));
    } else if (!pipelined && step_fused) {
      // Small batches on the tensor-core path: two launches per horizon step.  The ellipsoid step of horizon step t - 1
      // runs at the head of the kernel-row kernel of step t (gp_kstar_slices8_kernel<..., FUSE>), which takes the new
      // centres from registers; ellipsoids and segment sums alternate between two buffers because every CTA that works
      // on a candidate recomputes its step from the old ones.  Same arithmetic as the three-launch step below.
      if constexpr (sizeof(T) == 8) {
        const GpDev<T> gd = gp_dev<T>(g);
        T* pb[2] = {p, (T*)c->p2};
        T* qb[2] = {q, (T*)c->q2};
        char* kp[2] = {(char*)g->kpart, (char*)g->kpart + g->kpart_bytes / 2};
        auto step_io = [&](int t, const PostPart& pp) {  // the ellipsoid step of horizon step t, fed by pp's launches
          StepIO<T> io = make_io(t, 0);
          io.p_in = pb[t & 1], io.q_in = qb[t & 1], io.p_out = pb[(t + 1) & 1], io.q_out = qb[(t + 1) & 1];
          io.partial = (const T*)pp.partial, io.nb = pp.nb, io.rows_pad = pp.rows_pad;
          io.kpart = (const T*)pp.kpart, io.nseg = pp.nseg;
          for (int o = 0; o < ns; ++o) {
            io.gp_outputscale[o] = (T)g->outputscale[o], io.gp_noise[o] = (T)g->noise[o];
            for (int dd = 0; dd < g->d; ++dd) io.gp_inv_l2[o][dd] = (T)g->inv_l2[o][dd];
          }
          return io;
        };
        PostPart prev;
        for (int t = 0; t < T_len; ++t) {
          PostPart pp = post_part<T>(g, 0, (int)cand, 0);
          if (pp.nseg > 0) pp.kpart = kp[t & 1];
          const T* u_t = actions + (int64_t)t * nu;
          if (t == 0) {
            SE_TRY(post_kstar<T>(g, gd, pp, p, ns, u_t, tnu, mean, nullptr, st));
          } else {
            const StepIO<T> io = step_io(t - 1, prev);
            SE_TRY(post_kstar<T>(g, gd, pp, nullptr, ns, u_t, tnu, mean, jac, st, &pr, &io));
          }
          SE_TRY(post_gemm<T>(g, gd, pp, st));
          prev = pp;
        }
        SE_TRY(launch_reach_step<T>(pr, step_io(T_len - 1, prev), (int)cand, st));
      }
    } else if (!pipelined) {
      // per horizon step and workspace chunk: kernel rows -> variance GEMM -> ellipsoid step (which also reduces the
      // GEMM's partial sums, so there is no separate variance launch)
      const GpDev<T> gd = gp_dev<T>(g);
      const int64_t rows_chunk = chunk_rows(g, cand);
      for (int t = 0; t < T_len; ++t)
        for (int64_t row0 = 0; row0 < cand; row0 += rows_chunk) {
          const int nrows = (int)((cand - row0) < rows_chunk ? (cand - row0) : rows_chunk);
          PostPart pp = post_part<T>(g, row0, nrows, 0);
          SE_TRY(post_kstar<T>(g, gd, pp, p, ns, actions + (int64_t)t * nu, tnu, mean, t == 0 ? nullptr : jac, st));
          SE_TRY(post_gemm<T>(g, gd, pp, st));
          StepIO<T> io = make_io(t, row0);
          io.partial = (const T*)pp.partial, io.nb = pp.nb, io.rows_pad = pp.rows_pad;
          io.kpart = (const T*)pp.kpart, io.nseg = pp.nseg;
          for (int o = 0; o < ns; ++o) {
            io.gp_outputscale[o] = (T)g->outputscale[o], io.gp_noise[o] = (T)g->noise[o];
            for (int dd = 0; dd < g->d; ++dd) io.gp_inv_l2[o][dd] = (T)g->inv_l2[o][dd];
          }
          SE_TRY(launch_reach_step<T>(pr, io, nrows, st));
        }
    } else {
      // Two-part software pipeline: the population is cut in halves; the main stream runs nothing but the variance
      // GEMMs (tensor pipe) while the side stream prepares the other half (kernel rows: FP64 pipe) and advances the
      // ellipsoids of the half whose GEMM just finished.  A half's own chain stays sequential:
      //   kstar(h, t) -> gemm(h, t) -> var, reach(h, t) -> kstar(h, t + 1)
      SE_CUDA(cudaEventRecord(g->ev_fork, st));
      SE_CUDA(cudaStreamWaitEvent(g->side, g->ev_fork, 0));
      const GpDev<T> gd = gp_dev<T>(g);
      const int64_t half0 = half0_rows;
      PostPart part[2] = {post_part<T>(g, 0, (int)half0, 0), post_part<T>(g, half0, (int)(cand - half0), half0)};
      for (int t = 0; t < T_len; ++t)
        for (int h = 0; h < 2; ++h) {
          PostPart& pp = part[h];
          SE_TRY(post_kstar<T>(g, gd, pp, p, ns, actions + (int64_t)t * nu, tnu, mean, t == 0 ? nullptr : jac, g->side));
          SE_CUDA(cudaEventRecord(g->ev_k[h], g->side));
          SE_CUDA(cudaStreamWaitEvent(st, g->ev_k[h], 0));
          SE_TRY(post_gemm<T>(g, gd, pp, st));
          SE_CUDA(cudaEventRecord(g->ev_g[h], st));
          SE_CUDA(cudaStreamWaitEvent(g->side, g->ev_g[h], 0));
          SE_TRY(post_var<T>(g, gd, pp, var, mean, t == 0 ? nullptr : jac, g->side));
          SE_TRY(launch_reach_step<T>(pr, make_io(t, pp.row0), pp.nrows, g->side));
        }
      SE_CUDA(cudaEventRecord(g->ev_join, g->side));
      SE_CUDA(cudaStreamWaitEvent(st, g->ev_join, 0));
    }

    T* rec_obj = out->objective ? (T*)out->objective + (size_t)it * cand : nullptr;
    T* rec_con = out->constraint ? (T*)out->constraint + (size_t)it * cand : nullptr;
    if (!fused) SE_TRY(launch_cem_pack_costs<T>(objective, constraint, cand, pack_local, rec_obj, rec_con, st));
    if (c->world > 1) {
      // the one collective of the path: per-candidate (objective, constraint) costs of every rank
      const int r = g_nccl.AllGather(pack_local, pack_all, (size_t)(2 * cand), sizeof(T) == 8 ? 8 : 7, c->comm, st);
      SE_REQUIRE(r == 0, "ncclAllGather failed: %s", g_nccl.GetErrorString ? g_nccl.GetErrorString(r) : "?");
    }

    SelectArgs<T> sa;
    memset(&sa, 0, sizeof(sa));
    sa.costs = costs, sa.n_global = Ng, sa.num_elites = cf.num_elites, sa.src = src;
    sa.mean_next = c->mean + (size_t)(it + 1) * dist;
    sa.std_next = c->stdv + (size_t)(it + 1) * dist;
    sa.elite_ids = out->elite_ids ? out->elite_ids + (size_t)it * M * cf.num_elites : nullptr;
    sa.mean_hist = out->means ? (T*)out->means + (size_t)it * dist : nullptr;
    sa.std_hist = out->stds ? (T*)out->stds + (size_t)it * dist : nullptr;
    SE_TRY(launch_cem_select_refit<T>(sa, M, st));
  }
  // get_actions: rollouts of the LAST iteration (sampled from the distribution of iteration I-1)
  T* act_out = out->actions ? (T*)out->actions : (T*)c->act_out;
  int32_t* found = out->found ? out->found : c->found;
  int32_t* best = out->best_id ? out->best_id : c->best;
  SE_TRY(launch_cem_pick<T>(costs, Ng, cf.pick_best, src, M, act_out, found, best, st));
  return 0;
}

static int cem_solve_phase(se_cem_s* c, const void* dev_p0, const void* dev_noise, uint64_t seed,
                           const uint64_t* seed_dev, const se_cem_outputs* out, cudaStream_t st, int phase) {
  if (c->gp->dtype == SE_F64)
    return cem_solve_t<double>(c, c->pd, (const double*)dev_p0, (const double*)dev_noise, seed, seed_dev, out, st, phase);
  return cem_solve_t<float>(c, c->pf, (const float*)dev_p0, (const float*)dev_noise, seed, seed_dev, out, st, phase);
}

// Everything a captured solve bakes in: caller pointers and the library buffers its kernels were given.
static GraphKey graph_key(const se_cem_s* c, const void* p0, const void* noise, const se_cem_outputs* out) {
  GraphKey k;
  memset(&k, 0, sizeof(k));
  const se_gp_s* g = c->gp;
  const void* ptrs[] = {p0, noise, out->actions, out->found, out->best_id, out->means, out->stds, out->elite_ids,
                        out->objective, out->constraint, out->rollout_actions, out->trajectories, out->flags,
                        g->xt, g->alpha, g->w, g->kst, g->partial, g->kpart, g->post_mean, g->post_var, g->post_jac,
                        g->b_img, g->exp_ab, g->scale_b, g->wt};
  static_assert(sizeof(ptrs) / sizeof(ptrs[0]) <= sizeof(k.ptr) / sizeof(k.ptr[0]), "GraphKey too small");
  for (size_t i = 0; i < sizeof(ptrs) / sizeof(ptrs[0]); ++i) k.ptr[i] = ptrs[i];
  k.val[0] = (int64_t)g->impl, k.val[1] = g->tc_digits, k.val[2] = (int64_t)g->ws_limit, k.val[3] = g->impl_explicit;
  k.val[4] = (getenv("SE_ROLLOUT_FUSED") ? atoi(getenv("SE_ROLLOUT_FUSED")) : 1) +
             2 * (getenv("SE_STEP_FUSED") ? atoi(getenv("SE_STEP_FUSED")) : 1);
  k.val[5] = g->n_pad;
  k.val[6] = (int64_t)g->generation;  // kernels take GpDev (n, hyper-parameters) by value: a new model is a new graph
  k.val[7] = g->n;
  return k;
}

static void graph_cache_clear(se_cem_s* c) {
  for (auto& e : c->graphs)
    if (e.exec) cudaGraphExecDestroy(e.exec);
  c->graphs.clear();
}

// One MPC solve.  Small and mid-size solves are bounded by launch gaps (tools/graph_probe.py), so the enqueue phase is
// captured into a CUDA graph the second time the same buffers come back and replayed from then on; the candidate seed
// lives in device memory so that a replay draws new candidates.  SE_CEM_GRAPH=0 disables it.
extern "C" int se_cem_solve(se_cem_t c, const void* dev_p0, const void* dev_noise, uint64_t seed,
                            const se_cem_outputs* out, void* stream) {
  SE_REQUIRE(c && dev_p0 && out, "se_cem_solve: NULL argument");
  SE_REQUIRE(gp_live_generation(c->gp) == c->gp_generation,
             "se_cem_solve: the GP handle this CEM handle was created on has been destroyed or replaced "
             "(update_model); create a new CEM handle with se_cem_create");
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const int graph_env = getenv("SE_CEM_GRAPH") ? atoi(getenv("SE_CEM_GRAPH")) : 1;
  const int pipe_env = getenv("SE_PIPELINE") ? atoi(getenv("SE_PIPELINE")) : 0;
  const bool eligible = graph_env != 0 && !c->graph_failed && c->world == 1 && !c->gp->timing && pipe_env == 0;
  if (!eligible) return cem_solve_phase(c, dev_p0, dev_noise, seed, nullptr, out, st, 3);

  SE_TRY(cem_solve_phase(c, dev_p0, dev_noise, seed, nullptr, out, st, 1));  // may allocate / synchronise
  if (!c->seed_dev) SE_CUDA(cudaMalloc(&c->seed_dev, sizeof(uint64_t)));
  SE_TRY(launch_cem_set_seed(c->seed_dev, seed, st));
  const GraphKey key = graph_key(c, dev_p0, dev_noise, out);
  for (auto& e : c->graphs)
    if (e.exec && memcmp(&e.key, &key, sizeof(key)) == 0) {
      SE_CUDA(cudaGraphLaunch(e.exec, st));
      g_launches.fetch_add(e.kernels, std::memory_order_relaxed);
      e.stamp = ++c->graph_clock;
      return 0;
    }
  // not captured yet: the first sighting of a buffer set runs eagerly, the second is captured
  bool seen = false;
  for (auto& e : c->graphs)
    if (!e.exec && memcmp(&e.key, &key, sizeof(key)) == 0) seen = true;
  if (!seen) {
    if (c->graphs.size() >= 8) {  // drop the least recently used entry
      size_t lru = 0;
      for (size_t i = 1; i < c->graphs.size(); ++i)
        if (c->graphs[i].stamp < c->graphs[lru].stamp) lru = i;
      if (c->graphs[lru].exec) cudaGraphExecDestroy(c->graphs[lru].exec);
      c->graphs.erase(c->graphs.begin() + lru);
    }
    GraphEntry e;
    e.key = key, e.stamp = ++c->graph_clock;
    c->graphs.push_back(e);
    return cem_solve_phase(c, dev_p0, dev_noise, seed, c->seed_dev, out, st, 2);
  }
  if (!c->gstream) SE_CUDA(cudaStreamCreateWithFlags(&c->gstream, cudaStreamNonBlocking));
  cudaGraph_t graph = nullptr;
  const uint64_t launches_before = g_launches.load();
  cudaError_t ce = cudaStreamBeginCapture(c->gstream, cudaStreamCaptureModeRelaxed);
  int rc = -1;
  if (ce == cudaSuccess) {
    rc = cem_solve_phase(c, dev_p0, dev_noise, seed, c->seed_dev, out, c->gstream, 2);
    ce = cudaStreamEndCapture(c->gstream, &graph);
  }
  const uint64_t kernels = g_launches.load() - launches_before;
  cudaGraphExec_t exec = nullptr;
  if (ce == cudaSuccess && rc == 0 && graph) ce = cudaGraphInstantiate(&exec, graph, 0);
  if (graph) cudaGraphDestroy(graph);
  if (ce != cudaSuccess || rc != 0 || !exec) {
    // capture is an optimisation: fall back to plain launches for the rest of this handle's life
    cudaGetLastError();
    c->graph_failed = true;
    graph_cache_clear(c);
    g_launches.store(launches_before);
    return cem_solve_phase(c, dev_p0, dev_noise, seed, nullptr, out, st, 2);
  }
  for (auto& e : c->graphs)
    if (!e.exec && memcmp(&e.key, &key, sizeof(key)) == 0) {
      e.exec = exec, e.kernels = kernels, e.stamp = ++c->graph_clock;
      break;
    }
  SE_CUDA(cudaGraphLaunch(exec, st));
  return 0;
}

extern "C" int se_cem_noise(se_dtype dtype, uint64_t seed, int iteration, int num_instances, int num_rollouts_global,
                            int first_rollout, int count, int T, int n_u, void* dev_noise_out, void* stream) {
  SE_REQUIRE(dev_noise_out && count >= 0 && T >= 1 && n_u >= 1 && num_instances >= 1, "se_cem_noise: bad arguments");
  if (count == 0) return 0;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (dtype == SE_F64)
    return launch_cem_noise<double>(seed, iteration, num_instances, num_rollouts_global, first_rollout, count, T * n_u,
                                    (double*)dev_noise_out, st);
  return launch_cem_noise<float>(seed, iteration, num_instances, num_rollouts_global, first_rollout, count, T * n_u,
                                 (float*)dev_noise_out, st);
}

template <typename T>
static int cem_pick_t(const T* objective, const T* constraint, int N, int pick_best, int32_t* best_id, cudaStream_t st) {
  T* pack = nullptr;
  T* act = nullptr;
  int32_t* found = nullptr;
  SE_CUDA(cudaMalloc(&pack, (size_t)2 * N * sizeof(T)));
  SE_CUDA(cudaMalloc(&act, sizeof(T)));
  SE_CUDA(cudaMalloc(&found, sizeof(int32_t)));
  int rc = launch_cem_pack_costs<T>(objective, constraint, N, pack, nullptr, nullptr, st);
  ActionSrc<T> src;
  memset(&src, 0, sizeof(src));
  src.mode = 2, src.n_global = N, src.tnu = 0;  // no action sequence to copy out: the caller indexes its own tensor
  if (rc == 0) rc = launch_cem_pick<T>(CostView<T>{pack, 1, N}, N, pick_best, src, 1, act, found, best_id, st);
  cudaStreamSynchronize(st);
  cudaFree(pack);
  cudaFree(act);
  cudaFree(found);
  return rc;
}

extern "C" int se_cem_pick(se_dtype dtype, const void* dev_objective, const void* dev_constraint, int N, int pick_best,
                           int32_t* dev_best_id, void* stream) {
  SE_REQUIRE(dev_objective && dev_constraint && dev_best_id && N >= 1, "se_cem_pick: bad arguments");
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (dtype == SE_F64)
    return cem_pick_t<double>((const double*)dev_objective, (const double*)dev_constraint, N, pick_best, dev_best_id, st);
  return cem_pick_t<float>((const float*)dev_objective, (const float*)dev_constraint, N, pick_best, dev_best_id, st);
}

template <typename T>
static int select_refit_t(const T* objective, const T* constraint, const T* actions, int N, int T_len, int n_u, int E,
                          int32_t* elite_ids, T* mean_out, T* std_out, cudaStream_t st) {
  // temporary device buffers (test / integration entry point, not on the solve path)
  T* pack = nullptr;
  double* dist = nullptr;
  const int tnu = T_len * n_u;
  SE_CUDA(cudaMalloc(&pack, (size_t)2 * N * sizeof(T)));
  SE_CUDA(cudaMalloc(&dist, (size_t)4 * tnu * sizeof(double)));
  int rc = launch_cem_pack_costs<T>(objective, constraint, N, pack, nullptr, nullptr, st);
  if (rc == 0) rc = launch_cem_init_dist(dist, dist + tnu, tnu, st);
  SelectArgs<T> sa;
  memset(&sa, 0, sizeof(sa));
  sa.costs = CostView<T>{pack, 1, N};
  sa.n_global = N, sa.num_elites = E;
  sa.src.mode = 2, sa.src.n_global = N, sa.src.tnu = tnu, sa.src.ext = actions;
  sa.src.mean = dist, sa.src.stdv = dist + tnu;
  sa.mean_next = dist + 2 * tnu, sa.std_next = dist + 3 * tnu;
  sa.elite_ids = elite_ids, sa.mean_hist = mean_out, sa.std_hist = std_out;
  if (rc == 0) rc = launch_cem_select_refit<T>(sa, 1, st);
  cudaStreamSynchronize(st);
  cudaFree(pack);
  cudaFree(dist);
  return rc;
}

extern "C" int se_cem_select_refit(se_dtype dtype, const void* dev_objective, const void* dev_constraint,
                                   const void* dev_actions, int N, int T, int n_u, int E, int32_t* dev_elite_ids,
                                   void* dev_mean, void* dev_std, void* stream) {
  SE_REQUIRE(dev_objective && dev_constraint && dev_actions, "se_cem_select_refit: NULL argument");
  SE_REQUIRE(E >= 1 && E <= kMaxElites && E <= N, "se_cem_select_refit: num_elites = %d must be in [1, min(%d, N = %d)]",
             E, kMaxElites, N);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (dtype == SE_F64)
    return select_refit_t<double>((const double*)dev_objective, (const double*)dev_constraint,
                                  (const double*)dev_actions, N, T, n_u, E, dev_elite_ids, (double*)dev_mean,
                                  (double*)dev_std, st);
  return select_refit_t<float>((const float*)dev_objective, (const float*)dev_constraint, (const float*)dev_actions, N,
                               T, n_u, E, dev_elite_ids, (float*)dev_mean, (float*)dev_std, st);
}
