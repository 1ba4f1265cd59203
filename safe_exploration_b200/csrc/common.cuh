// common.cuh -- shared declarations of libsecem (sm_100a).  See include/secem.h for the C ABI.
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>

#include <atomic>
#include <cstdarg>
#include <cstdio>
#include <cstring>
#include <map>
#include <mutex>
#include <utility>

#include "../../include/secem.h"

#define SE_MAX_D (SE_MAX_NS + SE_MAX_NU)
#ifndef SE_KSTAR_UNROLL
#define SE_KSTAR_UNROLL 1  // 16-point steps of the kernel-row kernel unrolled together
#endif
#ifndef SE_KSTAR_TPS
#define SE_KSTAR_TPS 512  // resident threads per SM of the one-output-per-warp-group kernel-row kernel: 128 registers each (measured: C2 0.164 ms, C3 1.29 ms; 768 threads / 80 registers 0.168 / 1.40; 896 and 1024 slower still)
#endif
#define SE_HD __host__ __device__ __forceinline__

namespace se {

// ---- error plumbing ---------------------------------------------------------------------------------------------
extern thread_local char g_last_error[1024];
extern std::atomic<uint64_t> g_launches;

inline int fail(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_last_error, sizeof(g_last_error), fmt, ap);
  va_end(ap);
  return -1;
}

#define SE_CUDA(expr)                                                                                  \
  do {                                                                                                 \
    cudaError_t _e = (expr);                                                                           \
    if (_e != cudaSuccess)                                                                             \
      return ::se::fail("%s:%d: %s -> %s", __FILE__, __LINE__, #expr, cudaGetErrorString(_e));         \
  } while (0)

#define SE_TRY(expr)            \
  do {                          \
    int _r = (expr);            \
    if (_r != 0) return _r;     \
  } while (0)

#define SE_REQUIRE(cond, ...)                        \
  do {                                               \
    if (!(cond)) return ::se::fail(__VA_ARGS__);     \
  } while (0)

// every kernel launch of the library goes through this so bench.py can report `gpu_launches`
#define SE_LAUNCH(kernel, grid, block, smem, stream, ...)                                              \
  do {                                                                                                 \
    kernel<<<(grid), (block), (smem), (stream)>>>(__VA_ARGS__);                                        \
    ::se::g_launches.fetch_add(1, std::memory_order_relaxed);                                          \
    cudaError_t _e = cudaPeekAtLastError();                                                            \
    if (_e != cudaSuccess)                                                                             \
      return ::se::fail("%s:%d: launch %s -> %s", __FILE__, __LINE__, #kernel, cudaGetErrorString(_e)); \
  } while (0)

// Opt a kernel in to `bytes` of dynamic shared memory (> 48 KB needs it).  The attribute is per DEVICE, so it is
// tracked per (kernel, device): a process may drive several GPUs.
template <typename F>
inline int set_max_dyn_smem(F* func, int bytes) {
  static std::mutex mu;
  static std::map<std::pair<const void*, int>, int> have;
  int dev = 0;
  SE_CUDA(cudaGetDevice(&dev));
  std::lock_guard<std::mutex> lock(mu);
  int& cur = have[std::make_pair(reinterpret_cast<const void*>(func), dev)];
  if (bytes > cur) {
    SE_CUDA(cudaFuncSetAttribute(func, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes));
    cur = bytes;
  }
  return 0;
}

inline int64_t round_up(int64_t x, int64_t m) { return (x + m - 1) / m * m; }
inline int64_t ceil_div(int64_t x, int64_t m) { return (x + m - 1) / m; }

// ---- device-side views -------------------------------------------------------------------------------------------
// GP model in its padded device layout (DESIGN.md "Data layout in HBM")
template <typename T>
struct GpDev {
  int n;        // training points
  int n_pad;    // n rounded up to the GEMM tile (128); padded entries are zero
  int ns, nu, d;
  const T* xt;     // [d][n_pad]        training inputs, one contiguous row per input dimension
  const T* alpha;  // [ns][n_pad]       (K + noise I)^-1 y
  const T* w;      // [ns][n_pad][n_pad] L^-1, lower triangular, row i = i-th row of L^-1
  T inv_l2[SE_MAX_NS][SE_MAX_D];  // 1 / lengthscale^2
  T outputscale[SE_MAX_NS];
  T noise[SE_MAX_NS];
};

// problem constants in kernel-parameter form (converted from se_problem on the host)
template <typename T>
struct ProbDev {
  int ns, nu, m_safe, m_act, objective, all_states, m_obs;
  int clamp_nonpos;  // se_problem.clamp_nonpositive: every variance <= 0 becomes 1e-5 (and is counted)
  T k_fb[SE_MAX_NU][SE_MAX_NS];
  T a[SE_MAX_NS][SE_MAX_NS];
  T b[SE_MAX_NS][SE_MAX_NU];
  T bchol[SE_MAX_NS][SE_MAX_NS];  // lower Cholesky factor C of B = I + k_fb^T k_fb   (utils.py:175-177)
  T bmat[SE_MAX_NS][SE_MAX_NS];   // B itself (used by the closed-form n_s <= 2 path on q B)
  T l_mu[SE_MAX_NS], l_sigma[SE_MAX_NS];
  T c_safety;
  T h_mat[SE_MAX_FACES][SE_MAX_NS];
  T h_vec[SE_MAX_FACES];
  T act_a[SE_MAX_ACT_FACES][SE_MAX_NU];
  T act_b[SE_MAX_ACT_FACES];
  T act_penalty, state_penalty, objective_theta;
  T hobs_mat[SE_MAX_OBS_FACES][SE_MAX_NS];  // obstacle polytope of the intermediate ellipsoids (m_obs faces)
  T hobs_vec[SE_MAX_OBS_FACES];
};

// exp(x) in double for finite x <= 700, without the special-case branch of the library routine (which keeps every exp()
// of the kernel-row loop a serial 15-DFMA chain; without it the independent evaluations of a lane interleave).
// x below -700 is clamped (2^k stays a normal number, so the scaling is one integer add; the kernel rows are quantised
// to 2^-47 of the prior scale afterwards, so values below e^-700 are zero either way).  The clamp is an integer compare
// on the high word.  NaN / +Inf are the CALLER's business (the kernel-row kernels poison the row's mean instead).
// Polynomial: degree 11 through the Chebyshev nodes of |r| <= ln2/2, derived and checked by tools/exp_poly.py
// (evaluation error 1.4 x 2^-53).
__device__ __forceinline__ double exp_nb(double x) {
  const bool tiny = (unsigned)__double2hiint(x) >= 0xC085E000u;  // x <= -700 (or -Inf)
  const double xc = tiny ? -700.0 : x;
  const double t = fma(xc, 1.4426950408889634, 6755399441055744.0);  // 1.5 * 2^52: k = rint(x log2 e) in the low word
  const int k = __double2loint(t);
  const double kf = t - 6755399441055744.0;
  double r = fma(kf, -0.6931471805599453, xc);  // ln 2 = hi + lo
  r = fma(kf, -2.3190468138462996e-17, r);
  double p = 2.5110046444457486e-08;
  p = fma(p, r, 2.76326511328741e-07);
  p = fma(p, r, 2.7557240894691074e-06);
  p = fma(p, r, 2.4801485451264055e-05);
  p = fma(p, r, 0.00019841269890069431);
  p = fma(p, r, 0.0013888888952343797);
  p = fma(p, r, 0.008333333333319592);
  p = fma(p, r, 0.04166666666648799);
  p = fma(p, r, 0.1666666666666668);
  p = fma(p, r, 0.5000000000000019);
  p = fma(p, r, 1.0);
  p = fma(p, r, 1.0);
  return __hiloint2double(__double2hiint(p) + (k << 20), __double2loint(p));
}
// The same contract as exp_nb (finite x <= 700, clamped below -700) with a 64-entry table: exp(x) = 2^e T[j] exp(r),
// k = rint(64 x / ln 2) = 64 e + j, |r| <= ln 2 / 128, so a degree-6 Taylor polynomial is exact to 2^-64 and the serial
// DFMA chain is 6 long instead of 11 (11 FP64 instructions instead of 17).  `tab` = kExpTab copied to shared memory
// (a lane-dependent index into constant memory would serialise).  Error <= 2 ulp (table entry, polynomial, product).
__device__ const double kExpTab[64] = {
    1.0, 1.0108892860517005, 1.0218971486541166, 1.0330248790212284,
    1.0442737824274138, 1.0556451783605572, 1.0671404006768237, 1.0787607977571199,
    1.0905077326652577, 1.102382583307841, 1.1143867425958924, 1.1265216186082418,
    1.1387886347566916, 1.1511892299529827, 1.1637248587775775, 1.1763969916502812,
    1.189207115002721, 1.202156731452703, 1.215247359980469, 1.22848053610687,
    1.241857812073484, 1.255380757024691, 1.2690509571917332, 1.2828700160787783,
    1.2968395546510096, 1.3109612115247644, 1.3252366431597413, 1.339667524053303,
    1.3542555469368927, 1.3690024229745905, 1.383909881963832, 1.3989796725383112,
    1.4142135623730951, 1.42961333839197, 1.4451808069770467, 1.460917794180647,
    1.4768261459394993, 1.4929077282912648, 1.5091644275934228, 1.5255981507445384,
    1.5422108254079407, 1.559004400237837, 1.5759808451078865, 1.593142151342267,
    1.6104903319492543, 1.6280274218573478, 1.645755478153965, 1.6636765803267364,
    1.681792830507429, 1.7001063537185235, 1.718619298122478, 1.7373338352737062,
    1.7562521603732995, 1.7753764925265212, 1.7947090750031072, 1.8142521755003989,
    1.8340080864093424, 1.8539791250833855, 1.8741676341103, 1.8945759815869656,
    1.9152065613971474, 1.9360617934922943, 1.9571441241754002, 1.978456026387951};
__device__ __forceinline__ double exp_tab(double x, const double* __restrict__ tab) {
  const bool tiny = (unsigned)__double2hiint(x) >= 0xC085E000u;  // x <= -700 (or -Inf)
  const double xc = tiny ? -700.0 : x;
  const double t = fma(xc, 92.33248261689366, 6755399441055744.0);  // k = rint(64 x / ln 2) in the low word
  const int k = __double2loint(t);
  const double kf = t - 6755399441055744.0;
  double r = fma(kf, -0.010830424696249145, xc);  // ln 2 / 64 = hi + lo
  r = fma(kf, -3.623510646634843e-19, r);
  double p = 1.0 / 720.0;
  p = fma(p, r, 1.0 / 120.0);
  p = fma(p, r, 1.0 / 24.0);
  p = fma(p, r, 1.0 / 6.0);
  p = fma(p, r, 0.5);
  p = fma(p, r, 1.0);
  p = fma(p, r, 1.0);
  p *= tab[k & 63];
  return __hiloint2double(__double2hiint(p) + ((k >> 6) << 20), __double2loint(p));
}
// 2^shift exp(x), same contract: the power of two rides on the exponent add that exp_tab performs anyway, so the kernel-row
// kernel gets its fixed-point scaled value k 2^(63 - eA) without a multiply (|shift| and the result stay far inside the
// normal range: x >= -700, shift ~ 60 .. 80).
__device__ __forceinline__ double exp_tab_scaled(double x, const double* __restrict__ tab, int shift) {
  // clamp at -700 with ONE unsigned minimum on the high word: negative doubles order like their bit patterns, so
  // x <= -700 (or -Inf / NaN with the sign set) becomes -700 - (a fraction of 2^-11 from the untouched low word), and
  // non-negative x (high word below 2^31) pass through.  The clamped value only has to make the result vanish.
  const double xc = __hiloint2double((int)min((unsigned)__double2hiint(x), 0xC085E000u), __double2loint(x));
  const double t = fma(xc, 92.33248261689366, 6755399441055744.0);  // k = rint(64 x / ln 2) in the low word
  const int k = __double2loint(t);
  const double kf = t - 6755399441055744.0;
  double r = fma(kf, -0.010830424696249145, xc);  // ln 2 / 64 = hi + lo
  r = fma(kf, -3.623510646634843e-19, r);
  double p = 1.0 / 720.0;
  p = fma(p, r, 1.0 / 120.0);
  p = fma(p, r, 1.0 / 24.0);
  p = fma(p, r, 1.0 / 6.0);
  p = fma(p, r, 0.5);
  p = fma(p, r, 1.0);
  p = fma(p, r, 1.0);
  p *= tab[k & 63];
  return __hiloint2double(__double2hiint(p) + (((k >> 6) + shift) << 20), __double2loint(p));
}
__device__ __forceinline__ bool finite_f64(double x) { return (__double2hiint(x) & 0x7ff00000) != 0x7ff00000; }

template <typename T>
struct Num;
template <>
struct Num<float> {
  static constexpr se_dtype dtype = SE_F32;
  static __device__ __forceinline__ float exp_(float x) { return expf(x); }
  static __device__ __forceinline__ float sqrt_(float x) { return sqrtf(x); }
  static __device__ __forceinline__ float abs_(float x) { return fabsf(x); }
};
template <>
struct Num<double> {
  static constexpr se_dtype dtype = SE_F64;
  static __device__ __forceinline__ double exp_(double x) { return exp(x); }
  static __device__ __forceinline__ double sqrt_(double x) { return sqrt(x); }
  static __device__ __forceinline__ double abs_(double x) { return fabs(x); }
};

}  // namespace se
