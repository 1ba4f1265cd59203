// inst_reach.cuh -- launch_reach_step<SE_T>, launch_safety_distance<SE_T>, launch_rollout_small<SE_T>
#include "launchers.cuh"
#include "reach_kernels.cuh"
#include "rollout_kernels.cuh"

namespace se {

template <>
int launch_reach_step<SE_T>(const ProbDev<SE_T>& pr, const StepIO<SE_T>& io, int64_t N, cudaStream_t stream) {
  if (N <= 0) return 0;
  const int grid = (int)ceil_div(N, 128);
#define SE_CASE(NS_, NU_)                                                                         \
  if (pr.ns == NS_ && pr.nu == NU_) {                                                             \
    SE_LAUNCH((reach_step_kernel<SE_T, NS_, NU_>), grid, 128, 0, stream, pr, io, N);              \
    return 0;                                                                                     \
  }
  SE_FOR_NS_NU(SE_CASE)
#undef SE_CASE
  return fail("unsupported (n_s, n_u) = (%d, %d)", pr.ns, pr.nu);
}

template <>
int launch_var_zero_scan<SE_T>(const SE_T* var, int64_t count, int32_t* flags, cudaStream_t stream) {
  if (count <= 0) return 0;
  SE_CUDA(cudaMemsetAsync(flags + 2, 0, sizeof(int32_t), stream));
  SE_LAUNCH(var_zero_scan_kernel<SE_T>, (int)ceil_div(count, 256), 256, 0, stream, var, count, flags);
  return 0;
}

template <>
int launch_safety_distance<SE_T>(const ProbDev<SE_T>& pr, const SE_T* p, const SE_T* q, int64_t N, SE_T* d,
                                 uint8_t* inside, cudaStream_t stream) {
  if (N <= 0) return 0;
  const int grid = (int)ceil_div(N, 128);
  switch (pr.ns) {
#define SE_CASE(NS_)                                                                                   \
  case NS_:                                                                                            \
    SE_LAUNCH((safety_distance_kernel<SE_T, NS_>), grid, 128, 0, stream, pr, p, q, N, d, inside);      \
    return 0;
    SE_CASE(1) SE_CASE(2) SE_CASE(3) SE_CASE(4) SE_CASE(5) SE_CASE(6)
#undef SE_CASE
  }
  return fail("unsupported n_s = %d", pr.ns);
}

template <>
int launch_transpose_w<SE_T>(const SE_T* w, int n_pad, int ns, SE_T* wt, cudaStream_t stream) {
  dim3 grid(n_pad / 32, n_pad / 32, ns), block(32, 8);
  SE_LAUNCH(transpose_w_kernel<SE_T>, grid, block, 0, stream, w, n_pad, wt);
  return 0;
}

template <>
int launch_rollout_small<SE_T>(const GpDev<SE_T>& gp, const SE_T* wt, const ProbDev<SE_T>& pr, const RolloutIO<SE_T>& io,
                               cudaStream_t stream) {
  if (io.N <= 0) return 0;
  SE_REQUIRE(gp.n_pad == kSmallN, "fused rollout kernel: n_pad = %d, wants %d", gp.n_pad, kSmallN);
  SE_REQUIRE(io.N < (1ll << 31), "fused rollout kernel: %lld candidates exceed the grid limit", (long long)io.N);
  const int grid = (int)io.N;  // one CTA per candidate
#define SE_CASE(NS_, NU_)                                                                                      \
  if (pr.ns == NS_ && pr.nu == NU_) {                                                                          \
    SE_LAUNCH((rollout_small_kernel<SE_T, NS_, NU_>), grid, 32 * kRolloutWarps, 0, stream, gp, wt, pr, io);    \
    return 0;                                                                                                  \
  }
  SE_FOR_NS_NU(SE_CASE)
#undef SE_CASE
  return fail("unsupported (n_s, n_u) = (%d, %d)", pr.ns, pr.nu);
}

}  // namespace se
