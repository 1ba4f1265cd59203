// inst_kstar.cuh -- launch_kstar<SE_T> (included by inst_kstar_f32.cu / inst_kstar_f64.cu with SE_T defined)
#include "gp_kernels.cuh"
#include "launchers.cuh"

namespace se {

template <>
int launch_kstar<SE_T>(const GpDev<SE_T>& gp, const SE_T* p, int64_t p_stride, const SE_T* u, int64_t u_stride,
                       int64_t row0, int nrows, int rows_pad, SE_T* kst, SE_T* mean, SE_T* jac,
                       cudaStream_t stream) {
  constexpr int R = 2;
  const int warps = rows_pad / R;
  const int grid = (int)ceil_div(warps, 8);
  const bool want_jac = jac != nullptr;
#define SE_CASE(NS_, NU_)                                                                                       \
  if (gp.ns == NS_ && gp.nu == NU_) {                                                                           \
    if (want_jac)                                                                                               \
      SE_LAUNCH((gp_kstar_kernel<SE_T, NS_, NS_ + NU_, true, R>), grid, 256, 0, stream, gp, p, p_stride, u,     \
                u_stride, row0, nrows, rows_pad, kst, mean, jac);                                               \
    else                                                                                                        \
      SE_LAUNCH((gp_kstar_kernel<SE_T, NS_, NS_ + NU_, false, R>), grid, 256, 0, stream, gp, p, p_stride, u,    \
                u_stride, row0, nrows, rows_pad, kst, mean, jac);                                               \
    return 0;                                                                                                   \
  }
  SE_FOR_NS_NU(SE_CASE)
#undef SE_CASE
  return fail("unsupported (n_s, n_u) = (%d, %d)", gp.ns, gp.nu);
}

}  // namespace se
