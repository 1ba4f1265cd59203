// inst_kstar_opt1_s6.cu -- gp_kstar_slices8_kernel<..., OPT = 1, SFIX = 6>: the one-output-per-warp-group kernel-row
// kernel with the plane count fixed at compile time to 6, the default of float64-class models (own translation unit:
// the library builds in parallel).
#include "launchers.cuh"
#include "ozaki_kernels.cuh"

namespace se {

int launch_kstar_slices_opt1_s6(const GpDev<double>& gp, const double* p, int64_t p_stride, const double* u,
                                int64_t u_stride, int64_t row0, int nrows, int rows_pad, const int* exp_a, int8_t* a_img,
                                double* mean, double* jac, int spc, double* kpart, unsigned grid_x, unsigned grid_y,
                                cudaStream_t stream) {
  constexpr int S = 6;
  const dim3 grid(grid_x, grid_y);
#define SE_CASE(NS_, NU_)                                                                                            \
  if (NS_ >= 2 && gp.ns == NS_ && gp.nu == NU_) {                                                                    \
    if (jac)                                                                                                         \
      SE_LAUNCH((oz::gp_kstar_slices8_kernel<NS_, NS_ + NU_, true, false, (NS_ >= 2 ? 1 : NS_), 6>), grid, 64 * NS_, 0, \
                stream, gp, p, p_stride, u, u_stride, row0, nrows, rows_pad, exp_a, S, a_img, mean, jac, spc, kpart, \
                oz::FuseArgs<false>());                                                                              \
    else                                                                                                             \
      SE_LAUNCH((oz::gp_kstar_slices8_kernel<NS_, NS_ + NU_, false, false, (NS_ >= 2 ? 1 : NS_), 6>), grid, 64 * NS_,   \
                0, stream, gp, p, p_stride, u, u_stride, row0, nrows, rows_pad, exp_a, S, a_img, mean, jac, spc,     \
                kpart, oz::FuseArgs<false>());                                                                       \
    return 0;                                                                                                        \
  }
  SE_FOR_NS_NU(SE_CASE)
#undef SE_CASE
  return fail("unsupported (n_s, n_u) = (%d, %d)", gp.ns, gp.nu);
}

}  // namespace se
