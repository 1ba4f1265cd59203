// inst_ozaki_cluster.cu -- the cluster flavour of the tcgen05 variance GEMM (gp_trisumsq_ozaki_cluster_kernel), in its
// own translation unit so the library builds in parallel.
#include "launchers.cuh"
#include "ozaki_kernels.cuh"

namespace se {

// cluster flavour: CA x CB CTAs share kernel-row / L^-1 digit planes through multicast bulk copies
template <int S, int CA, int CB>
static int launch_oz_gemm_cluster_s(const int8_t* a_img, const int8_t* b_img, const double* scale_b, int n, int n_pad,
                                    int rows_pad, const TriOrder& od, int sm_count, double* partial,
                                    cudaStream_t stream) {
  using C = oz::Cfg<S>;
  TriOrder og = od;
  og.nb = (od.nb + CA - 1) / CA;
  og.rt_count = od.rt_count / CB;
  og.rc = od.rc / CB > 0 ? od.rc / CB : 1;
  const int64_t items = (int64_t)og.nb * og.rt_count * od.ns;
  const int clusters = (int)(items < sm_count / (CA * CB) ? items : sm_count / (CA * CB));
  SE_TRY(set_max_dyn_smem((oz::gp_trisumsq_ozaki_cluster_kernel<S, CA, CB>), C::SMEM_BYTES));
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3((unsigned)(clusters * CA * CB), 1, 1);
  cfg.blockDim = dim3(oz::THREADS, 1, 1);
  cfg.dynamicSmemBytes = C::SMEM_BYTES;
  cfg.stream = stream;
  cudaLaunchAttribute attr;
  attr.id = cudaLaunchAttributeClusterDimension;
  attr.val.clusterDim.x = CA * CB;
  attr.val.clusterDim.y = 1;
  attr.val.clusterDim.z = 1;
  cfg.attrs = &attr;
  cfg.numAttrs = 1;
  SE_CUDA(cudaLaunchKernelEx(&cfg, oz::gp_trisumsq_ozaki_cluster_kernel<S, CA, CB>, a_img, b_img, scale_b, n, n_pad,
                             rows_pad, og, od.rt_count, od.nb, partial));
  g_launches.fetch_add(1, std::memory_order_relaxed);
  return 0;
}
template <int S>
static int launch_oz_gemm_cluster(int shape, const int8_t* a_img, const int8_t* b_img, const double* scale_b, int n,
                                  int n_pad, int rows_pad, const TriOrder& od, int sm_count, double* partial,
                                  cudaStream_t stream) {
  switch (shape) {
    case 12: return launch_oz_gemm_cluster_s<S, 1, 2>(a_img, b_img, scale_b, n, n_pad, rows_pad, od, sm_count, partial, stream);
    case 21: return launch_oz_gemm_cluster_s<S, 2, 1>(a_img, b_img, scale_b, n, n_pad, rows_pad, od, sm_count, partial, stream);
  }
  return fail("tcgen05 variance GEMM: cluster shape %d not compiled (12, 21)", shape);
}

int launch_trisumsq_ozaki_cluster(int S, int shape, const int8_t* a_img, const int8_t* b_img, const double* scale_b,
                                  int n, int n_pad, int rows_pad, const TriOrder& od, int sm_count, double* partial,
                                  cudaStream_t stream) {
#define SE_OZ_CL_CASE(S_) \
  case S_:                \
    return launch_oz_gemm_cluster<S_>(shape, a_img, b_img, scale_b, n, n_pad, rows_pad, od, sm_count, partial, stream);
  switch (S) {
    SE_OZ_CL_CASE(3)
    SE_OZ_CL_CASE(4)
    SE_OZ_CL_CASE(5)
    SE_OZ_CL_CASE(6)
    SE_OZ_CL_CASE(7)
    SE_OZ_CL_CASE(8)
  }
#undef SE_OZ_CL_CASE
  return fail("tcgen05 variance GEMM: digit planes S = %d not compiled (3 .. 8)", S);
}

}  // namespace se
