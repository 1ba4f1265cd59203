// inst_ozaki.cu -- launchers of the tcgen05 (kind::i8) digit-sliced variance GEMM and of the kernels feeding it
#include <cstdlib>

#include "launchers.cuh"
#include "ozaki_kernels.cuh"

namespace se {

int oz_image_bytes_b(int n_pad, int ns, int S, size_t* out) {
  *out = (size_t)ns * n_pad * (size_t)n_pad * S;
  return 0;
}

// exp_a[o] = exponent with outputscale_o < 2^exp_a; per-row exponents / scales of L^-1; digit image of L^-1
int launch_oz_pack_model(const double* w, int n_pad, int ns, const double* host_outputscale, int S, int* exp_a,
                         int* exp_b, double* scale_b, int8_t* b_img, cudaStream_t stream) {
  SE_REQUIRE(S >= 3 && S <= oz::MAX_S, "digit planes S = %d outside [3, %d]", S, oz::MAX_S);
  int ea[SE_MAX_NS] = {};
  for (int o = 0; o < ns; ++o) {
    SE_REQUIRE(host_outputscale[o] > 0.0, "outputscale[%d] must be > 0", o);
    ea[o] = oz::digit_exponent(host_outputscale[o]);
  }
  SE_CUDA(cudaMemcpyAsync(exp_a, ea, ns * sizeof(int), cudaMemcpyHostToDevice, stream));
  const int64_t rows = (int64_t)ns * n_pad;
  SE_LAUNCH(oz::oz_row_scale_kernel, (int)ceil_div(rows * 32, 256), 256, 0, stream, w, n_pad, ns, exp_a, exp_b, scale_b);
  dim3 grid((unsigned)ceil_div(n_pad / 4, 128), (unsigned)n_pad, (unsigned)ns);
  SE_LAUNCH(oz::oz_pack_b_kernel, grid, 128, 0, stream, w, n_pad, exp_b, S, b_img);
  return 0;
}

int launch_kstar_slices(const GpDev<double>& gp, const double* p, int64_t p_stride, const double* u, int64_t u_stride,
                        int64_t row0, int nrows, int rows_pad, const int* exp_a, int S, int8_t* a_img, double* mean,
                        double* jac, int seg_per_cta, double* kpart, cudaStream_t stream, const ProbDev<double>* fuse_pr,
                        const StepIO<double>* fuse_io) {
  // SE_KSTAR_R = 1 / 2: the one-warp-per-row(-pair) kernels, kept for comparison; default: the row-group kernel
  static const int r_env = getenv("SE_KSTAR_R") ? atoi(getenv("SE_KSTAR_R")) : 8;
  const bool want_jac = jac != nullptr;
  if (r_env != 1 && r_env != 2) {
    static const int nostore = getenv("SE_KSTAR_NOSTORE") ? atoi(getenv("SE_KSTAR_NOSTORE")) : 0;  // measurement only
    if (nostore) S = 0;
    // seg_per_cta > 0: the training set is split over blockIdx.y in runs of seg_per_cta 128-point segments (small
    // batches); 0: one CTA walks all of it
    const int segs = gp.n_pad / 128;
    const int spc = seg_per_cta > 0 ? seg_per_cta : segs;
    dim3 grid8((unsigned)ceil_div(rows_pad / 8, 2), (unsigned)ceil_div(segs, spc));
    // One output per warp group (NS x the warps per CTA, 80 registers each): the default for plain launches of models
    // with two or more outputs -- C3 kernel rows 2.45 -> 1.83 ms, C2 0.221 -> 0.208 ms (SE_KSTAR_OPT=0: all outputs per
    // thread).  The fused-step launches of small batches keep one thread per cell.
    static const int opt_env = getenv("SE_KSTAR_OPT") ? atoi(getenv("SE_KSTAR_OPT")) : 1;
    if (opt_env == 1 && !fuse_io && gp.ns >= 2)
      return launch_kstar_slices_opt1(gp, p, p_stride, u, u_stride, row0, nrows, rows_pad, exp_a, S, a_img, mean, jac,
                                      spc, kpart, grid8.x, grid8.y, stream);
#define SE_CASE8(NS_, NU_)                                                                                          \
  if (gp.ns == NS_ && gp.nu == NU_) {                                                                               \
    if (fuse_io) {                                                                                                  \
      SE_REQUIRE(want_jac && fuse_pr, "the fused ellipsoid step precedes a kernel-row launch with Jacobians");      \
      oz::FuseArgs<true> fa;                                                                                        \
      fa.pr = *fuse_pr, fa.io = *fuse_io;                                                                           \
      SE_LAUNCH((oz::gp_kstar_slices8_kernel<NS_, NS_ + NU_, true, true>), grid8, 64, 0, stream, gp, p, p_stride,   \
                u, u_stride, row0, nrows, rows_pad, exp_a, S, a_img, mean, jac, spc, kpart, fa);                    \
      return 0;                                                                                                     \
    }                                                                                                               \
    if (want_jac)                                                                                                   \
      SE_LAUNCH((oz::gp_kstar_slices8_kernel<NS_, NS_ + NU_, true>), grid8, 64, 0, stream, gp, p, p_stride, u,      \
                u_stride, row0, nrows, rows_pad, exp_a, S, a_img, mean, jac, spc, kpart, oz::FuseArgs<false>());    \
    else                                                                                                            \
      SE_LAUNCH((oz::gp_kstar_slices8_kernel<NS_, NS_ + NU_, false>), grid8, 64, 0, stream, gp, p, p_stride, u,     \
                u_stride, row0, nrows, rows_pad, exp_a, S, a_img, mean, jac, spc, kpart, oz::FuseArgs<false>());    \
    return 0;                                                                                                       \
  }
    SE_FOR_NS_NU(SE_CASE8)
#undef SE_CASE8
    return fail("unsupported (n_s, n_u) = (%d, %d)", gp.ns, gp.nu);
  }
  const int R = r_env == 2 ? 2 : 1;
  const int warps = rows_pad / R;
  const int grid = (int)ceil_div(warps, 8);
#define SE_CASE(NS_, NU_)                                                                                        \
  if (gp.ns == NS_ && gp.nu == NU_) {                                                                            \
    if (want_jac && R == 1)                                                                                      \
      SE_LAUNCH((oz::gp_kstar_slices_kernel<NS_, NS_ + NU_, true, 1>), grid, 256, 0, stream, gp, p, p_stride, u, \
                u_stride, row0, nrows, rows_pad, exp_a, S, a_img, mean, jac);                                    \
    else if (want_jac)                                                                                           \
      SE_LAUNCH((oz::gp_kstar_slices_kernel<NS_, NS_ + NU_, true, 2>), grid, 256, 0, stream, gp, p, p_stride, u, \
                u_stride, row0, nrows, rows_pad, exp_a, S, a_img, mean, jac);                                    \
    else if (R == 1)                                                                                             \
      SE_LAUNCH((oz::gp_kstar_slices_kernel<NS_, NS_ + NU_, false, 1>), grid, 256, 0, stream, gp, p, p_stride,   \
                u, u_stride, row0, nrows, rows_pad, exp_a, S, a_img, mean, jac);                                 \
    else                                                                                                         \
      SE_LAUNCH((oz::gp_kstar_slices_kernel<NS_, NS_ + NU_, false, 2>), grid, 256, 0, stream, gp, p, p_stride,   \
                u, u_stride, row0, nrows, rows_pad, exp_a, S, a_img, mean, jac);                                 \
    return 0;                                                                                                    \
  }
  SE_FOR_NS_NU(SE_CASE)
#undef SE_CASE
  return fail("unsupported (n_s, n_u) = (%d, %d)", gp.ns, gp.nu);
}

template <int S, int MODE>
static int launch_oz_gemm_s(const int8_t* a_img, const int8_t* b_img, const double* scale_b, int n, int n_pad,
                            int rows_pad, const TriOrder& od, int grid, double* partial, cudaStream_t stream) {
  using C = oz::Cfg<S>;
  SE_TRY(set_max_dyn_smem((oz::gp_trisumsq_ozaki_kernel<S, MODE>), C::SMEM_BYTES));
  static const int dbg = getenv("SE_OZ_DBG") ? atoi(getenv("SE_OZ_DBG")) : 0;  // measurement knobs, see the kernel
  SE_LAUNCH((oz::gp_trisumsq_ozaki_kernel<S, MODE>), grid, oz::THREADS, C::SMEM_BYTES, stream, a_img, b_img, scale_b, n,
            n_pad, rows_pad, od, partial, dbg);
  return 0;
}

int launch_trisumsq_ozaki(const int8_t* a_img, const int8_t* b_img, const double* scale_b, int n, int n_pad,
                          int rows_pad, int ns, int S, double* partial, int sm_count, int* nb_out, cudaStream_t stream) {
  // candidate digit tiles kept L2-resident per chunk: 24 MB measured best at C2 (0.934 vs 0.943 ms at 32, 0.971 at 96)
  static const int64_t l2_budget = getenv("SE_OZ_L2MB") ? atoll(getenv("SE_OZ_L2MB")) << 20 : (24ll << 20);
  TriOrder od;
  od.nb = (n + oz::TN - 1) / oz::TN;  // row tiles that hold at least one training point
  od.rt_count = rows_pad / oz::TM;
  od.ns = ns;
  int64_t rc = l2_budget / ((int64_t)oz::TM * n_pad * S * ns);
  if (l2_budget <= 0 || rc > od.rt_count) rc = od.rt_count;
  if (rc < 1) rc = 1;
  od.rc = (int)rc;
  const int64_t items = (int64_t)od.nb * od.rt_count * ns;
  const int grid = (int)(items < sm_count ? items : sm_count);
  *nb_out = 2 * od.nb;  // the epilogue writes one partial block per 32 columns
  static const int reuse = getenv("SE_OZ_REUSE") ? atoi(getenv("SE_OZ_REUSE")) : 2;
  // SE_OZ_CLUSTER = (CTAs sharing a candidate tile)(CTAs sharing a row tile); 0: one CTA per item.  Default 12: pairs of
  // CTAs on one row tile of L^-1 and two neighbouring candidate tiles, the L^-1 digit planes multicast to both (measured,
  // isolated launches: C2 0.892 -> 0.884 ms, C3 (n = 5000, n_s = 4: the 629 MB image of L^-1 does not fit L2) 12.9 ->
  // 11.9 ms).  21 (pairs on one candidate tile, kernel rows multicast, the shallower row tile padded with the zeros above
  // the diagonal) is slower: C2 0.934, C3 12.6.  Clusters of four (22, 14, 41) ran at half speed and were removed.
  static const int shape_env = getenv("SE_OZ_CLUSTER") ? atoi(getenv("SE_OZ_CLUSTER")) : 12;
  int shape = (shape_env == 12 || shape_env == 21) ? shape_env : 0;
  if (shape == 12 && od.rt_count % 2 != 0) shape = 0;  // an odd number of candidate tiles: no pairs
  if (sm_count < 2) shape = 0;
  if (shape > 0 && reuse == 2)
    return launch_trisumsq_ozaki_cluster(S, shape, a_img, b_img, scale_b, n, n_pad, rows_pad, od, sm_count, partial, stream);
#define SE_OZ_CASE(S_)                                                                                               \
  case S_:                                                                                                           \
    return reuse == 2 ? launch_oz_gemm_s<S_, 2>(a_img, b_img, scale_b, n, n_pad, rows_pad, od, grid, partial, stream)  \
           : reuse ? launch_oz_gemm_s<S_, 1>(a_img, b_img, scale_b, n, n_pad, rows_pad, od, grid, partial, stream)     \
                   : launch_oz_gemm_s<S_, 0>(a_img, b_img, scale_b, n, n_pad, rows_pad, od, grid, partial, stream);
  switch (S) {
    SE_OZ_CASE(3)
    SE_OZ_CASE(4)
    SE_OZ_CASE(5)
    SE_OZ_CASE(6)
    SE_OZ_CASE(7)
    SE_OZ_CASE(8)
  }
#undef SE_OZ_CASE
  return fail("tcgen05 variance GEMM: digit planes S = %d not compiled (3 .. 8)", S);
}

template <int S>
static int launch_probe_s(const int8_t* a_img, const int8_t* b_img, int nk, uint32_t lbo, uint32_t sbo, int32_t* out,
                          cudaStream_t stream) {
  const int smem = S * (oz::A_BLK + oz::B_BLK) + 1024 + 64;
  static const int reuse = getenv("SE_OZ_REUSE") ? atoi(getenv("SE_OZ_REUSE")) : 2;
  if (reuse == 2) {
    SE_CUDA(cudaFuncSetAttribute((oz::oz_probe_kernel<S, 2>), cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    SE_LAUNCH((oz::oz_probe_kernel<S, 2>), 1, 128, smem, stream, a_img, b_img, nk, lbo, sbo, out);
  } else if (reuse) {
    SE_CUDA(cudaFuncSetAttribute((oz::oz_probe_kernel<S, 1>), cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    SE_LAUNCH((oz::oz_probe_kernel<S, 1>), 1, 128, smem, stream, a_img, b_img, nk, lbo, sbo, out);
  } else {
    SE_CUDA(cudaFuncSetAttribute((oz::oz_probe_kernel<S, 0>), cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    SE_LAUNCH((oz::oz_probe_kernel<S, 0>), 1, 128, smem, stream, a_img, b_img, nk, lbo, sbo, out);
  }
  return 0;
}

int launch_oz_probe(const int8_t* a_img, const int8_t* b_img, int nk, int S, int lbo, int sbo, int32_t* out,
                    cudaStream_t stream) {
  switch (S) {
    case 3: return launch_probe_s<3>(a_img, b_img, nk, lbo, sbo, out, stream);
    case 4: return launch_probe_s<4>(a_img, b_img, nk, lbo, sbo, out, stream);
    case 5: return launch_probe_s<5>(a_img, b_img, nk, lbo, sbo, out, stream);
    case 6: return launch_probe_s<6>(a_img, b_img, nk, lbo, sbo, out, stream);
    case 7: return launch_probe_s<7>(a_img, b_img, nk, lbo, sbo, out, stream);
    case 8: return launch_probe_s<8>(a_img, b_img, nk, lbo, sbo, out, stream);
  }
  return fail("probe: S = %d not compiled", S);
}

static __global__ void debug_exp_kernel(const double* __restrict__ x, int64_t n, double* __restrict__ out, int table) {
  const int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  __shared__ double etab[64];
  if (threadIdx.x < 64) etab[threadIdx.x] = kExpTab[threadIdx.x];
  __syncthreads();
  if (i < n) out[i] = table ? exp_tab(x[i], etab) : exp_nb(x[i]);
}
int launch_debug_exp(const double* x, int64_t n, double* out, int table, cudaStream_t stream) {
  if (n <= 0) return 0;
  SE_LAUNCH(debug_exp_kernel, (int)ceil_div(n, (int64_t)256), 256, 0, stream, x, n, out, table);
  return 0;
}

// measured int8 tensor-pipe peak in OP/s (2 x MAC).  sustained = 0: best of a few short launches (burst clocks);
// sustained = 1: launches back to back for ~1.5 s and reports the last third (the 1 kW power cap has settled by then).
int measure_peak_i8(int sustained, int n_mma, int issuers, double* ops_out) {
  int dev = 0, sms = 0;
  SE_CUDA(cudaGetDevice(&dev));
  SE_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
  const int smem = 14 * 1024, iters = 40000 * (256 / n_mma);
  SE_CUDA(cudaFuncSetAttribute(oz::oz_peak_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
  cudaEvent_t e0, e1;
  SE_CUDA(cudaEventCreate(&e0));
  SE_CUDA(cudaEventCreate(&e1));
  const double ops = (double)sms * iters * (2.0 * 128 * n_mma * 32);
  double result = 0.0;
  if (!sustained) {
    double best_ms = 1e30;
    for (int rep = 0; rep < 4; ++rep) {
      SE_CUDA(cudaEventRecord(e0, 0));
      SE_LAUNCH(oz::oz_peak_kernel, sms, 128, smem, 0, iters, n_mma, issuers, nullptr);
      SE_CUDA(cudaEventRecord(e1, 0));
      SE_CUDA(cudaEventSynchronize(e1));
      float ms = 0.f;
      SE_CUDA(cudaEventElapsedTime(&ms, e0, e1));
      if (rep > 0 && ms < best_ms) best_ms = ms;
    }
    result = ops / (best_ms * 1e-3);
  } else {
    const int reps = 450;  // ~3 ms each
    for (int rep = 0; rep < reps; ++rep) {
      if (rep == reps * 2 / 3) SE_CUDA(cudaEventRecord(e0, 0));
      SE_LAUNCH(oz::oz_peak_kernel, sms, 128, smem, 0, iters, n_mma, issuers, nullptr);
    }
    SE_CUDA(cudaEventRecord(e1, 0));
    SE_CUDA(cudaEventSynchronize(e1));
    float ms = 0.f;
    SE_CUDA(cudaEventElapsedTime(&ms, e0, e1));
    result = ops * (reps - reps * 2 / 3) / (ms * 1e-3);
  }
  cudaEventDestroy(e0);
  cudaEventDestroy(e1);
  *ops_out = result;
  return 0;
}

}  // namespace se
