// inst_gemm.cu -- variance GEMM launchers, model packing and the pipe-peak micro-benchmarks
#include "gp_kernels.cuh"
#include "launchers.cuh"

namespace se {

template <typename T>
static int launch_trisumsq_simt(const T* kst, const T* w, int n_pad, int rows_pad, int ns, T* partial, int* counter,
                                int sm_count, int* nb_out, cudaStream_t stream) {
  const int nb = n_pad / 64;
  const int64_t items = (int64_t)nb * (rows_pad / 64) * ns;
  const int grid = (int)(items < (int64_t)sm_count * 4 ? items : (int64_t)sm_count * 4);
  SE_CUDA(cudaMemsetAsync(counter, 0, sizeof(int), stream));
  SE_LAUNCH((gp_trisumsq_simt_kernel<T>), grid, 256, 0, stream, kst, w, n_pad, rows_pad, ns, partial, counter);
  *nb_out = nb;
  return 0;
}

template <>
int launch_trisumsq<float>(se_gemm_impl impl, const float* kst, const float* w, int n_pad, int rows_pad, int ns,
                           float* partial, int* counter, int sm_count, int* nb_out, cudaStream_t stream) {
  (void)impl;
  return launch_trisumsq_simt<float>(kst, w, n_pad, rows_pad, ns, partial, counter, sm_count, nb_out, stream);
}

template <>
int launch_trisumsq<double>(se_gemm_impl impl, const double* kst, const double* w, int n_pad, int rows_pad, int ns,
                            double* partial, int* counter, int sm_count, int* nb_out, cudaStream_t stream) {
  if (impl == SE_GEMM_SIMT)
    return launch_trisumsq_simt<double>(kst, w, n_pad, rows_pad, ns, partial, counter, sm_count, nb_out, stream);
  using C = DmmaCfg;
  static bool attr_set = false;
  if (!attr_set) {
    SE_CUDA(cudaFuncSetAttribute(gp_trisumsq_dmma_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                 C::SMEM_BYTES));
    attr_set = true;
  }
  const int nb = n_pad / C::BN;
  const int64_t items = (int64_t)nb * (rows_pad / C::BM) * ns;
  const int grid = (int)(items < sm_count ? items : sm_count);
  SE_CUDA(cudaMemsetAsync(counter, 0, sizeof(int), stream));
  SE_LAUNCH(gp_trisumsq_dmma_kernel, grid, C::THREADS, C::SMEM_BYTES, stream, kst, w, n_pad, rows_pad, ns, partial,
            counter);
  *nb_out = nb;
  return 0;
}

template <typename T>
int launch_var(const GpDev<T>& gp, const T* partial, int nb, int rows_pad, int64_t row0, int nrows, T* var,
               cudaStream_t stream) {
  const int64_t total = (int64_t)nrows * gp.ns;
  SE_LAUNCH((gp_var_kernel<T>), (int)ceil_div(total, 256), 256, 0, stream, gp, partial, nb, rows_pad, row0, nrows,
            var);
  return 0;
}
template int launch_var<float>(const GpDev<float>&, const float*, int, int, int64_t, int, float*, cudaStream_t);
template int launch_var<double>(const GpDev<double>&, const double*, int, int, int64_t, int, double*, cudaStream_t);

template <typename T>
int launch_pack_model(const T* x, const T* alpha, const T* linv, int n, int d, int ns, int n_pad, T* xt, T* alpha_p,
                      T* w, cudaStream_t stream) {
  SE_LAUNCH((pack_xt_kernel<T>), (int)ceil_div((int64_t)d * n_pad, 256), 256, 0, stream, x, n, d, n_pad, xt);
  SE_LAUNCH((pack_rows_kernel<T>), (int)ceil_div((int64_t)ns * n_pad, 256), 256, 0, stream, alpha, ns, n, n_pad,
            alpha_p);
  dim3 grid((unsigned)ceil_div(n_pad, 256), (unsigned)n_pad, (unsigned)ns);
  SE_LAUNCH((pack_w_kernel<T>), grid, 256, 0, stream, linv, n, n_pad, w);
  return 0;
}
template int launch_pack_model<float>(const float*, const float*, const float*, int, int, int, int, float*, float*,
                                      float*, cudaStream_t);
template int launch_pack_model<double>(const double*, const double*, const double*, int, int, int, int, double*,
                                       double*, double*, cudaStream_t);

// ------------------------------------------------------------------------------------------------------------------
// pipe peaks: the roofline denominators bench.py quotes for the FP32 / FP64 pipes (MEASURED_PEAKS.json only holds
// HBM and bf16 tensor numbers)
// ------------------------------------------------------------------------------------------------------------------
template <typename T>
__global__ void __launch_bounds__(256) peak_fma_kernel(T* out, int iters, T seed) {
  T a[16];
#pragma unroll
  for (int i = 0; i < 16; ++i) a[i] = seed + T(i) * T(1e-3) + T(threadIdx.x) * T(1e-6);
  const T b = T(0.999999), c = T(1e-7);
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int i = 0; i < 16; ++i) a[i] = a[i] * b + c;
  }
  T s = T(0);
#pragma unroll
  for (int i = 0; i < 16; ++i) s += a[i];
  if (s == T(123.456)) out[0] = s;
}

__global__ void __launch_bounds__(256) peak_dmma_kernel(double* out, int iters) {
  double c[8][2];
#pragma unroll
  for (int i = 0; i < 8; ++i) c[i][0] = c[i][1] = 0.0;
  const double a = 1.0 + threadIdx.x * 1e-9, b = 1e-9;
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int i = 0; i < 8; ++i) dmma_m8n8k4(c[i][0], c[i][1], a, b);
  }
  double s = 0.0;
#pragma unroll
  for (int i = 0; i < 8; ++i) s += c[i][0] + c[i][1];
  if (s == 123.456) out[0] = s;
}

int measure_peak(int kind, double* flops_out) {
  int dev = 0, sms = 0;
  SE_CUDA(cudaGetDevice(&dev));
  SE_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
  double* scratch = nullptr;
  SE_CUDA(cudaMalloc(&scratch, 64));
  cudaEvent_t e0, e1;
  SE_CUDA(cudaEventCreate(&e0));
  SE_CUDA(cudaEventCreate(&e1));
  const int grid = sms * 8, iters = kind == 0 ? 20000 : 4000;
  double best_ms = 1e30;
  for (int rep = 0; rep < 4; ++rep) {
    SE_CUDA(cudaEventRecord(e0, 0));
    if (kind == 0) {
      SE_LAUNCH((peak_fma_kernel<float>), grid, 256, 0, 0, reinterpret_cast<float*>(scratch), iters, 1.0f);
    } else if (kind == 1) {
      SE_LAUNCH((peak_fma_kernel<double>), grid, 256, 0, 0, scratch, iters, 1.0);
    } else {
      SE_LAUNCH(peak_dmma_kernel, grid, 256, 0, 0, scratch, iters);
    }
    SE_CUDA(cudaEventRecord(e1, 0));
    SE_CUDA(cudaEventSynchronize(e1));
    float ms = 0.f;
    SE_CUDA(cudaEventElapsedTime(&ms, e0, e1));
    if (rep > 0 && ms < best_ms) best_ms = ms;
  }
  double flop;
  if (kind == 2)
    flop = (double)grid * 8 /*warps*/ * (double)iters * 8 /*mma*/ * 512.0 /*2*8*8*4*/;
  else
    flop = (double)grid * 256.0 * (double)iters * 16.0 * 2.0;
  *flops_out = flop / (best_ms * 1e-3);
  cudaEventDestroy(e0);
  cudaEventDestroy(e1);
  cudaFree(scratch);
  return 0;
}

}  // namespace se
