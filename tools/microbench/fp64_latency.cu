// Dependent-issue latencies on one warp: DFMA chain, LDS.64 chain (pointer chase), DFMA + LDS + STS round trip, __syncthreads.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o fp64_latency fp64_latency.cu && ./fp64_latency
#include <cstdio>
#include <cuda_runtime.h>
__global__ void k(double* out, long long* cyc, int threads_active) {
  __shared__ double sm[1024];
  __shared__ int chase[1024];
  for (int i = threadIdx.x; i < 1024; i += blockDim.x) sm[i] = 1.0 + 1e-9 * i, chase[i] = (i * 37 + 11) & 1023;
  __syncthreads();
  double a = 1.0000001, b = 0.999999, c = out[0];
  long long t0 = clock64();
#pragma unroll 1
  for (int i = 0; i < 256; ++i) { c = fma(a, c, b); c = fma(a, c, b); c = fma(a, c, b); c = fma(a, c, b); }
  long long t1 = clock64();
  int p = threadIdx.x & 1023;
#pragma unroll 1
  for (int i = 0; i < 256; ++i) { p = chase[p]; p = chase[p]; p = chase[p]; p = chase[p]; }
  long long t2 = clock64();
  int idx = threadIdx.x;
#pragma unroll 1
  for (int i = 0; i < 256; ++i) {  // load -> fma -> store -> (next load depends on the stored value's address chain)
    double v = sm[idx]; v = fma(a, v, b); sm[idx] = v; idx = (idx + 32) & 1023;
  }
  long long t3 = clock64();
#pragma unroll 1
  for (int i = 0; i < 256; ++i) __syncthreads();
  long long t4 = clock64();
  double r = 1.0 / (c + 2.0), s = sqrt(c + 3.0), q = rsqrt(c + 4.0);
  long long t5 = clock64();
  if (threadIdx.x == 0) {
    cyc[0] = (t1 - t0) / 1024, cyc[1] = (t2 - t1) / 1024, cyc[2] = (t3 - t2) / 256, cyc[3] = (t4 - t3) / 256, cyc[4] = t5 - t4;
    out[1] = c + p + sm[5] + r + s + q;
  }
}
int main() {
  double* out; long long* cyc;
  cudaMalloc(&out, 64); cudaMalloc(&cyc, 64); cudaMemset(out, 0, 64);
  for (int threads : {32, 512}) {
    k<<<1, threads>>>(out, cyc, threads);
    long long h[5];
    cudaMemcpy(h, cyc, sizeof(h), cudaMemcpyDeviceToHost);
    printf("threads %d: DFMA dependent %lld cycles, LDS dependent %lld cycles, LDS+DFMA+STS iteration %lld cycles, __syncthreads %lld cycles, div+sqrt+rsqrt %lld cycles\n",
           threads, h[0], h[1], h[2], h[3], h[4]);
  }
  return 0;
}
