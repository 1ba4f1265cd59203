// inst_cem.cu -- launchers of the CEM kernels (sampling, selection + refit, pick)
#include "cem_kernels.cuh"
#include "launchers.cuh"

namespace se {

template <typename T>
int launch_cem_sample(const ActionSrc<T>& src, int num_inst, int first, int n_local, int ns, const T* p0, T* actions,
                      T* p, T* rec_actions, cudaStream_t stream) {
  const int pairs = (src.tnu + 1) / 2;
  const int64_t total = (int64_t)num_inst * n_local * pairs;
  SE_LAUNCH((cem_sample_kernel<T>), (int)ceil_div(total, 256), 256, 0, stream, src, num_inst, first, n_local, ns, p0,
            actions, p, rec_actions);
  return 0;
}

template <typename T>
int launch_cem_noise(uint64_t seed, int iteration, int num_inst, int n_global, int first, int count, int tnu, T* out,
                     cudaStream_t stream) {
  const int pairs = (tnu + 1) / 2;
  const int64_t total = (int64_t)num_inst * count * pairs;
  SE_LAUNCH((cem_noise_kernel<T>), (int)ceil_div(total, 256), 256, 0, stream, seed, iteration, num_inst, n_global,
            first, count, tnu, out);
  return 0;
}

// 0 = never, 1 (default) = populations above kSortAllMax run the 8-CTA cluster kernel
static int select_cluster_mode() {
  static const int mode = getenv("SE_SELECT_CLUSTER") ? atoi(getenv("SE_SELECT_CLUSTER")) : 1;
  return mode;
}

template <typename T>
static int launch_cem_select_cluster(const SelectArgs<T>& a, int num_inst, cudaStream_t stream) {
  const int chunk = (int)round_up(ceil_div(a.n_global, kSelCluster), 32);
  const int cached = chunk <= kSelChunkCache ? 1 : 0;
  const int smem = kMaxElites * 12 + (cached ? chunk * 8 : 0);
  SE_TRY(set_max_dyn_smem(cem_select_cluster_kernel<T>, smem));
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3((unsigned)(num_inst * kSelCluster), 1, 1);
  cfg.blockDim = dim3(kSelThreads, 1, 1);
  cfg.dynamicSmemBytes = smem;
  cfg.stream = stream;
  cudaLaunchAttribute attr;
  attr.id = cudaLaunchAttributeClusterDimension;
  attr.val.clusterDim.x = kSelCluster;
  attr.val.clusterDim.y = 1;
  attr.val.clusterDim.z = 1;
  cfg.attrs = &attr;
  cfg.numAttrs = 1;
  SE_CUDA(cudaLaunchKernelEx(&cfg, cem_select_cluster_kernel<T>, a, chunk, cached));
  g_launches.fetch_add(1, std::memory_order_relaxed);
  return 0;
}

template <typename T>
int launch_cem_select_refit(const SelectArgs<T>& a, int num_inst, cudaStream_t stream) {
  if (a.n_global > kSortAllMax && select_cluster_mode() != 0) return launch_cem_select_cluster<T>(a, num_inst, stream);
  // selected keys + indices, and (populations of 4097 .. 16384) a cache of every key
  const int cache_n = a.n_global > kSortAllMax && a.n_global <= 16384 ? a.n_global : 0;
  const int smem = kMaxElites * 12 + cache_n * 8;
  SE_TRY(set_max_dyn_smem(cem_select_refit_kernel<T>, smem));
  SE_LAUNCH((cem_select_refit_kernel<T>), num_inst, kSelThreads, smem, stream, a, cache_n);
  return 0;
}

template <typename T>
int launch_cem_pick(const CostView<T>& costs, int n_global, int pick_best, const ActionSrc<T>& src, int num_inst,
                    T* actions_out, int32_t* found, int32_t* best_id, cudaStream_t stream) {
  SE_LAUNCH((cem_pick_kernel<T>), num_inst, 256, 0, stream, costs, n_global, pick_best, src, actions_out, found,
            best_id);
  return 0;
}

template <typename T>
int launch_cem_pack_costs(const T* objective, const T* constraint, int64_t count, T* pack, T* rec_obj, T* rec_con,
                          cudaStream_t stream) {
  SE_LAUNCH((cem_pack_costs_kernel<T>), (int)ceil_div(count, 256), 256, 0, stream, objective, constraint, count, pack,
            rec_obj, rec_con);
  return 0;
}

int launch_cem_set_seed(uint64_t* dst, uint64_t seed, cudaStream_t stream) {
  SE_LAUNCH(cem_set_seed_kernel, 1, 1, 0, stream, dst, seed);
  return 0;
}

int launch_cem_init_dist(double* mean, double* stdv, int count, cudaStream_t stream) {
  SE_LAUNCH(cem_init_dist_kernel, (int)ceil_div(count, 256), 256, 0, stream, mean, stdv, count);
  return 0;
}

#define SE_INST(T)                                                                                                  \
  template int launch_cem_sample<T>(const ActionSrc<T>&, int, int, int, int, const T*, T*, T*, T*, cudaStream_t);   \
  template int launch_cem_noise<T>(uint64_t, int, int, int, int, int, int, T*, cudaStream_t);                       \
  template int launch_cem_select_refit<T>(const SelectArgs<T>&, int, cudaStream_t);                                 \
  template int launch_cem_pick<T>(const CostView<T>&, int, int, const ActionSrc<T>&, int, T*, int32_t*, int32_t*,   \
                                  cudaStream_t);                                                                    \
  template int launch_cem_pack_costs<T>(const T*, const T*, int64_t, T*, T*, T*, cudaStream_t);
SE_INST(float)
SE_INST(double)
#undef SE_INST

}  // namespace se
