// cem_kernels.cuh -- the CEM iteration on the device (a1-a6 in SURVEY.md 8a): candidate sampling, feasibility
// partition + top-E selection with an index tie-break, Gaussian refit, and the final pick.
//
// Restates constrained-cem-mpc/constrained_cem_mpc/constrained_cem_mpc.py:158-170 (sampling), :222-257 (elite
// selection and refit) and :259-280 (get_actions).  No host round trips: every decision is taken by these kernels.
#pragma once
#include <cooperative_groups.h>

#include "common.cuh"

namespace se {

// ---- Philox4x32-10 (Salmon et al., SC'11) + Box-Muller; restated bit-for-bit in oracle/oracle.py -----------------
__device__ __forceinline__ void philox4x32_10(uint32_t (&c)[4], uint32_t k0, uint32_t k1) {
#pragma unroll
  for (int r = 0; r < 10; ++r) {
    const uint32_t hi0 = __umulhi(0xD2511F53u, c[0]), lo0 = 0xD2511F53u * c[0];
    const uint32_t hi1 = __umulhi(0xCD9E8D57u, c[2]), lo1 = 0xCD9E8D57u * c[2];
    const uint32_t n0 = hi1 ^ c[1] ^ k0, n2 = hi0 ^ c[3] ^ k1;
    c[0] = n0, c[1] = lo1, c[2] = n2, c[3] = lo0;
    k0 += 0x9E3779B9u;
    k1 += 0xBB67AE85u;
  }
}

// two standard normals for (candidate, pair m, iteration)
__device__ __forceinline__ void philox_normal_pair(uint64_t seed, uint32_t iteration, uint32_t candidate, uint32_t m,
                                                   double& z0, double& z1) {
  uint32_t c[4] = {candidate, m, iteration, 0x5ECE3u};
  philox4x32_10(c, (uint32_t)(seed & 0xFFFFFFFFull), (uint32_t)(seed >> 32));
  const uint64_t m1 = ((uint64_t)c[0] << 21) | (c[1] >> 11);
  const uint64_t m2 = ((uint64_t)c[2] << 21) | (c[3] >> 11);
  const double u1 = ((double)m1 + 0.5) * 0x1p-53;
  const double u2 = ((double)m2 + 0.5) * 0x1p-53;
  const double rad = sqrt(-2.0 * log(u1));
  double sn, cs;
  sincospi(2.0 * u2, &sn, &cs);
  z0 = rad * cs;
  z1 = rad * sn;
}

// where the candidates of an iteration come from
template <typename T>
struct ActionSrc {
  int mode;          // 0 = Philox, 1 = external standard-normal noise, 2 = explicit action tensor
  uint64_t seed;
  const uint64_t* seed_dev;  // when set, the seed is read from device memory (a replayed CUDA graph gets a new one)
  int iteration;
  int n_global;      // rollouts per instance over all ranks
  int inst_offset;   // global index of local instance 0 (instance sharding): the RNG is keyed by GLOBAL instance id
  int tnu;           // T * n_u
  const T* ext;      // mode 1: noise [M, N_global, T*n_u] of this iteration; mode 2: actions [M, N_global, T*n_u]
  const double* mean;  // [M, T*n_u] sampling distribution of this iteration
  const double* stdv;
};

template <typename T>
__device__ __forceinline__ T candidate_action(const ActionSrc<T>& s, int inst, int gid, int e) {
  const int64_t flat = ((int64_t)inst * s.n_global + gid) * s.tnu + e;
  if (s.mode == 2) return s.ext[flat];
  double z;
  if (s.mode == 1) {
    z = (double)s.ext[flat];
  } else {
    double z0, z1;
    philox_normal_pair(s.seed_dev ? *s.seed_dev : s.seed, (uint32_t)s.iteration,
                       (uint32_t)((inst + s.inst_offset) * s.n_global + gid), (uint32_t)(e >> 1), z0, z1);
    z = (e & 1) ? z1 : z0;
  }
  // actions = means + stds * eps  ==  Normal(means, stds).sample()   (constrained_cem_mpc.py:167)
  return (T)(s.mean[(int64_t)inst * s.tnu + e] + s.stdv[(int64_t)inst * s.tnu + e] * z);
}

// actions[cl, :] for the local shard and p[cl, :] = p0[inst, :]; one thread per (local candidate, pair)
template <typename T>
__global__ void cem_sample_kernel(ActionSrc<T> src, int num_inst, int first, int n_local, int ns,
                                  const T* __restrict__ p0, T* __restrict__ actions, T* __restrict__ p,
                                  T* __restrict__ rec_actions) {
  const int pairs = (src.tnu + 1) >> 1;
  const int64_t idx = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (idx >= (int64_t)num_inst * n_local * pairs) return;
  const int m = idx % pairs;
  const int64_t cl = idx / pairs;  // local candidate = inst * n_local + l
  const int inst = cl / n_local, gid = first + (int)(cl % n_local);
#pragma unroll
  for (int h = 0; h < 2; ++h) {
    const int e = 2 * m + h;
    if (e < src.tnu) {
      const T a = candidate_action<T>(src, inst, gid, e);
      actions[cl * src.tnu + e] = a;
      if (rec_actions) rec_actions[cl * src.tnu + e] = a;
    }
  }
  if (m == 0)
    for (int i = 0; i < ns; ++i) p[cl * ns + i] = p0[inst * ns + i];
}

template <typename T>
__global__ void cem_noise_kernel(uint64_t seed, int iteration, int num_inst, int n_global, int first, int count,
                                 int tnu, T* __restrict__ out) {
  const int pairs = (tnu + 1) >> 1;
  const int64_t idx = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (idx >= (int64_t)num_inst * count * pairs) return;
  const int m = idx % pairs;
  const int64_t cl = idx / pairs;
  const int inst = cl / count, gid = first + (int)(cl % count);
  double z0, z1;
  philox_normal_pair(seed, (uint32_t)iteration, (uint32_t)(inst * n_global + gid), (uint32_t)m, z0, z1);
  if (2 * m < tnu) out[cl * tnu + 2 * m] = (T)z0;
  if (2 * m + 1 < tnu) out[cl * tnu + 2 * m + 1] = (T)z1;
}

// ---- selection ------------------------------------------------------------------------------------------------------
// costs of all ranks after the all-gather: [W][2][M][N_local]; plane 0 = objective, plane 1 = constraint
template <typename T>
struct CostView {
  const T* base;
  int num_inst, n_local;
  __device__ __forceinline__ T get(int plane, int inst, int gid) const {
    const int r = gid / n_local, l = gid % n_local;
    return base[(((int64_t)r * 2 + plane) * num_inst + inst) * n_local + l];
  }
};

// monotone map double -> uint64 (total order; -0 folded into +0; NaN sorts last like torch.sort)
__device__ __forceinline__ uint64_t orderable(double x) {
  x = x + 0.0;
  const uint64_t b = (uint64_t)__double_as_longlong(x);
  return (b >> 63) ? ~b : (b | 0x8000000000000000ull);
}

constexpr int kSelThreads = 1024;
constexpr int kMaxElites = 4096;
constexpr int kSortAllMax = kMaxElites;  // populations up to this size go through the sorting network whole (smem arrays hold kMaxElites keys)

template <typename T>
struct SelectArgs {
  CostView<T> costs;
  int n_global, num_elites;
  ActionSrc<T> src;       // regenerates the elites' action sequences (no second collective)
  double* mean_next;      // [M, T*n_u]
  double* std_next;
  int32_t* elite_ids;     // [M, E] or null
  T* mean_hist;           // [M, T*n_u] slice of this iteration, or null
  T* std_hist;
};

__device__ __forceinline__ int block_sum_int(int v, int* s_warp) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
  for (int s = 16; s > 0; s >>= 1) v += __shfl_xor_sync(0xffffffffu, v, s);
  __syncthreads();
  if (lane == 0) s_warp[warp] = v;
  __syncthreads();
  int t = 0;
  for (int w = 0; w < (int)(blockDim.x >> 5); ++w) t += s_warp[w];
  return t;
}

// Tail shared by the selection kernels: bitonic sort of the P (power of two, padded with ~0 keys) selected (key, index)
// pairs by ascending (cost, global index), elite ids out, then the refit -- mean and Bessel-corrected std over the
// elites (constrained_cem_mpc.py:234-235), two passes in fp64, one warp per element of the [T, n_u] action sequence,
// fixed summation order.  Called by all kSelThreads threads of ONE CTA.
template <typename T>
__device__ __forceinline__ void sort_and_refit(const SelectArgs<T>& a, int inst, uint64_t* s_key, int* s_idx, int P) {
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int E = a.num_elites;
  __syncthreads();
  for (int size = 2; size <= P; size <<= 1) {
    for (int stride = size >> 1; stride > 0; stride >>= 1) {
      for (int t = tid; t < (P >> 1); t += kSelThreads) {
        const int lo = 2 * t - (t & (stride - 1));
        const int hi = lo + stride;
        const bool ascending = (lo & size) == 0;
        const uint64_t kl = s_key[lo], kh = s_key[hi];
        const int il = s_idx[lo], ih = s_idx[hi];
        const bool lo_gt_hi = kl > kh || (kl == kh && il > ih);
        if (lo_gt_hi == ascending) {
          s_key[lo] = kh, s_key[hi] = kl;
          s_idx[lo] = ih, s_idx[hi] = il;
        }
      }
      __syncthreads();
    }
  }
  if (a.elite_ids)
    for (int i = tid; i < E; i += kSelThreads) a.elite_ids[(int64_t)inst * E + i] = s_idx[i];

  const int tnu = a.src.tnu;
  for (int e = warp; e < tnu; e += kSelThreads / 32) {
    double s = 0.0;
    for (int k = lane; k < E; k += 32) s += (double)candidate_action<T>(a.src, inst, s_idx[k], e);
#pragma unroll
    for (int sh = 16; sh > 0; sh >>= 1) s += __shfl_xor_sync(0xffffffffu, s, sh);
    const double mean = s / (double)E;
    double v = 0.0;
    for (int k = lane; k < E; k += 32) {
      const double dlt = (double)candidate_action<T>(a.src, inst, s_idx[k], e) - mean;
      v += dlt * dlt;
    }
#pragma unroll
    for (int sh = 16; sh > 0; sh >>= 1) v += __shfl_xor_sync(0xffffffffu, v, sh);
    const double sd = sqrt(v / (double)(E - 1));  // E == 1 -> NaN, like torch.std (quirk Q5)
    if (lane == 0) {
      a.mean_next[(int64_t)inst * tnu + e] = mean;
      a.std_next[(int64_t)inst * tnu + e] = sd;
      if (a.mean_hist) a.mean_hist[(int64_t)inst * tnu + e] = (T)mean;
      if (a.std_hist) a.std_hist[(int64_t)inst * tnu + e] = (T)sd;
    }
  }
}

// One CTA per MPC instance.  Elite order = ascending (cost, global index): the reference's torch.sort is unstable and
// ties are the norm (constraint costs are small integers), so the index tie-break is fixed here and in the oracle.
template <typename T>
__global__ void __launch_bounds__(kSelThreads)
cem_select_refit_kernel(SelectArgs<T> a, int cache_n) {
  extern __shared__ __align__(16) unsigned char sel_smem[];
  uint64_t* s_key = reinterpret_cast<uint64_t*>(sel_smem);          // [P]
  int* s_idx = reinterpret_cast<int*>(s_key + kMaxElites);           // [P]
  uint64_t* s_all = reinterpret_cast<uint64_t*>(s_idx + kMaxElites); // [n] keys of the whole population when cache_n >= n
  __shared__ int s_hist[256];
  __shared__ int s_warp[kSelThreads / 32];
  __shared__ int s_warp2[kSelThreads / 32];
  __shared__ uint64_t s_prefix;
  __shared__ int s_remaining;

  const int inst = blockIdx.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int n = a.n_global, E = a.num_elites;

  // feasible = constraint_costs == 0   (constrained_cem_mpc.py:247)
  int cnt = 0;
  for (int g = tid; g < n; g += kSelThreads) cnt += ((double)a.costs.get(1, inst, g) == 0.0) ? 1 : 0;
  const int num_feasible = block_sum_int(cnt, s_warp);
  const bool by_objective = num_feasible >= E;  // :248

  auto key_of = [&](int g) -> uint64_t {
    if (by_objective) {
      if ((double)a.costs.get(1, inst, g) != 0.0) return ~0ull;  // infeasible: never selected
      return orderable((double)a.costs.get(0, inst, g));
    }
    return orderable((double)a.costs.get(1, inst, g));
  };

  int P = 1;  // length of the (key, index) array the bitonic network below sorts
  if (n <= kSortAllMax) {
    // Populations up to 4096 (the reference's own sizes are N = 20 .. 500): no selection pass at all, the whole
    // population goes through the sorting network (55 compare-exchange stages of one comparator per thread at
    // N <= 1024, 78 of two at 4096).  Measured alternatives: ranking every key by counting costs N^2 64-bit comparisons
    // on this one SM (9 us at N = 500), the radix select below re-reads the costs eight times (39 us at N = 2048).
    while (P < n) P <<= 1;
    for (int g = tid; g < P; g += kSelThreads) {
      s_key[g] = g < n ? key_of(g) : ~0ull;
      s_idx[g] = g < n ? g : 0x7FFFFFFF;
    }
  } else {
  // radix select: the E-th smallest key, most significant byte first.  The keys are formed once and kept in shared
  // memory when the population fits (cache_n >= n: up to 16384 candidates), so the eight passes and the compaction do
  // not go back to the cost planes in global memory.
  const bool cached = cache_n >= n;
  if (cached)
    for (int g = tid; g < n; g += kSelThreads) s_all[g] = key_of(g);
  auto key_at = [&](int g) -> uint64_t { return cached ? s_all[g] : key_of(g); };
  if (tid == 0) {
    s_prefix = 0;
    s_remaining = E;
  }
  uint64_t mask = 0;
  for (int pass = 0; pass < 8; ++pass) {
    const int shift = 56 - 8 * pass;
    if (tid < 256) s_hist[tid] = 0;
    __syncthreads();
    const uint64_t prefix = s_prefix;
    // neighbouring costs share their leading bytes, so on the first passes every key lands in ONE bin: a thread counts
    // runs of equal bins in a register and issues one atomic per run (16 same-address atomics per thread otherwise)
    int run_bin = -1, run_len = 0;
    for (int g = tid; g < n; g += kSelThreads) {
      const uint64_t k = key_at(g);
      if ((k & mask) != prefix) continue;
      const int bin = (int)((k >> shift) & 255);
      if (bin != run_bin) {
        if (run_len) atomicAdd(&s_hist[run_bin], run_len);
        run_bin = bin, run_len = 0;
      }
      ++run_len;
    }
    if (run_len) atomicAdd(&s_hist[run_bin], run_len);
    __syncthreads();
    if (tid == 0) {
      int rem = s_remaining, b = 0;
      for (; b < 255; ++b) {
        if (s_hist[b] >= rem) break;
        rem -= s_hist[b];
      }
      s_remaining = rem;
      s_prefix = prefix | ((uint64_t)b << shift);
    }
    mask |= 0xFFull << shift;
    __syncthreads();
  }
  const uint64_t vstar = s_prefix;
  const int take_equal = s_remaining;  // how many keys == v* belong to the elite set (lowest indices first)
  const int num_less = E - take_equal;

  // ordered compaction in index order
  int base_less = 0, base_eq = 0;
  for (int g0 = 0; g0 < n; g0 += kSelThreads) {
    const int g = g0 + tid;
    uint64_t k = ~0ull;
    bool f_less = false, f_eq = false;
    if (g < n) {
      k = key_at(g);
      f_less = k < vstar;
      f_eq = k == vstar;
    }
    const unsigned b_less = __ballot_sync(0xffffffffu, f_less), b_eq = __ballot_sync(0xffffffffu, f_eq);
    __syncthreads();
    if (lane == 0) {
      s_warp[warp] = __popc(b_less);
      s_warp2[warp] = __popc(b_eq);
    }
    __syncthreads();
    int off_less = base_less, off_eq = base_eq, tot_less = 0, tot_eq = 0;
    for (int w = 0; w < kSelThreads / 32; ++w) {
      if (w < warp) {
        off_less += s_warp[w];
        off_eq += s_warp2[w];
      }
      tot_less += s_warp[w];
      tot_eq += s_warp2[w];
    }
    const unsigned lt = (1u << lane) - 1u;
    if (f_less) {
      const int slot = off_less + __popc(b_less & lt);
      s_key[slot] = k;
      s_idx[slot] = g;
    }
    if (f_eq) {
      const int slot = off_eq + __popc(b_eq & lt);
      if (slot < take_equal) {
        s_key[num_less + slot] = k;
        s_idx[num_less + slot] = g;
      }
    }
    base_less += tot_less;
    base_eq += tot_eq;
  }
  // pad the E selected keys to a power of two
  while (P < E) P <<= 1;
  __syncthreads();
  for (int i = E + tid; i < P; i += kSelThreads) {
    s_key[i] = ~0ull;
    s_idx[i] = 0x7FFFFFFF;
  }
  }  // n > kSortAllMax
  sort_and_refit<T>(a, inst, s_key, s_idx, P);
}

// Populations above kSortAllMax: the same selection spread over a thread-block CLUSTER per instance (8 CTAs, one GPC).
// CTA r of the cluster owns the contiguous index range [r chunk, (r + 1) chunk) of the global population and keeps its
// keys in its own shared memory; the 8 radix passes build one 256-bin histogram per CTA and every CTA adds the eight of
// them through distributed shared memory (two cluster barriers per pass), so each takes the same decision; the ordered
// compaction writes the selected (key, index) pairs straight into the leader's arrays at offsets given by the
// per-CTA counts (index order = rank order), and the leader sorts the E elites and refits.  The cost of the selection no
// longer grows with the world size: 131072 keys are 16 per thread and pass (measured: see DESIGN.md section 6).
constexpr int kSelCluster = 8;
constexpr int kSelChunkCache = 16384;  // keys a CTA can keep in shared memory (128 KB)

template <typename T>
__global__ void __launch_bounds__(kSelThreads)
cem_select_cluster_kernel(SelectArgs<T> a, int chunk, int cached) {
  namespace cg = cooperative_groups;
  cg::cluster_group cluster = cg::this_cluster();
  extern __shared__ __align__(16) unsigned char sel_smem[];
  uint64_t* s_key = reinterpret_cast<uint64_t*>(sel_smem);          // [kMaxElites]  (used on the leader)
  int* s_idx = reinterpret_cast<int*>(s_key + kMaxElites);           // [kMaxElites]
  uint64_t* s_all = reinterpret_cast<uint64_t*>(s_idx + kMaxElites); // [chunk] this CTA's keys when cached
  __shared__ int s_hist[256];
  __shared__ int s_ghist[256];
  __shared__ int s_warp[kSelThreads / 32];
  __shared__ int s_warp2[kSelThreads / 32];
  __shared__ int s_cnt[2];  // read by the other CTAs of the cluster: (feasible | less, equal)
  __shared__ uint64_t s_prefix;
  __shared__ int s_remaining;

  const int crank = (int)cluster.block_rank(), csize = (int)cluster.num_blocks();
  const int inst = blockIdx.x / csize, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int n = a.n_global, E = a.num_elites;
  const int g0 = min(n, crank * chunk), g1 = min(n, g0 + chunk);

  // feasible = constraint_costs == 0   (constrained_cem_mpc.py:247), counted over the whole population
  int cnt = 0;
  for (int g = g0 + tid; g < g1; g += kSelThreads) cnt += ((double)a.costs.get(1, inst, g) == 0.0) ? 1 : 0;
  cnt = block_sum_int(cnt, s_warp);
  if (tid == 0) s_cnt[0] = cnt;
  cluster.sync();
  int num_feasible = 0;
  for (int r = 0; r < csize; ++r) num_feasible += cluster.map_shared_rank(s_cnt, r)[0];
  const bool by_objective = num_feasible >= E;  // :248
  cluster.sync();  // everyone has read s_cnt before it is reused

  auto key_of = [&](int g) -> uint64_t {
    if (by_objective) {
      if ((double)a.costs.get(1, inst, g) != 0.0) return ~0ull;  // infeasible: never selected
      return orderable((double)a.costs.get(0, inst, g));
    }
    return orderable((double)a.costs.get(1, inst, g));
  };
  if (cached)
    for (int g = g0 + tid; g < g1; g += kSelThreads) s_all[g - g0] = key_of(g);
  auto key_at = [&](int g) -> uint64_t { return cached ? s_all[g - g0] : key_of(g); };
  if (tid == 0) {
    s_prefix = 0;
    s_remaining = E;
  }
  uint64_t mask = 0;
  for (int pass = 0; pass < 8; ++pass) {
    const int shift = 56 - 8 * pass;
    if (tid < 256) s_hist[tid] = 0;
    __syncthreads();
    const uint64_t prefix = s_prefix;
    int run_bin = -1, run_len = 0;
    for (int g = g0 + tid; g < g1; g += kSelThreads) {
      const uint64_t k = key_at(g);
      if ((k & mask) != prefix) continue;
      const int bin = (int)((k >> shift) & 255);
      if (bin != run_bin) {
        if (run_len) atomicAdd(&s_hist[run_bin], run_len);
        run_bin = bin, run_len = 0;
      }
      ++run_len;
    }
    if (run_len) atomicAdd(&s_hist[run_bin], run_len);
    cluster.sync();  // all eight histograms are complete
    if (tid < 256) {
      int t = 0;
      for (int r = 0; r < csize; ++r) t += cluster.map_shared_rank(s_hist, r)[tid];
      s_ghist[tid] = t;
    }
    __syncthreads();
    if (tid == 0) {  // the same scan on every CTA: identical decisions without a broadcast
      int rem = s_remaining, b = 0;
      for (; b < 255; ++b) {
        if (s_ghist[b] >= rem) break;
        rem -= s_ghist[b];
      }
      s_remaining = rem;
      s_prefix = prefix | ((uint64_t)b << shift);
    }
    mask |= 0xFFull << shift;
    cluster.sync();  // remote reads of s_hist are done before the next pass clears it (also publishes s_prefix locally)
  }
  const uint64_t vstar = s_prefix;
  const int take_equal = s_remaining;  // how many keys == v* belong to the elite set (lowest indices first)
  const int num_less = E - take_equal;

  // per-CTA totals -> offsets in index (= rank) order
  int c_less = 0, c_eq = 0;
  for (int g = g0 + tid; g < g1; g += kSelThreads) {
    const uint64_t k = key_at(g);
    c_less += k < vstar ? 1 : 0;
    c_eq += k == vstar ? 1 : 0;
  }
  c_less = block_sum_int(c_less, s_warp);
  c_eq = block_sum_int(c_eq, s_warp2);
  if (tid == 0) s_cnt[0] = c_less, s_cnt[1] = c_eq;
  cluster.sync();
  int base_less = 0, base_eq = 0;
  for (int r = 0; r < crank; ++r) {
    const int* rc = cluster.map_shared_rank(s_cnt, r);
    base_less += rc[0];
    base_eq += rc[1];
  }
  uint64_t* l_key = cluster.map_shared_rank(s_key, 0);
  int* l_idx = cluster.map_shared_rank(s_idx, 0);
  // ordered compaction of this CTA's range into the leader's arrays
  for (int gb = g0; gb < g1; gb += kSelThreads) {
    const int g = gb + tid;
    uint64_t k = ~0ull;
    bool f_less = false, f_eq = false;
    if (g < g1) {
      k = key_at(g);
      f_less = k < vstar;
      f_eq = k == vstar;
    }
    const unsigned b_less = __ballot_sync(0xffffffffu, f_less), b_eq = __ballot_sync(0xffffffffu, f_eq);
    __syncthreads();
    if (lane == 0) {
      s_warp[warp] = __popc(b_less);
      s_warp2[warp] = __popc(b_eq);
    }
    __syncthreads();
    int off_less = base_less, off_eq = base_eq, tot_less = 0, tot_eq = 0;
    for (int w = 0; w < kSelThreads / 32; ++w) {
      if (w < warp) {
        off_less += s_warp[w];
        off_eq += s_warp2[w];
      }
      tot_less += s_warp[w];
      tot_eq += s_warp2[w];
    }
    const unsigned lt = (1u << lane) - 1u;
    if (f_less) {
      const int slot = off_less + __popc(b_less & lt);
      l_key[slot] = k;
      l_idx[slot] = g;
    }
    if (f_eq) {
      const int slot = off_eq + __popc(b_eq & lt);
      if (slot < take_equal) {
        l_key[num_less + slot] = k;
        l_idx[num_less + slot] = g;
      }
    }
    base_less += tot_less;
    base_eq += tot_eq;
  }
  cluster.sync();  // every CTA's elites have landed in the leader's shared memory
  if (crank != 0) return;
  int P = 1;
  while (P < E) P <<= 1;
  for (int i = E + tid; i < P; i += kSelThreads) {
    s_key[i] = ~0ull;
    s_idx[i] = 0x7FFFFFFF;
  }
  sort_and_refit<T>(a, inst, s_key, s_idx, P);
}

// get_actions (constrained_cem_mpc.py:272-280): from the LAST iteration's rollouts return the first feasible one
// (quirk Q3: the reference's sort runs along a size-1 dimension), or the arg-min objective one when pick_best != 0.
template <typename T>
__global__ void __launch_bounds__(256)
cem_pick_kernel(CostView<T> costs, int n_global, int pick_best, ActionSrc<T> src, T* __restrict__ actions_out,
                int32_t* __restrict__ found, int32_t* __restrict__ best_id) {
  __shared__ unsigned long long s_best[8];
  __shared__ int s_bid[8];
  const int inst = blockIdx.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  uint64_t bk = ~0ull;
  int bi = 0x7FFFFFFF;
  for (int g = tid; g < n_global; g += blockDim.x) {
    if ((double)costs.get(1, inst, g) == 0.0) {
      const uint64_t k = pick_best ? orderable((double)costs.get(0, inst, g)) : 0ull;
      if (k < bk || (k == bk && g < bi)) bk = k, bi = g;
    }
  }
#pragma unroll
  for (int s = 16; s > 0; s >>= 1) {
    const uint64_t ok = __shfl_xor_sync(0xffffffffu, bk, s);
    const int oi = __shfl_xor_sync(0xffffffffu, bi, s);
    if (ok < bk || (ok == bk && oi < bi)) bk = ok, bi = oi;
  }
  if (lane == 0) s_best[warp] = bk, s_bid[warp] = bi;
  __syncthreads();
  for (int w = 0; w < (int)(blockDim.x >> 5); ++w) {
    const uint64_t ok = s_best[w];
    const int oi = s_bid[w];
    if (ok < bk || (ok == bk && oi < bi)) bk = ok, bi = oi;
  }
  const bool ok = bi != 0x7FFFFFFF;
  if (tid == 0) {
    found[inst] = ok ? 1 : 0;
    best_id[inst] = ok ? bi : -1;
  }
  for (int e = tid; e < src.tnu; e += blockDim.x)
    actions_out[(int64_t)inst * src.tnu + e] = ok ? candidate_action<T>(src, inst, bi, e) : T(0);
}

// local (objective, constraint) -> packed [2][M][N_local] send buffer (+ optional per-iteration records)
template <typename T>
__global__ void cem_pack_costs_kernel(const T* __restrict__ objective, const T* __restrict__ constraint,
                                      int64_t count, T* __restrict__ pack, T* __restrict__ rec_obj,
                                      T* __restrict__ rec_con) {
  const int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x;
  if (i >= count) return;
  const T o = objective[i], c = constraint[i];
  pack[i] = o;
  pack[count + i] = c;
  if (rec_obj) rec_obj[i] = o;
  if (rec_con) rec_con[i] = c;
}

static __global__ void cem_set_seed_kernel(uint64_t* dst, uint64_t seed) { *dst = seed; }

static __global__ void cem_init_dist_kernel(double* mean, double* stdv, int count) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < count) {
    mean[i] = 0.0;  // means = zeros, stds = ones   (constrained_cem_mpc.py:227-228)
    stdv[i] = 1.0;
  }
}

}  // namespace se
