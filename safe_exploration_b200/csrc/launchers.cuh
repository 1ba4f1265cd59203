// launchers.cuh -- host-side launch functions; the template-heavy kernels are instantiated in separate translation
// units (inst_*.cu) so the library builds in parallel.
#pragma once
#include "common.cuh"

namespace se {

template <typename T>
struct StepIO;
template <typename T>
struct ActionSrc;
template <typename T>
struct SelectArgs;
template <typename T>
struct CostView;

// K1: kernel rows + mean (+ Jacobian when jac != nullptr) for rows [row0, row0 + nrows) of the candidate batch
template <typename T>
int launch_kstar(const GpDev<T>& gp, const T* p, int64_t p_stride, const T* u, int64_t u_stride, int64_t row0,
                 int nrows, int rows_pad, T* kst, T* mean, T* jac, cudaStream_t stream);

// K2: partial row sums of squares of L^-1 k*; returns the number of column blocks through *nb_out
template <typename T>
int launch_trisumsq(se_gemm_impl impl, const T* kst, const T* w, int n_pad, int rows_pad, int ns, T* partial,
                    int* counter, int sm_count, int* nb_out, cudaStream_t stream);

template <typename T>
int launch_var(const GpDev<T>& gp, const T* partial, int nb, int rows_pad, int64_t row0, int nrows, T* var,
               const T* kpart, int nseg, T* mean, T* jac, cudaStream_t stream);

template <typename T>
int launch_reach_step(const ProbDev<T>& pr, const StepIO<T>& io, int64_t N, cudaStream_t stream);
template <typename T>
int launch_var_zero_scan(const T* var, int64_t count, int32_t* flags, cudaStream_t stream);
template <typename T>
struct RolloutIO;
template <typename T>
int launch_transpose_w(const T* w, int n_pad, int ns, T* wt, cudaStream_t stream);
template <typename T>
int launch_rollout_small(const GpDev<T>& gp, const T* wt, const ProbDev<T>& pr, const RolloutIO<T>& io,
                         cudaStream_t stream);

template <typename T>
int launch_safety_distance(const ProbDev<T>& pr, const T* p, const T* q, int64_t N, T* d, uint8_t* inside,
                           cudaStream_t stream);

template <typename T>
int launch_pack_model(const T* x, const T* alpha, const T* linv, int n, int d, int ns, int n_pad, T* xt, T* alpha_p,
                      T* w, cudaStream_t stream);

// CEM
template <typename T>
int launch_cem_sample(const ActionSrc<T>& src, int num_inst, int first, int n_local, int ns, const T* p0, T* actions,
                      T* p, T* rec_actions, cudaStream_t stream);
template <typename T>
int launch_cem_noise(uint64_t seed, int iteration, int num_inst, int n_global, int first, int count, int tnu, T* out,
                     cudaStream_t stream);
template <typename T>
int launch_cem_select_refit(const SelectArgs<T>& a, int num_inst, cudaStream_t stream);
template <typename T>
int launch_cem_pick(const CostView<T>& costs, int n_global, int pick_best, const ActionSrc<T>& src, int num_inst,
                    T* actions_out, int32_t* found, int32_t* best_id, cudaStream_t stream);
template <typename T>
int launch_cem_pack_costs(const T* objective, const T* constraint, int64_t count, T* pack, T* rec_obj, T* rec_con,
                          cudaStream_t stream);
int launch_cem_init_dist(double* mean, double* stdv, int count, cudaStream_t stream);
int launch_cem_set_seed(uint64_t* dst, uint64_t seed, cudaStream_t stream);

int measure_peak(int kind, double* flops_out);
int measure_peak_i8(int sustained, int n_mma, int issuers, double* ops_out);

// tcgen05 (kind::i8) digit-sliced variance GEMM (ozaki_kernels.cuh)
int launch_oz_pack_model(const double* w, int n_pad, int ns, const double* host_outputscale, int S, int* exp_a,
                         int* exp_b, double* scale_b, int8_t* b_img, cudaStream_t stream);
int launch_kstar_slices(const GpDev<double>& gp, const double* p, int64_t p_stride, const double* u, int64_t u_stride,
                        int64_t row0, int nrows, int rows_pad, const int* exp_a, int S, int8_t* a_img, double* mean,
                        double* jac, int seg_per_cta, double* kpart, cudaStream_t stream,
                        const ProbDev<double>* fuse_pr = nullptr, const StepIO<double>* fuse_io = nullptr);
int launch_kstar_slices_opt1(const GpDev<double>& gp, const double* p, int64_t p_stride, const double* u,
                             int64_t u_stride, int64_t row0, int nrows, int rows_pad, const int* exp_a, int S,
                             int8_t* a_img, double* mean, double* jac, int spc, double* kpart, unsigned grid_x,
                             unsigned grid_y, cudaStream_t stream);
int launch_kstar_slices_opt1_s6(const GpDev<double>& gp, const double* p, int64_t p_stride, const double* u,
                                int64_t u_stride, int64_t row0, int nrows, int rows_pad, const int* exp_a, int8_t* a_img,
                                double* mean, double* jac, int spc, double* kpart, unsigned grid_x, unsigned grid_y,
                                cudaStream_t stream);
int launch_trisumsq_ozaki(const int8_t* a_img, const int8_t* b_img, const double* scale_b, int n, int n_pad,
                          int rows_pad, int ns, int S, double* partial, int sm_count, int* nb_out, cudaStream_t stream);
struct TriOrder;
int launch_trisumsq_ozaki_cluster(int S, int shape, const int8_t* a_img, const int8_t* b_img, const double* scale_b,
                                  int n, int n_pad, int rows_pad, const TriOrder& od, int sm_count, double* partial,
                                  cudaStream_t stream);
int launch_oz_probe(const int8_t* a_img, const int8_t* b_img, int nk, int S, int lbo, int sbo, int32_t* out,
                    cudaStream_t stream);
int launch_debug_exp(const double* x, int64_t n, double* out, int table, cudaStream_t stream);

}  // namespace se

// (n_s, n_u) combinations with compiled register-resident kernels
#define SE_FOR_NS_NU(X) \
  X(1, 1) X(1, 2) X(1, 3) X(2, 1) X(2, 2) X(2, 3) X(3, 1) X(3, 2) X(3, 3) X(4, 1) X(4, 2) X(4, 3) X(5, 1) X(5, 2) \
      X(5, 3) X(6, 1) X(6, 2) X(6, 3)
