// Launchers of the CUDA-core kernels (elementwise.cu, admm.cu).
#pragma once
#include "kernels.cuh"

namespace nmfb {

cudaError_t ingest_data_launch(const void* src, int is_f64, long src_ld, int row0, int nrows, int ncols,
                               __nv_bfloat16* rowp, long rowp_plane, int rowp_rows_p, __nv_bfloat16* colp,
                               long colp_plane, int colp_rows_p, int nv, unsigned int* stats,
                               const unsigned int* rowmax, const unsigned int* colmax, int row_lo, int row_n,
                               int col_lo, int col_n, cudaStream_t stream);
cudaError_t data_absmax_launch(const void* src, int is_f64, long src_ld, int row0, int nrows, int ncols,
                               unsigned int* rowmax, unsigned int* colmax, cudaStream_t stream);
cudaError_t fixed_planes_launch(const float* master, int KP, int Lp, __nv_bfloat16* cpl, long cpl_plane,
                                cudaStream_t stream);
cudaError_t refresh_planes_launch(const float* master, float* adopt, int KP, int Lp, __nv_bfloat16* cpl,
                                  long cpl_plane, __nv_bfloat16* rpl, long rpl_plane, int skip_lo, int skip_hi,
                                  cudaStream_t stream);
cudaError_t row_planes3_launch(const float* master, int KP, int Lp, __nv_bfloat16* out, cudaStream_t stream);
cudaError_t zero_words_guarded_launch(unsigned int* a, int na, unsigned int* b, int nb, const int* done_flag,
                                      cudaStream_t stream);
cudaError_t import_factor_launch(const void* src, int is_f64, long stride_l, long stride_c, int L, int k,
                                 float* master, int KP, int Lp, cudaStream_t stream);
cudaError_t export_factor_launch(const float* master, int KP, int Lp, void* dst, int is_f64, long stride_l,
                                 long stride_c, int L, int k, cudaStream_t stream);
cudaError_t gram_reduce_launch(const float* part, int slots, int KP, int Lp, float* g32, double* g64,
                               cudaStream_t stream);
cudaError_t mur_eu_update_launch(const float* part, int slots, const float* gram, float* master,
                                 __nv_bfloat16* cpl, long cpl_plane, __nv_bfloat16* rpl, long rpl_plane,
                                 int KP, int k, int L, int Lp, int nrows_p, float lambda, cudaStream_t stream);
cudaError_t objective_simt_launch(const __nv_bfloat16* vpl, long v_plane, int v_rows_p, int nv, const float* wt,
                                  int Mp, const float* ht, int Np, int k, int m, int n, double* partial,
                                  int* nparts, cudaStream_t stream);
cudaError_t sum_partials_launch(const double* partial, int n, double scale, double* out,
                                cudaStream_t stream);
cudaError_t kl_vaux_launch(const float* prod, float* d_out, float* dual_v, long ld, int rows, int cols,
                           const __nv_bfloat16* vpl, long v_plane, int v_rows_p, int nv, const int* done_flag,
                           cudaStream_t stream);

}  // namespace nmfb
