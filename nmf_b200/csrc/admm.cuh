// Launchers of the ADMM / AO-ADMM kernels (admm.cu).
#pragma once
#include "kernels.cuh"

namespace nmfb {

// G = F F^T in fp64 of the factor held by three bf16 cplanes [3][256][Lp]; scratch >= gram_fp64_scratch_bytes(KP, Lp)
cudaError_t gram_fp64_launch(const __nv_bfloat16* cpl, long plane, int KP, int Lp, double* scratch, float* g32,
                             double* g64, cudaStream_t stream);
size_t gram_fp64_scratch_bytes(int KP, int Lp);

cudaError_t chol_inverse_launch(const double* gram, int KP, int k, int rho_mode, double rho_in,
                                double* gscratch, double* ainv, double* rho_out, int* flags,
                                cudaStream_t stream);

struct AdmmApplyArgs {
    const float* part;        // numerator partial slots [slots][KP][Lp] (fp32) or nullptr (use c64)
    int slots;
    double* c64;              // fp64 [KP][Lp] reduced numerator: written when part != nullptr, read otherwise
    const double* ainv;       // fp64 [KP][KP]
    const double* rho;        // device scalar
    float lambda;
    int prox;                 // 0 = nn, 1 = l1n, 2 = l2n (binv != nullptr)
    float* x;                 // fp32 master [KP][Lp] of the constrained factor (in/out)
    float* u;                 // fp32 dual [KP][Lp] (in/out)
    float* aux_master;        // optional: auxiliary master (ADMM)
    __nv_bfloat16* aux_cpl;   // optional: auxiliary cplanes, 3 planes (ADMM)
    long aux_cpl_plane;
    __nv_bfloat16* x_cpl;     // optional: cplanes of x, 3 planes (AO-ADMM: x is the next GEMM operand)
    long x_cpl_plane;
    __nv_bfloat16* x_rpl;     // optional: rplanes of x, 2 planes (objective operands)
    long x_rpl_plane;
    double* norm_partial;     // optional: [grid][4]  |x-aux|^2, |x|^2, |x-x_prev|^2, |u|^2
    const int* done_flag;     // optional: skip the launch's work when *done_flag != 0
    int KP, k, L, Lp;         // L = live rows of the window being updated, Lp = pitch of the factor
    int nrows_p;              // padded rows of the window (grid = nrows_p / 64); 0 = Lp
    const double* binv;       // fp64 [KP][KP] inverse Tikhonov system (prox l2n) or nullptr; LAST on purpose:
                              // a field in the middle moved the hot kernel from 128 to 132 registers (one CTA/SM)
};
cudaError_t admm_apply_launch(const AdmmApplyArgs& a, cudaStream_t stream);
// out[KP][KP] = (lambda / *rho) * T^T T, T = tridiag(-1, 2, -1) of order k (admm.py:142-146)
cudaError_t tikhonov_gram_launch(int k, int KP, double lambda, const double* rho, double* out, cudaStream_t stream);

// Reduce norm partials, raise flags[1] (= inner loop done) and count the executed inner iterations
// in flags[used_idx]; no-op when already done.
cudaError_t admm_inner_flag_launch(const double* norm_partial, int nblocks, int* flags, int used_idx,
                                   cudaStream_t stream);
// flags bookkeeping between factor updates / outer iterations.
cudaError_t admm_flags_reset_launch(int* flags, int mask, cudaStream_t stream);
// mailbox[1..3] = (double) flags[2], flags[3], flags[0]
cudaError_t admm_flags_publish_launch(const int* flags, double* mailbox, cudaStream_t stream);

}  // namespace nmfb
