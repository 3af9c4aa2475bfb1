// Launch interface of the Kullback-Leibler kernels (kl.cu).
#pragma once
#include "kernels.cuh"

namespace nmfb {

struct alignas(64) KlArgs {
    CUtensorMap xr[2];       // rplanes of the factor being updated [Lp][KP], box {64, 128}
    CUtensorMap yc[2];       // cplanes of the other factor [256][Tp], box {64, 128}
    CUtensorMap d[2];        // K-blocked data planes, box {64, 128}
    float* out;              // partial slots [nsplit][KP][Lp]
    double* obj_partial;     // optional: one fp64 slot per CTA for the KL objective of the incoming pair
    int KP, Lp;
    int npanels, ntiles, nsplit, tiles_per_split;
    int d_rows_p;
    const __nv_bfloat16* x_rpl[2];   // the row planes of X themselves (same base as xr[])
};

struct KlPlan {
    int npanels, ntiles, nsplit, tiles_per_split, slots, smem_bytes, grid;
};

int kl_plan(int KP, int NV, int R, int T, int num_sms, KlPlan* plan);
cudaError_t kl_panel_launch(const KlArgs& args, const KlPlan& plan, int NV, int num_sms, cudaStream_t stream);
// kl_ts.cu (KP = 64): X as the TMEM-resident A operand of MMA1
cudaError_t kl_panel_ts_launch(const KlArgs& args, const KlPlan& plan, int NV, cudaStream_t stream);
// kl_ts2.cu (KP = 128): the quotient Q too, stored over S, as the A operand of MMA2
cudaError_t kl_panel_ts2_launch(const KlArgs& args, const KlPlan& plan, int NV, cudaStream_t stream);

// sums[c] = sum over the first L entries of master row c (fp64): `np.ones_like(x) @ h.T` (mur.py:26)
// and `w.T @ np.ones_like(x)` (mur.py:42) collapse to these k-vectors.
cudaError_t factor_rowsum_launch(const float* master, int KP, int L, int Lp, double* sums, cudaStream_t stream);

// x <- 2a / (b + sqrt(b^2 + 4 lambda a)),  a = x * sum_slots(part),  b = bsum[c]     (mur.py:25-27, 41-43)
cudaError_t mur_kl_update_launch(const float* part, int slots, const double* bsum, float* master,
                                 __nv_bfloat16* cpl, long cpl_plane, __nv_bfloat16* rpl, long rpl_plane,
                                 int KP, int k, int L, int Lp, int nrows_p, float lambda, cudaStream_t stream);

}  // namespace nmfb
