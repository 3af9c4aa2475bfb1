// Fused GEMM -> elementwise -> GEMM kernel of the Kullback-Leibler multiplicative updates
// (mur.py:24-27 W side, mur.py:40-43 H side); W.H is formed tile by tile in TMEM and never written.
//
// For one side, with data D (R x T; V for the W update, V^T for the H update), the factor being
// updated X (R x k) and the other factor Y (T x k):
//     S[r][t]   = sum_c X[r][c] Y[t][c]                    MMA1   (tile 128 r x 64 t, K = k)
//     Q[r][t]   = D[r][t] / (S[r][t] + 1e-9)               epilogue warps, registers -> shared memory
//     Out[c][r] = sum_t Y[t][c] Q[r][t]                    MMA2   (128 c x 128 r, K = 64 t)
// Out is the bracket of mur.py:25 / :41; the update itself (x * Out, the row / column sums and the
// regularised quotient) is the small kernel at the end of this file.
//
// The same pass can also emit the generalised KL objective of the pair (X, Y) it STARTS from
// (utils.py:23-26): S and D are in registers anyway, so the objective of iteration i costs no second
// sweep over V when it is taken from the W-side pass of iteration i + 1 (OBJ = true).
//
// The tcgen05 kernels live in kl_ts2.cu (KP = 128: X and Q as TMEM-resident A operands) and kl_ts.cu (KP = 64: X in
// TMEM, Q through shared memory); this file holds the plan shared by both, the dispatch, the row-sum kernel and the
// update kernel.  The SS-form kernel of round 1 (all operands from shared memory) lost the round-2 A/B and is gone.
#include "kernels.cuh"
#include "kl.cuh"
#include "ptx.cuh"

namespace nmfb {

namespace {
constexpr int kOutChunk = 8;           // t-tiles accumulated in TMEM before a drain (as in kl_ts.cu / kl_ts2.cu)
}  // namespace

int kl_plan(int KP, int NV, int R, int T, int num_sms, KlPlan* plan) {
    if (KP != 64 && KP != 128) return -1;            // k <= 128 in this build
    if (NV < 1 || NV > 2) return -1;
    const int npanels = (R + 127) / 128;
    const int ntiles = (T + 63) / 64;
    int nsplit = 1;
    double best = 1e30;
    for (int s = 1; s <= 64; ++s) {
        if (s > 1 && ntiles / s < 2 * kOutChunk) break;
        const long items = static_cast<long>(npanels) * s;
        const long waves = (items + num_sms - 1) / num_sms;
        const double cost = static_cast<double>(waves) * ((ntiles + s - 1) / s + 6);
        if (cost < best * 0.97) { best = cost; nsplit = s; }
    }
    int per = (ntiles + nsplit - 1) / nsplit;
    nsplit = (ntiles + per - 1) / per;
    plan->npanels = npanels;
    plan->ntiles = ntiles;
    plan->nsplit = nsplit;
    plan->tiles_per_split = per;
    plan->slots = nsplit;
    plan->smem_bytes = 0;                            // each kernel file knows its own layout
    const int nitems = npanels * nsplit;
    plan->grid = nitems < num_sms ? nitems : num_sms;
    return 0;
}

// KP = 128: X and the quotient Q both as TMEM-resident A operands (kl_ts2.cu); KP = 64: X only (kl_ts.cu).
// Round-2 A/B at 20000 x 50000, k = 128 (profiles/r02a_*): SS-form kernel 255.6 it/s (deleted), X in TMEM 265.5,
// X and Q in TMEM 304.2.
cudaError_t kl_panel_launch(const KlArgs& args, const KlPlan& plan, int NV, int num_sms, cudaStream_t stream) {
    (void)num_sms;
    if (args.KP == 128) return kl_panel_ts2_launch(args, plan, NV, stream);
    if (args.KP == 64) return kl_panel_ts_launch(args, plan, NV, stream);
    return cudaErrorInvalidValue;
}

// ------------------------------------------------------------------ row sums of a factor
// One block per component: sum of master[c][0..L) in fp64 (warp shuffles + one smem hop).
__global__ void __launch_bounds__(256)
factor_rowsum_kernel(const float* __restrict__ master, int L, int Lp, double* __restrict__ sums) {
    __shared__ double red[8];
    const float* rowp = master + static_cast<long>(blockIdx.x) * Lp;
    double t = 0.0;
    for (int i = threadIdx.x * 4; i < L; i += 256 * 4) {
        const float4 v = *reinterpret_cast<const float4*>(rowp + i);       // pads beyond L are zero
        t += static_cast<double>(v.x) + static_cast<double>(v.y) + static_cast<double>(v.z) + static_cast<double>(v.w);
    }
    for (int o = 16; o; o >>= 1) t += __shfl_xor_sync(0xffffffffu, t, o);
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = t;
    __syncthreads();
    if (threadIdx.x == 0) {
        double s = 0.0;
        for (int w = 0; w < 8; ++w) s += red[w];
        sums[blockIdx.x] = s;
    }
}

cudaError_t factor_rowsum_launch(const float* master, int KP, int L, int Lp, double* sums, cudaStream_t stream) {
    factor_rowsum_kernel<<<KP, 256, 0, stream>>>(master, L, Lp, sums);
    count_launch();
    return cudaGetLastError();
}

// ------------------------------------------------------------------ KL multiplicative update
// One block owns 64 consecutive r and all KP components (same thread tile as the Euclidean update,
// so the plane stores are identical).
template <int CPT>
__global__ void __launch_bounds__(256)
mur_kl_update_kernel(const float* __restrict__ part, int slots, const double* __restrict__ bsum,
                     float* __restrict__ master, __nv_bfloat16* __restrict__ cpl, long cpl_plane,
                     __nv_bfloat16* __restrict__ rpl, long rpl_plane, int k, int L, int Lp, float lambda) {
    constexpr int KP = CPT * 16;
    const int tid = threadIdx.x;
    const int tr = tid & 15, tc = tid >> 4;
    const int r0 = blockIdx.x * 64;
    __align__(16) __nv_bfloat16 rh[4][CPT], rl[4][CPT];
#pragma unroll
    for (int a = 0; a < CPT; ++a) {
        const int c = tc * CPT + a;
        const long off = static_cast<long>(c) * Lp + r0 + tr * 4;
        float4 num = make_float4(0.f, 0.f, 0.f, 0.f);
        for (int s = 0; s < slots; ++s) {
            const float4 p = *reinterpret_cast<const float4*>(part + static_cast<long>(s) * KP * Lp + off);
            num.x += p.x; num.y += p.y; num.z += p.z; num.w += p.w;
        }
        const float4 x = *reinterpret_cast<const float4*>(master + off);
        const float b = static_cast<float>(bsum[c]);
        const float xv[4] = {x.x, x.y, x.z, x.w}, nv[4] = {num.x, num.y, num.z, num.w};
        float xn[4];
        __align__(8) __nv_bfloat16 ch[4], cl[4];
#pragma unroll
        for (int q = 0; q < 4; ++q) {
            const float av = xv[q] * nv[q];                               // a = w * (...)      mur.py:25
            xn[q] = 2.f * av / (b + sqrtf(b * b + 4.f * lambda * av));    //                    mur.py:27
            if (c >= k || r0 + tr * 4 + q >= L) xn[q] = 0.f;
            split2(xn[q], ch[q], cl[q]);
            rh[q][a] = ch[q];
            rl[q][a] = cl[q];
        }
        *reinterpret_cast<float4*>(master + off) = make_float4(xn[0], xn[1], xn[2], xn[3]);
        *reinterpret_cast<uint2*>(cpl + off) = *reinterpret_cast<const uint2*>(ch);
        *reinterpret_cast<uint2*>(cpl + cpl_plane + off) = *reinterpret_cast<const uint2*>(cl);
    }
#pragma unroll
    for (int q = 0; q < 4; ++q) {
        const long ro = static_cast<long>(r0 + tr * 4 + q) * KP + tc * CPT;
#pragma unroll
        for (int a = 0; a < CPT; a += 4) {
            *reinterpret_cast<uint2*>(rpl + ro + a) = *reinterpret_cast<const uint2*>(&rh[q][a]);
            *reinterpret_cast<uint2*>(rpl + rpl_plane + ro + a) = *reinterpret_cast<const uint2*>(&rl[q][a]);
        }
    }
}

cudaError_t mur_kl_update_launch(const float* part, int slots, const double* bsum, float* master,
                                 __nv_bfloat16* cpl, long cpl_plane, __nv_bfloat16* rpl, long rpl_plane,
                                 int KP, int k, int L, int Lp, int nrows_p, float lambda, cudaStream_t stream) {
    const int grid = nrows_p / 64;
    if (KP == 64)
        mur_kl_update_kernel<4><<<grid, 256, 0, stream>>>(part, slots, bsum, master, cpl, cpl_plane, rpl, rpl_plane, k, L, Lp, lambda);
    else if (KP == 128)
        mur_kl_update_kernel<8><<<grid, 256, 0, stream>>>(part, slots, bsum, master, cpl, cpl_plane, rpl, rpl_plane, k, L, Lp, lambda);
    else
        mur_kl_update_kernel<16><<<grid, 256, 0, stream>>>(part, slots, bsum, master, cpl, cpl_plane, rpl, rpl_plane, k, L, Lp, lambda);
    count_launch();
    return cudaGetLastError();
}

}  // namespace nmfb
