// Internal launch interfaces shared by the translation units of libnmf_b200.so.
#pragma once
#include <cstddef>
#include <cstdint>
#include <cstdio>
#include <cuda.h>
#include <cuda_bf16.h>
#include <cuda_runtime.h>

namespace nmfb {

// Number of kernels this process has launched through the library (bench.py `gpu_launches`).
extern unsigned long long g_launch_count;
inline void count_launch(int n = 1) { g_launch_count += static_cast<unsigned long long>(n); }

// ------------------------------------------------------------------ panel GEMM (panel_gemm.cu)
struct alignas(64) PanelGemmArgs {
    CUtensorMap fac[3];      // factor planes, component-major [>=KP rows][ld], box {64, 64|128}
    CUtensorMap dat[2];      // data planes [rows][ld], box {64, BN}
    float* out;              // partial slots [slots][KP][Lp]
    int KP, NB, NA, BN;
    int npanels, nsplit, kb_per_split, kblocks_total;
    int Lp;
    int nstages;
};

struct PanelGemmPlan {
    int BN, nstages, npanels, nsplit, kb_per_split, kblocks_total, slots;
    int smem_bytes;
};

int panel_gemm_plan(int KP, int NB, int NA, int L, int R, int force_bn, int force_split, int num_sms,
                    PanelGemmPlan* plan);
cudaError_t panel_gemm_launch(const PanelGemmArgs& args, const PanelGemmPlan& plan, int num_sms,
                              cudaStream_t stream);

// ------------------------------------------------------------------ residual objective (residual.cu)
struct alignas(64) ResidualArgs {
    CUtensorMap wr[2];       // W row-major planes [Mp][KP], box {64, 128}
    CUtensorMap hr[2];       // H^T row-major planes [Np][KP], box {64, BN}
    CUtensorMap v[2];        // V planes [Mp][ldv], box {64, 128}
    double* partial;         // one slot per CTA
    int KP, NV, BN;
    int m, n;                // true extents (tail masking)
    int mtiles, ntiles;
    int kl;                  // 0: sum (v-s)^2 ; 1: generalised KL terms
};
cudaError_t residual_launch(const ResidualArgs& args, int grid, cudaStream_t stream);
int residual_smem_bytes(int KP, int NV, int BN);

}  // namespace nmfb
