// Internal launch interfaces shared by the translation units of libnmf_b200.so.
#pragma once
#include <cstddef>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cuda.h>
#include <cuda_bf16.h>
#include <cuda_runtime.h>

namespace nmfb {

// Number of kernels this process has launched through the library (bench.py `gpu_launches`).
extern unsigned long long g_launch_count;
inline void count_launch(int n = 1) { g_launch_count += static_cast<unsigned long long>(n); }

// Data planes (V, V^T) are stored K-blocked: [col / 64][rows_p][64], so that the [rows x 64] box a
// TMA load fetches for one contraction block is ONE contiguous run of HBM (rows * 128 bytes).
__host__ __device__ inline long blocked_index(int row, int col, int rows_p) {
    return (static_cast<long>(col >> 6) * rows_p + row) * 64 + (col & 63);
}

// ------------------------------------------------------------------ panel GEMM (panel_gemm.cu)
struct alignas(64) PanelGemmArgs {
    CUtensorMap fac[3];      // factor planes, component-major [>=KP rows][ld], box {64, 64|128}
    CUtensorMap dat[3];      // data planes [rows][ld], box {64, BN}
    float* out;              // partial slots [slots][KP][Lp]
    int KP, NB, NA, BN;
    int npanels, nsplit, kb_per_split, kblocks_total;
    int Lp;
    int nstages;
    int dat_rows_p;          // > 0: data planes are K-blocked with this many rows per block; 0: row-major
    int cluster, ngroups;    // CTAs per cluster; panel groups (= ceil(npanels / cluster))
    int chunk_kb;            // contraction blocks per TMEM accumulation chunk
    int dat_box_rows;        // rows per TMA box of the data planes (BN is fetched as BN / dat_box_rows boxes)
    int prefetch_kb;         // L2 prefetch distance of the data stream, in contraction blocks (0 = off)
};

struct PanelGemmPlan {
    int BN, nstages, npanels, nsplit, kb_per_split, kblocks_total, slots;
    int smem_bytes;
    int cluster, ngroups, chunk_kb, max_clusters, dat_box_rows, prefetch_kb;
};

int panel_gemm_plan(int KP, int NB, int NA, int L, int R, int force_bn, int force_split, int num_sms,
                    PanelGemmPlan* plan);
cudaError_t panel_gemm_launch(const PanelGemmArgs& args, const PanelGemmPlan& plan, int num_sms,
                              cudaStream_t stream);
// exact-hi variant (panel_gemm_exact.cu): NB = 3 fixed-point factor planes, NA >= 2 fixed-point data planes
int panel_gemm_exact_plan(int KP, int NA, int L, int R, int blocked, int num_sms, PanelGemmPlan* plan);
cudaError_t panel_gemm_exact_launch(const PanelGemmArgs& args, const PanelGemmPlan& plan, cudaStream_t stream);

// ------------------------------------------------------------------ residual objective (residual_n256.cu, residual_ts.cu)
struct alignas(64) ResidualArgs {
    CUtensorMap wr[2];       // W row-major planes [Mp][KP], box {64, 128}
    CUtensorMap hr[2];       // H^T row-major planes [Np][KP], box {64, BN}
    CUtensorMap v[2];        // V planes [Mp][ldv], box {64, 128}
    double* partial;         // one slot per CTA
    int KP, NV, BN;
    int m, n;                // true extents (tail masking)
    int mtiles, ntiles;
    int kl;                  // 0: sum (v-s)^2 ; 1: generalised KL terms
    int v_rows_p;            // rows per K-block of the blocked V planes
    int cluster;             // CTAs per cluster (H^T stages are multicast across it)
    const __nv_bfloat16* w_rpl[2];   // TS form only: the W row planes [rows][KP] themselves (same base as wr[])
};
int residual_max_clusters(int cs, int num_sms);
// wide tiles (residual_n256.cu): 128 x 256 tiles, args.BN == 256, args.ntiles in 256-column tiles; KP <= 128, one V plane
cudaError_t residual_n256_launch(const ResidualArgs& args, int grid, cudaStream_t stream);
int residual_n256_smem_bytes(int KP, int NV);
// TS form (residual_ts.cu): W planes of the tile row as the TMEM-resident A operand, args.BN == 128; every KP, 1-2 V planes
cudaError_t residual_ts_launch(const ResidualArgs& args, int grid, cudaStream_t stream);
int residual_ts_smem_bytes(int KP, int NV);

// ------------------------------------------------------------------ KL data step of ADMM / AO-ADMM (kl_data.cu)
struct alignas(64) KlDataArgs {
    CUtensorMap br[3];       // B operand: three row-major planes [Np][KP] of the right factor, box {64, 128 / cluster}
    CUtensorMap v[2];        // V planes (K-blocked), box {64, 128}
    const __nv_bfloat16* a_rpl[3];   // A operand: three row-major planes [rows][KP] of the left factor (this rank's rows)
    float* dual_v;           // K-blocked fp32 [ld / 64][rows_p][64], updated in place
    float* d_out;            // K-blocked fp32: D = v_aux + dual_v
    long rows_p;
    unsigned int* rowmax;    // |max| bits of D per row (index row_origin + local row) and per column
    unsigned int* colmax;
    const int* done_flag;    // AO-ADMM: != 0 -> the kernel does nothing
    int row_origin;
    int KP, NV;
    int mtiles, ntiles;
    int v_rows_p;
    int cluster;
};
cudaError_t kl_data_launch(const KlDataArgs& args, int grid, cudaStream_t stream);
int kl_data_smem_bytes(int KP, int NV);

}  // namespace nmfb
