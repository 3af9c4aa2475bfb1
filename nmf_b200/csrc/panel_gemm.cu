// Split-precision skinny GEMM for the NMF contractions (V.H^T, W^T.V, Grams) on tcgen05.
//
//   P^T[c][r] = sum_t  F_j[c][t] * D_i[r][t]      (all allowed plane pairs (i,j) summed)
//
// D = "data" planes (V or V^T, bf16, row-major [rows][ld], contraction index t contiguous),
// F = "factor" planes (the other factor, bf16, component-major [KP][ld], t contiguous).
// fp32-grade products come from bf16 hi/lo(/lo2) planes: every MMA is an exact bf16 x bf16
// product accumulated in fp32 in TMEM.  The factor is the UMMA *A* operand (M = 128 components,
// or two stacked 64-row planes when KP = 64) and a panel of BN data rows is the *B* operand
// (N = BN), so one shared-memory stage of the big operand feeds every plane MMA.
//
// Persistent, warp-specialised: warp 0 = TMA producer, warp 1 = MMA issuer (+TMEM owner),
// warps 2..5 = epilogue (TMEM -> registers -> global partial sums).  Work items are
// (K-split, row-panel) pairs walked round-robin; each item writes its own fp32 partial slot
// (no atomics => bit-reproducible), which the update kernels reduce in a fixed order.
//
// Reference call sites replaced: mur.py:29 `x @ h.T`, mur.py:45 `w.T @ x`, admm.py:221
// `other_aux.T @ data_aux`, ao_admm.py:53,56 `w.T @ w`, `w.T @ y`.
#include "kernels.cuh"
#include "ptx.cuh"

namespace nmfb {

namespace {

constexpr int kBK = 64;                 // contraction elements per stage (= one 128 B swizzle span)
constexpr int kATileBytes = 128 * 128;  // one 128-row A tile
constexpr int kThreads = 192;

struct StageLayout {
    int a_tiles;       // 128-row A tiles per stage
    int a_bytes;       // a_tiles * 16 KB
    int b_plane_bytes; // BN * 128
    int stage_bytes;
};

__host__ __device__ inline StageLayout stage_layout(int KP, int NB, int NA, int BN) {
    StageLayout s;
    s.a_tiles = (KP == 64) ? (NB + 1) / 2 : (KP == 128 ? NB : 2 * NB);
    s.a_bytes = s.a_tiles * kATileBytes;
    s.b_plane_bytes = BN * 128;
    s.stage_bytes = s.a_bytes + NA * s.b_plane_bytes;
    return s;
}

}  // namespace

__global__ void __launch_bounds__(kThreads, 1)
panel_gemm_kernel(const __grid_constant__ PanelGemmArgs args) {
    extern __shared__ __align__(1024) uint8_t smem_raw[];
    // 1024-byte alignment is required by SWIZZLE_128B; the dynamic smem base is only
    // guaranteed 16 B aligned, so align by hand (host adds 1 KB of slack).
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) &
                                               ~static_cast<uintptr_t>(1023));
    const int KP = args.KP, NB = args.NB, NA = args.NA, BN = args.BN;
    const StageLayout sl = stage_layout(KP, NB, NA, BN);
    const int nstages = args.nstages;
    uint8_t* stage_base = smem;
    uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + static_cast<size_t>(nstages) * sl.stage_bytes);
    uint64_t* empty_bar = full_bar + nstages;
    uint64_t* acc_full = empty_bar + nstages;     // [2]
    uint64_t* acc_empty = acc_full + 2;           // [2]
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(acc_empty + 2);

    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;
    const int acc_cols = BN * (KP == 256 ? 2 : 1);      // TMEM columns per accumulator stage
    const int nacc = (2 * acc_cols <= 512) ? 2 : 1;

    const int nitems = args.npanels * args.nsplit;
    const int kb_per = args.kb_per_split;

    // KP == 64 with an odd plane count leaves the upper half of the last A tile unused:
    // zero it once (TMA never writes there), so the stacked MMA adds exact zeros.
    if (KP == 64 && (NB & 1)) {
        for (int s = 0; s < nstages; ++s) {
            uint4* z = reinterpret_cast<uint4*>(stage_base + static_cast<size_t>(s) * sl.stage_bytes +
                                                (sl.a_tiles - 1) * kATileBytes + 8192);
            for (int i = threadIdx.x; i < 8192 / 16; i += kThreads) z[i] = make_uint4(0, 0, 0, 0);
        }
        fence_async_smem();
    }

    if (warp == 0 && lane == 0) {
        for (int j = 0; j < NB; ++j) tma_prefetch_desc(&args.fac[j]);
        for (int i = 0; i < NA; ++i) tma_prefetch_desc(&args.dat[i]);
        for (int s = 0; s < nstages; ++s) {
            mbar_init(&full_bar[s], 1);
            mbar_init(&empty_bar[s], 1);
        }
        for (int a = 0; a < 2; ++a) {
            mbar_init(&acc_full[a], 1);
            mbar_init(&acc_empty[a], 4);      // one arrive per epilogue warp
        }
        fence_mbar_init();
    }
    if (warp == 1) {
        tmem_alloc(tmem_slot, 512);
        tmem_relinquish();
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    if (warp == 0) {
        // ------------------------------------------------------------ TMA producer
        if (lane == 0) {
            const int a_box_rows = (KP == 64) ? 64 : 128;
            const int a_box_bytes = a_box_rows * 128;
            uint32_t tx_bytes = NA * sl.b_plane_bytes;
            // number of A boxes per stage
            int n_a_boxes = (KP == 64) ? NB : sl.a_tiles;
            tx_bytes += n_a_boxes * a_box_bytes;
            int stage = 0;
            uint32_t phase = 0;
            for (int item = blockIdx.x; item < nitems; item += gridDim.x) {
                const int split = item / args.npanels;
                const int panel = item - split * args.npanels;
                const int kb0 = split * kb_per;
                const int kb1 = min(kb0 + kb_per, args.kblocks_total);
                for (int kb = kb0; kb < kb1; ++kb) {
                    mbar_wait(&empty_bar[stage], phase ^ 1);
                    uint8_t* st = stage_base + static_cast<size_t>(stage) * sl.stage_bytes;
                    mbar_expect_tx(&full_bar[stage], tx_bytes);
                    const int t0 = kb * kBK;
                    if (KP == 64) {
                        for (int j = 0; j < NB; ++j)
                            tma_load_2d_hint(st + (j >> 1) * kATileBytes + (j & 1) * 8192, &args.fac[j],
                                             &full_bar[stage], t0, 0, kPolicyEvictLast);
                    } else if (KP == 128) {
                        for (int j = 0; j < NB; ++j)
                            tma_load_2d_hint(st + j * kATileBytes, &args.fac[j], &full_bar[stage], t0, 0,
                                             kPolicyEvictLast);
                    } else {
                        for (int t = 0; t < sl.a_tiles; ++t)
                            tma_load_2d_hint(st + t * kATileBytes, &args.fac[t >> 1], &full_bar[stage], t0,
                                             (t & 1) * 128, kPolicyEvictLast);
                    }
                    for (int i = 0; i < NA; ++i)
                        tma_load_2d_hint(st + sl.a_bytes + i * sl.b_plane_bytes, &args.dat[i],
                                         &full_bar[stage], t0, panel * BN, kPolicyEvictFirst);
                    if (++stage == nstages) { stage = 0; phase ^= 1; }
                }
            }
        }
    } else if (warp == 1) {
        // ------------------------------------------------------------ MMA issuer
        if (lane == 0) {
            const uint32_t idesc = umma_idesc_bf16(128, BN);
            int stage = 0;
            uint32_t phase = 0;
            int acc = 0;
            uint32_t acc_phase = 0;
            for (int item = blockIdx.x; item < nitems; item += gridDim.x) {
                const int split = item / args.npanels;
                const int kb0 = split * kb_per;
                const int kb1 = min(kb0 + kb_per, args.kblocks_total);
                mbar_wait(&acc_empty[acc], acc_phase ^ 1);
                tc_fence_after();
                const uint32_t d_base = tmem_base + acc * acc_cols;
                uint32_t started0 = 0, started1 = 0;       // per accumulator block: has it been written?
                for (int kb = kb0; kb < kb1; ++kb) {
                    mbar_wait(&full_bar[stage], phase);
                    tc_fence_after();
                    const uint32_t st = smem_u32(stage_base + static_cast<size_t>(stage) * sl.stage_bytes);
#pragma unroll
                    for (int ks = 0; ks < kBK / 16; ++ks) {
                        for (int t = 0; t < sl.a_tiles; ++t) {
                            const int plane = (KP == 64) ? 0 : (KP == 128 ? t : (t >> 1));
                            const int rb = (KP == 256) ? (t & 1) : 0;
                            const uint64_t da = umma_desc_k_sw128(st + t * kATileBytes + ks * 32);
                            for (int i = 0; i < NA; ++i) {
                                // plane pair filter: drop products below the lowest kept plane
                                if (KP != 64 && (i + plane > NB - 1)) continue;
                                const uint64_t db =
                                    umma_desc_k_sw128(st + sl.a_bytes + i * sl.b_plane_bytes + ks * 32);
                                uint32_t& started = rb ? started1 : started0;
                                umma_bf16(d_base + rb * BN, da, db, idesc, started);
                                started = 1;
                            }
                        }
                    }
                    umma_commit(&empty_bar[stage]);        // frees the smem stage when the MMAs retire
                    if (++stage == nstages) { stage = 0; phase ^= 1; }
                }
                umma_commit(&acc_full[acc]);               // accumulator complete -> epilogue
                if (++acc == nacc) { acc = 0; acc_phase ^= 1; }
            }
        }
    } else {
        // ------------------------------------------------------------ epilogue (warps 2..5)
        const int q = warp & 3;                              // TMEM lane quarter this warp may read
        int acc = 0;
        uint32_t acc_phase = 0;
        const int sps = (KP == 64) ? 2 : 1;                  // partial slots per split
        for (int item = blockIdx.x; item < nitems; item += gridDim.x) {
            const int split = item / args.npanels;
            const int panel = item - split * args.npanels;
            mbar_wait(&acc_full[acc], acc_phase);
            tc_fence_after();
            const int l = q * 32 + lane;                     // accumulator lane == stacked component row
            const int nblk = (KP == 256) ? 2 : 1;
            for (int rb = 0; rb < nblk; ++rb) {
                int c, slot;
                if (KP == 64) { c = l & 63; slot = split * sps + (l >> 6); }
                else          { c = rb * 128 + l; slot = split; }
                float* dst = args.out + (static_cast<size_t>(slot) * KP + c) * args.Lp +
                             static_cast<size_t>(panel) * BN;
                const uint32_t taddr = tmem_base + (static_cast<uint32_t>(q * 32) << 16) +
                                       acc * acc_cols + rb * BN;
                for (int col = 0; col < BN; col += 32) {
                    uint32_t r[32];
                    tmem_ld32(taddr + col, r);
                    tmem_ld_wait();
                    float4* d4 = reinterpret_cast<float4*>(dst + col);
#pragma unroll
                    for (int v = 0; v < 8; ++v)
                        d4[v] = make_float4(__uint_as_float(r[4 * v]), __uint_as_float(r[4 * v + 1]),
                                            __uint_as_float(r[4 * v + 2]), __uint_as_float(r[4 * v + 3]));
                }
            }
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(&acc_empty[acc]);
            if (++acc == nacc) { acc = 0; acc_phase ^= 1; }
        }
    }

    tc_fence_before();
    __syncthreads();
    if (warp == 1) {
        tc_fence_after();
        tmem_dealloc(tmem_base, 512);
    }
}

// ---------------------------------------------------------------------------------------------
// host side
// ---------------------------------------------------------------------------------------------
int panel_gemm_plan(int KP, int NB, int NA, int L, int R, int force_bn, int force_split, int num_sms,
                    PanelGemmPlan* plan) {
    if (!(KP == 64 || KP == 128 || KP == 256) || NB < 1 || NB > 3 || NA < 1 || NA > 2) return -1;
    const int smem_budget = 227 * 1024 - 1024 /*align slack*/ - 256 /*barriers*/;
    int BN = force_bn ? force_bn : 256;
    StageLayout sl = stage_layout(KP, NB, NA, BN);
    if (smem_budget / sl.stage_bytes < 2 && !force_bn) {
        BN = 128;
        sl = stage_layout(KP, NB, NA, BN);
    }
    int nstages = smem_budget / sl.stage_bytes;
    if (nstages < 2) return -2;
    if (nstages > 8) nstages = 8;
    const int npanels = (L + BN - 1) / BN;
    const int kblocks = (R + kBK - 1) / kBK;
    // K-split: fill the machine (items ~ multiple of #SMs) while keeping >= 16 k-blocks per item.
    int nsplit = 1;
    if (force_split > 0) {
        nsplit = force_split;
    } else {
        double best = 1e30;
        const int max_split = 64;
        for (int s = 1; s <= max_split; ++s) {
            if (s > 1 && kblocks / s < 16) break;
            const long items = static_cast<long>(npanels) * s;
            const long waves = (items + num_sms - 1) / num_sms;
            // cost ~ waves * k-blocks per item  (+ epilogue/partial traffic penalty per split)
            const double cost = static_cast<double>(waves) * ((kblocks + s - 1) / s + 8);
            if (cost < best * 0.97) { best = cost; nsplit = s; }
        }
    }
    if (nsplit > kblocks) nsplit = kblocks;
    int kb_per = (kblocks + nsplit - 1) / nsplit;
    nsplit = (kblocks + kb_per - 1) / kb_per;            // no empty splits
    plan->BN = BN;
    plan->nstages = nstages;
    plan->npanels = npanels;
    plan->nsplit = nsplit;
    plan->kb_per_split = kb_per;
    plan->kblocks_total = kblocks;
    plan->slots = nsplit * (KP == 64 ? 2 : 1);
    plan->smem_bytes = nstages * sl.stage_bytes + 1024 + 256;
    return 0;
}

cudaError_t panel_gemm_launch(const PanelGemmArgs& args, const PanelGemmPlan& plan, int num_sms,
                              cudaStream_t stream) {
    static bool attr_set = false;
    if (!attr_set) {
        cudaError_t e = cudaFuncSetAttribute(panel_gemm_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                             227 * 1024);
        if (e != cudaSuccess) return e;
        attr_set = true;
    }
    const int nitems = plan.npanels * plan.nsplit;
    const int grid = nitems < num_sms ? nitems : num_sms;
    panel_gemm_kernel<<<grid, kThreads, plan.smem_bytes, stream>>>(args);
    count_launch();
    return cudaGetLastError();
}

}  // namespace nmfb
