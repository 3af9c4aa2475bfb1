// "Exact-hi" variant of the panel GEMM for the ADMM / AO-ADMM right-hand sides and Grams.
//
// The solves of admm.py:230 / ao_admm.py:59 amplify errors of the right-hand side A^T V: on the
// reference's own trajectories a RANDOM relative perturbation of 1e-8 per element already costs
// 6e-5 in the objective a few iterations later (a uniform bias is harmless), and an fp32 tensor-core
// accumulation chain of ~100 truncating adds scatters by ~3e-7.  So here plane 0 of BOTH operands is
// a fixed-point slice (ptx.cuh: fixed_unit / fixed_slice): its products are integers of <= 2^14
// common units and a chunk of <= 512 of them accumulates in TMEM WITHOUT ANY rounding (accumulator
// E).  All remaining plane pairs are 2^-7 smaller and go to a second accumulator (F), whose rounding
// scatter is 2^-7 smaller too.  Each chunk is drained into double-float running sums (error-free
// TwoSum on the CUDA cores) and written as a (hi, lo) pair of fp32 partial slots, which the
// consumers add in fp64.
//
// Structure (roles, pipeline, cluster multicast of the factor tiles) as in panel_gemm.cu.
#include "kernels.cuh"
#include "ptx.cuh"

namespace nmfb {

namespace {
constexpr int kBKx = 64;
constexpr int kATileX = 128 * 128;
constexpr int kThreadsX = 320;
}  // namespace

template <int KP, int NA, int BN>
__global__ void __launch_bounds__(kThreadsX, 1)
panel_gemm_exact_kernel(const __grid_constant__ PanelGemmArgs args) {
    extern __shared__ __align__(1024) uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) &
                                               ~static_cast<uintptr_t>(1023));
    constexpr int NB = 3;
    constexpr int A_TILES = (KP == 64) ? 2 : (KP == 128 ? 3 : 6);
    constexpr int A_BYTES = A_TILES * kATileX;
    constexpr int B_PLANE = BN * 128;
    constexpr int STAGE = A_BYTES + NA * B_PLANE;
    constexpr int ACCW = BN * (KP == 256 ? 2 : 1);       // columns of ONE accumulator (E or F)
    static_assert(4 * ACCW <= 512, "2 buffers x (E, F) must fit TMEM");
    static_assert(NA >= 2, "exact mode needs the fixed-point plane and at least one remainder plane");
    const int CS = args.cluster;
    const int nstages = args.nstages;
    uint8_t* stage_base = smem;
    uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + static_cast<size_t>(nstages) * STAGE);
    uint64_t* empty_bar = full_bar + nstages;
    uint64_t* acc_full = empty_bar + nstages;     // [2]
    uint64_t* acc_empty = acc_full + 2;           // [2]
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(acc_empty + 2);

    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;
    const uint32_t crank = (CS > 1) ? cluster_ctarank() : 0;
    const uint16_t cmask = static_cast<uint16_t>((1u << CS) - 1);
    const int ncl = gridDim.x / CS;
    const int cid = blockIdx.x / CS;
    const int nitems = args.ngroups * args.nsplit;
    const int kb_per = args.kb_per_split;
    const int chunk_kb = args.chunk_kb;

    if (KP == 64) {      // upper half of the second stacked tile (plane 2 | nothing) stays zero
        for (int s = 0; s < nstages; ++s) {
            uint4* z = reinterpret_cast<uint4*>(stage_base + static_cast<size_t>(s) * STAGE + kATileX + 8192);
            for (int i = threadIdx.x; i < 8192 / 16; i += kThreadsX) z[i] = make_uint4(0, 0, 0, 0);
        }
        fence_async_smem();
    }
    if (warp == 0 && lane == 0) {
        for (int j = 0; j < NB; ++j) tma_prefetch_desc(&args.fac[j]);
        for (int i = 0; i < NA; ++i) tma_prefetch_desc(&args.dat[i]);
        for (int s = 0; s < nstages; ++s) {
            mbar_init(&full_bar[s], 1);
            mbar_init(&empty_bar[s], CS);
        }
        for (int a = 0; a < 2; ++a) {
            mbar_init(&acc_full[a], 1);
            mbar_init(&acc_empty[a], 8);
        }
        fence_mbar_init();
    }
    if (warp == 1) {
        tmem_alloc(tmem_slot, 512);
        tmem_relinquish();
    }
    tc_fence_before();
    __syncthreads();
    if (CS > 1) cluster_sync_all();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    if (warp == 0) {
        // ------------------------------------------------------------ TMA producer
        if (lane == 0) {
            constexpr int PROW = (KP == 64) ? 64 : 128;
            constexpr int N_A_BOXES = (KP == 64) ? NB : A_TILES;
            constexpr uint32_t TX_BYTES = NA * B_PLANE + N_A_BOXES * PROW * 128;
            const int srow = PROW / CS;
            const int rows_p = args.dat_rows_p;
            int stage = 0;
            uint32_t phase = 0;
            for (int item = cid; item < nitems; item += ncl) {
                const int split = item / args.ngroups;
                const int group = item - split * args.ngroups;
                const int panel = group * CS + static_cast<int>(crank);
                const int kb0 = split * kb_per;
                const int kb1 = min(kb0 + kb_per, args.kblocks_total);
                for (int kb = kb0; kb < kb1; ++kb) {
                    const int pk = kb + args.prefetch_kb;
                    if (args.prefetch_kb > 0 && pk < kb1)
#pragma unroll
                        for (int i = 0; i < NA; ++i)
                            tma_prefetch_2d(&args.dat[i], rows_p ? 0 : pk * kBKx,
                                            (rows_p ? pk * rows_p : 0) + panel * BN);
                    mbar_wait(&empty_bar[stage], phase ^ 1);
                    uint8_t* st = stage_base + static_cast<size_t>(stage) * STAGE;
                    mbar_expect_tx(&full_bar[stage], TX_BYTES);
                    const int t0 = kb * kBKx;
#pragma unroll
                    for (int b = 0; b < N_A_BOXES; ++b) {
                        const int plane = (KP == 256) ? (b >> 1) : b;
                        const int row0 = (KP == 256) ? (b & 1) * 128 : 0;
                        const int off = (KP == 64) ? (b >> 1) * kATileX + (b & 1) * 8192 : b * kATileX;
                        uint8_t* dst = st + off + crank * srow * 128;
                        if (CS > 1)
                            tma_load_2d_mcast(dst, &args.fac[plane], &full_bar[stage], t0, row0 + crank * srow,
                                              cmask, kPolicyEvictLast);
                        else
                            tma_load_2d_hint(dst, &args.fac[plane], &full_bar[stage], t0, row0, kPolicyEvictLast);
                    }
                    const int dc0 = rows_p ? 0 : t0;
                    const int dc1 = (rows_p ? kb * rows_p : 0) + panel * BN;
#pragma unroll
                    for (int i = 0; i < NA; ++i)
                        tma_load_2d_hint(st + A_BYTES + i * B_PLANE, &args.dat[i], &full_bar[stage], dc0, dc1,
                                         kPolicyEvictFirst);
                    if (++stage == nstages) { stage = 0; phase ^= 1; }
                }
            }
        }
    } else if (warp == 1) {
        // ------------------------------------------------------------ MMA issuer
        if (lane == 0) {
            constexpr uint32_t idesc = umma_idesc_bf16(128, BN);
            const uint64_t desc0 = umma_desc_k_sw128(0);
            const uint32_t smem0 = smem_u32(stage_base);
            int stage = 0;
            uint32_t phase = 0;
            int acc = 0;
            uint32_t acc_phase = 0;
            for (int item = cid; item < nitems; item += ncl) {
                const int split = item / args.ngroups;
                const int kb0 = split * kb_per;
                const int kb1 = min(kb0 + kb_per, args.kblocks_total);
                for (int cb = kb0; cb < kb1; cb += chunk_kb) {
                    const int ce = min(cb + chunk_kb, kb1);
                    mbar_wait(&acc_empty[acc], acc_phase ^ 1);
                    tc_fence_after();
                    const uint32_t e_base = tmem_base + acc * (2 * ACCW);      // exact accumulator
                    const uint32_t f_base = e_base + ACCW;                      // floating accumulator
                    for (int kb = cb; kb < ce; ++kb) {
                        const uint32_t fresh = (kb == cb) ? 0u : 1u;
                        mbar_wait(&full_bar[stage], phase);
                        tc_fence_after();
                        const uint64_t da0 = desc0 + ((smem0 + stage * STAGE) >> 4);
                        const uint64_t db0 = da0 + (A_BYTES >> 4);
#pragma unroll
                        for (int ks = 0; ks < kBKx / 16; ++ks) {
#pragma unroll
                            for (int t = 0; t < A_TILES; ++t) {
                                const int plane = (KP == 64) ? (t == 0 ? 0 : 2) : (KP == 128 ? t : (t >> 1));
                                const int rb = (KP == 256) ? (t & 1) : 0;
#pragma unroll
                                for (int i = 0; i < NA; ++i) {
                                    if (plane + i > 2) continue;                 // below the kept precision
                                    const bool exact = (plane == 0 && i == 0);   // fixed x fixed -> E
                                    // first MMA into (E | F) of this row block within the chunk's first stage
                                    const bool first = (ks == 0) && (plane == 0) && (exact ? true : (i == 1));
                                    umma_bf16((exact ? e_base : f_base) + rb * BN,
                                              da0 + ((t * kATileX + ks * 32) >> 4),
                                              db0 + ((i * B_PLANE + ks * 32) >> 4), idesc, first ? fresh : 1u);
                                }
                            }
                        }
                        if (CS > 1) umma_commit_mcast(&empty_bar[stage], cmask);
                        else        umma_commit(&empty_bar[stage]);
                        if (++stage == nstages) { stage = 0; phase ^= 1; }
                    }
                    umma_commit(&acc_full[acc]);
                    acc ^= 1;
                    if (acc == 0) acc_phase ^= 1;
                }
            }
        }
    } else {
        // ------------------------------------------------------------ drain (TwoSum) + epilogue
        constexpr int NG = ACCW / 64;                        // 32-column groups per thread (1 or 2)
        const int ew = warp - 2;
        const int q = warp & 3;
        const int half = ew >> 2;
        const int l = q * 32 + lane;
        int acc = 0;
        uint32_t acc_phase = 0;
        for (int item = cid; item < nitems; item += ncl) {
            const int split = item / args.ngroups;
            const int group = item - split * args.ngroups;
            const int panel = group * CS + static_cast<int>(crank);
            const int kb0 = split * kb_per;
            const int kb1 = min(kb0 + kb_per, args.kblocks_total);
            float hi[NG * 32], lo[NG * 32];
#pragma unroll
            for (int e = 0; e < NG * 32; ++e) { hi[e] = 0.f; lo[e] = 0.f; }
            for (int cb = kb0; cb < kb1; cb += chunk_kb) {
                mbar_wait(&acc_full[acc], acc_phase);
                tc_fence_after();
                const uint32_t taddr = tmem_base + (static_cast<uint32_t>(q * 32) << 16) + acc * (2 * ACCW) +
                                       half * (ACCW / 2);
#pragma unroll
                for (int g = 0; g < NG; ++g) {
                    uint32_t re[32], rf[32];
                    tmem_ld32(taddr + g * 32, re);
                    tmem_ld32(taddr + ACCW + g * 32, rf);
                    tmem_ld_wait();
#pragma unroll
                    for (int e = 0; e < 32; ++e) {
                        const float x = __uint_as_float(re[e]);
                        const float s = __fadd_rn(hi[g * 32 + e], x);            // TwoSum (Knuth): s + err == hi + x
                        const float bb = __fsub_rn(s, hi[g * 32 + e]);
                        const float err = __fadd_rn(__fsub_rn(hi[g * 32 + e], __fsub_rn(s, bb)), __fsub_rn(x, bb));
                        hi[g * 32 + e] = s;
                        lo[g * 32 + e] = __fadd_rn(lo[g * 32 + e], __fadd_rn(err, __uint_as_float(rf[e])));
                    }
                }
                tc_fence_before();
                __syncwarp();
                if (lane == 0) mbar_arrive(&acc_empty[acc]);
                acc ^= 1;
                if (acc == 0) acc_phase ^= 1;
            }
            if (panel < args.npanels) {
                int c, slot, rb, r0;
                if (KP == 256) { rb = half; r0 = 0; } else { rb = 0; r0 = half * (BN / 2); }
                if (KP == 64) { c = l & 63; slot = (split * 2 + (l >> 6)) * 2; }
                else          { c = rb * 128 + l; slot = split * 2; }
                float* dst = args.out + (static_cast<size_t>(slot) * KP + c) * args.Lp +
                             static_cast<size_t>(panel) * BN + r0;
                float* dst_lo = dst + static_cast<size_t>(KP) * args.Lp;
#pragma unroll
                for (int g = 0; g < NG; ++g)
#pragma unroll
                    for (int v = 0; v < 8; ++v) {
                        reinterpret_cast<float4*>(dst + g * 32)[v] =
                            make_float4(hi[g * 32 + 4 * v], hi[g * 32 + 4 * v + 1], hi[g * 32 + 4 * v + 2], hi[g * 32 + 4 * v + 3]);
                        reinterpret_cast<float4*>(dst_lo + g * 32)[v] =
                            make_float4(lo[g * 32 + 4 * v], lo[g * 32 + 4 * v + 1], lo[g * 32 + 4 * v + 2], lo[g * 32 + 4 * v + 3]);
                    }
            }
        }
    }

    tc_fence_before();
    __syncthreads();
    if (CS > 1) cluster_sync_all();
    if (warp == 1) {
        tc_fence_after();
        tmem_dealloc(tmem_base, 512);
    }
}

namespace {
using KernelFnX = void (*)(const PanelGemmArgs);
KernelFnX find_exact(int KP, int NA, int BN) {
    if (KP == 64 && NA == 2 && BN == 128) return panel_gemm_exact_kernel<64, 2, 128>;
    if (KP == 64 && NA == 3 && BN == 128) return panel_gemm_exact_kernel<64, 3, 128>;
    if (KP == 128 && NA == 2 && BN == 128) return panel_gemm_exact_kernel<128, 2, 128>;
    if (KP == 128 && NA == 3 && BN == 128) return panel_gemm_exact_kernel<128, 3, 128>;
    if (KP == 256 && NA == 2 && BN == 64) return panel_gemm_exact_kernel<256, 2, 64>;
    return nullptr;
}
}  // namespace

int panel_gemm_exact_plan(int KP, int NA, int L, int R, int blocked, int num_sms, PanelGemmPlan* plan) {
    const int BN = (KP == 256) ? 64 : 128;
    if (!find_exact(KP, NA, BN)) return -1;
    const int a_tiles = (KP == 64) ? 2 : (KP == 128 ? 3 : 6);
    const int stage = a_tiles * kATileX + NA * BN * 128;
    const int smem_budget = 227 * 1024 - 1024 - 256;
    int nstages = smem_budget / stage;
    if (nstages < 2) return -2;
    if (nstages > 6) nstages = 6;
    const int npanels = (L + BN - 1) / BN;
    const int kblocks = (R + kBKx - 1) / kBKx;
    int cluster = 2;
    while (cluster > 1 && npanels < cluster) cluster >>= 1;
    const int ngroups = (npanels + cluster - 1) / cluster;
    PanelGemmPlan probe;                      // resident-cluster count of the sibling kernel (same footprint)
    int ncl = num_sms / cluster;
    if (cluster > 1 && panel_gemm_plan(128, 2, 1, 1 << 20, 1 << 16, 0, 1, num_sms, &probe) == 0 && probe.cluster == cluster)
        ncl = probe.max_clusters;
    else if (cluster > 1)
        ncl = num_sms / cluster * 3 / 4;
    if (ncl > num_sms / cluster) ncl = num_sms / cluster;     // a caller may hand over fewer SMs than the device has
    int nsplit = 1;
    double best = 1e30;
    for (int s = 1; s <= 64; ++s) {
        if (s > 1 && kblocks / s < 16) break;
        const long items = static_cast<long>(ngroups) * s;
        const long waves = (items + ncl - 1) / ncl;
        const double cost = static_cast<double>(waves) * ((kblocks + s - 1) / s + 8);
        if (cost < best * 0.97) { best = cost; nsplit = s; }
    }
    int kb_per = (kblocks + nsplit - 1) / nsplit;
    nsplit = (kblocks + kb_per - 1) / kb_per;
    plan->BN = BN;
    plan->nstages = nstages;
    plan->npanels = npanels;
    plan->nsplit = nsplit;
    plan->kb_per_split = kb_per;
    plan->kblocks_total = kblocks;
    plan->slots = nsplit * (KP == 64 ? 2 : 1) * 2;
    plan->smem_bytes = nstages * stage + 1024 + 256;
    plan->cluster = cluster;
    plan->ngroups = ngroups;
    plan->chunk_kb = 8;                       // <= 512 contraction elements: exact in the E accumulator
    plan->max_clusters = ncl;
    plan->dat_box_rows = BN;
    plan->prefetch_kb = blocked ? 6 : 0;
    return 0;
}

cudaError_t panel_gemm_exact_launch(const PanelGemmArgs& args, const PanelGemmPlan& plan, cudaStream_t stream) {
    KernelFnX fn = find_exact(args.KP, args.NA, args.BN);
    if (!fn) return cudaErrorInvalidValue;
    cudaError_t e0 = cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024);
    if (e0 != cudaSuccess) return e0;
    const int CS = plan.cluster;
    const int nitems = plan.ngroups * plan.nsplit;
    int ncl = plan.max_clusters;
    if (nitems < ncl) ncl = nitems;
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(ncl * CS);
    cfg.blockDim = dim3(kThreadsX);
    cfg.dynamicSmemBytes = plan.smem_bytes;
    cfg.stream = stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = CS;
    attr[0].val.clusterDim.y = 1;
    attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    cudaError_t e = cudaLaunchKernelEx(&cfg, fn, args);
    if (e != cudaSuccess) return e;
    count_launch();
    return cudaGetLastError();
}

}  // namespace nmfb
