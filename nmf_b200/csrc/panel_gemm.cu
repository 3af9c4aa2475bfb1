// Split-precision skinny GEMM for the NMF contractions (V.H^T, W^T.V, Grams) on tcgen05.
//
//   P^T[c][r] = sum_t  F_j[c][t] * D_i[r][t]      (all allowed plane pairs (i,j) summed)
//
// D = "data" planes (V or V^T, bf16, K-blocked so that a [rows x 64] box is contiguous in HBM),
// F = "factor" planes (the other factor, bf16, component-major [KP][ld], t contiguous).
// fp32-grade products come from bf16 hi/lo(/lo2) planes: every MMA is an exact bf16 x bf16
// product accumulated in fp32 in TMEM.  The factor is the UMMA *A* operand (M = 128 components,
// or two stacked 64-row planes when KP = 64) and a panel of BN data rows is the *B* operand
// (N = BN), so one shared-memory stage of the big operand feeds every plane MMA.
//
// Persistent, warp-specialised, clustered:
//   warp 0      TMA producer.  The CTAs of a cluster work on adjacent row panels over the SAME
//               contraction range, so each CTA fetches 1/CS of every factor tile and multicasts it
//               to the whole cluster: the factor planes cross L2->SM once per cluster, not once
//               per CTA (they are as many bytes as the V panel itself).
//   warp 1      MMA issuer (+ TMEM owner).  Accumulates CHUNKS of `chunk_kb` contraction blocks
//               into one of two TMEM accumulators.
//   warps 2..9  drain each finished chunk into fp32 running sums held in registers (round to
//               nearest), then write the panel once per work item.  The tensor core adds into its
//               fp32 accumulator with truncation; keeping the chains short bounds that bias
//               (measured: 1.6e-5 relative after 4096 contraction elements, ~1e-6 after 512).
// Work items are (K-split, panel group) pairs walked round-robin by cluster; each item writes its
// own fp32 partial slot (no atomics => bit-reproducible), which the update kernels reduce in a
// fixed order.
//
// Reference call sites replaced: mur.py:29 `x @ h.T`, mur.py:45 `w.T @ x`, admm.py:221
// `other_aux.T @ data_aux`, ao_admm.py:53,56 `w.T @ w`, `w.T @ y`.
#include "kernels.cuh"
#include "ptx.cuh"

namespace nmfb {

namespace {

constexpr int kBK = 64;                 // contraction elements per stage (= one 128 B swizzle span)
constexpr int kATileBytes = 128 * 128;  // one 128-row A tile
constexpr int kEpiWarps = 8;
constexpr int kThreads = 64 + 32 * kEpiWarps;

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

// Compile-time shape <KP, NB, NA, BN>: the MMA issuer is ONE thread, and at ~6 cycles per dependent
// instruction every run-time loop bound or descriptor field it has to compute shows up directly as
// tensor-pipe idle time (measured: 3600 cycles per contraction block with run-time bounds against
// 1024 cycles of MMA work).  With everything static an MMA issue is an add and the instruction.
template <int KP, int NB, int NA, int BN>
__global__ void __launch_bounds__(kThreads, 1)
panel_gemm_kernel(const __grid_constant__ PanelGemmArgs args) {
    extern __shared__ __align__(1024) uint8_t smem_raw[];
    // SWIZZLE_128B wants 1024-byte alignment; the dynamic smem base is the same in every CTA of the
    // kernel, so aligning by hand keeps the multicast destination offsets identical cluster-wide.
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) &
                                               ~static_cast<uintptr_t>(1023));
    constexpr int A_TILES = (KP == 64) ? (NB + 1) / 2 : (KP == 128 ? NB : 2 * NB);
    constexpr int A_BYTES = A_TILES * kATileBytes;
    constexpr int B_PLANE = BN * 128;
    constexpr int STAGE = A_BYTES + NA * B_PLANE;
    constexpr int ACC_COLS = BN * (KP == 256 ? 2 : 1);   // TMEM columns per accumulator
    static_assert(2 * ACC_COLS <= 512, "two accumulators must fit TMEM");
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

    const int ncl = gridDim.x / CS;                      // clusters in the grid
    const int cid = blockIdx.x / CS;
    const int nitems = args.ngroups * args.nsplit;       // items per cluster walk
    const int kb_per = args.kb_per_split;
    const int chunk_kb = args.chunk_kb;

    // KP == 64 with an odd plane count leaves the upper half of the last A tile unused:
    // zero it once (TMA never writes there), so the stacked MMA adds exact zeros.
    if (KP == 64 && (NB & 1)) {
        for (int s = 0; s < nstages; ++s) {
            uint4* z = reinterpret_cast<uint4*>(stage_base + static_cast<size_t>(s) * STAGE +
                                                (A_TILES - 1) * kATileBytes + 8192);
            for (int i = threadIdx.x; i < 8192 / 16; i += kThreads) z[i] = make_uint4(0, 0, 0, 0);
        }
        fence_async_smem();
    }

    if (warp == 0 && lane == 0) {
        for (int j = 0; j < NB; ++j) tma_prefetch_desc(&args.fac[j]);
        for (int i = 0; i < NA; ++i) tma_prefetch_desc(&args.dat[i]);
        for (int s = 0; s < nstages; ++s) {
            mbar_init(&full_bar[s], 1);
            mbar_init(&empty_bar[s], CS);     // one MMA-retired arrive from every CTA of the cluster
        }
        for (int a = 0; a < 2; ++a) {
            mbar_init(&acc_full[a], 1);
            mbar_init(&acc_empty[a], kEpiWarps);
        }
        fence_mbar_init();
    }
    if (warp == 1) {
        tmem_alloc(tmem_slot, 512);
        tmem_relinquish();
    }
    tc_fence_before();
    __syncthreads();
    if (CS > 1) cluster_sync_all();           // peers' barriers are initialised before any multicast
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    if (warp == 0) {
        // ------------------------------------------------------------ TMA producer
        if (lane == 0) {
            constexpr int PROW = (KP == 64) ? 64 : 128;          // rows of one factor box (whole)
            constexpr int N_A_BOXES = (KP == 64) ? NB : A_TILES;
            constexpr uint32_t TX_BYTES = NA * B_PLANE + N_A_BOXES * PROW * 128;
            const int srow = PROW / CS;                          // rows this CTA fetches of each box
            const int dbr = args.dat_box_rows;
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
                    mbar_wait(&empty_bar[stage], phase ^ 1);
                    uint8_t* st = stage_base + static_cast<size_t>(stage) * STAGE;
                    mbar_expect_tx(&full_bar[stage], TX_BYTES);
                    const int t0 = kb * kBK;
#pragma unroll
                    for (int b = 0; b < N_A_BOXES; ++b) {
                        constexpr int dummy = 0; (void)dummy;
                        const int plane = (KP == 256) ? (b >> 1) : b;
                        const int row0 = (KP == 256) ? (b & 1) * 128 : 0;
                        const int off = (KP == 64) ? (b >> 1) * kATileBytes + (b & 1) * 8192 : b * kATileBytes;
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
                        for (int sub = 0; sub < BN; sub += dbr)
                            tma_load_2d_hint(st + A_BYTES + i * B_PLANE + sub * 128, &args.dat[i],
                                             &full_bar[stage], dc0, dc1 + sub, kPolicyEvictFirst);
                    const int pk = kb + args.prefetch_kb;            // pull the V stream into L2 ahead of use
                    if (args.prefetch_kb > 0 && pk < kb1)
                        for (int i = 0; i < NA; ++i)
                            for (int sub = 0; sub < BN; sub += dbr)
                                tma_prefetch_2d(&args.dat[i], rows_p ? 0 : pk * kBK,
                                                (rows_p ? pk * rows_p : 0) + panel * BN + sub);
                    if (++stage == nstages) { stage = 0; phase ^= 1; }
                }
            }
        }
    } else if (warp == 1) {
        // ------------------------------------------------------------ MMA issuer
        if (lane == 0) {
            constexpr uint32_t idesc = umma_idesc_bf16(128, BN);
            const uint64_t desc0 = umma_desc_k_sw128(0);           // layout fields; address added per use
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
                    const uint32_t d_base = tmem_base + acc * ACC_COLS;
                    for (int kb = cb; kb < ce; ++kb) {
                        const uint32_t fresh = (kb == cb) ? 0u : 1u;   // first block of a chunk overwrites
                        mbar_wait(&full_bar[stage], phase);
                        tc_fence_after();
                        const uint64_t da0 = desc0 + ((smem0 + stage * STAGE) >> 4);
                        const uint64_t db0 = da0 + (A_BYTES >> 4);
#pragma unroll
                        for (int ks = 0; ks < kBK / 16; ++ks) {
#pragma unroll
                            for (int t = 0; t < A_TILES; ++t) {
                                constexpr int dummy = 0; (void)dummy;
                                const int plane = (KP == 64) ? 0 : (KP == 128 ? t : (t >> 1));
                                const int rb = (KP == 256) ? (t & 1) : 0;
#pragma unroll
                                for (int i = 0; i < NA; ++i) {
                                    // plane pair filter: drop products below the lowest kept plane
                                    if (KP != 64 && (i + plane > NB - 1)) continue;
                                    // first MMA into this accumulator block within the stage?
                                    const bool first = (ks == 0) && (i == 0) && (KP == 256 ? (t < 2) : (t == 0));
                                    umma_bf16(d_base + rb * BN, da0 + ((t * kATileBytes + ks * 32) >> 4),
                                              db0 + ((i * B_PLANE + ks * 32) >> 4), idesc, first ? fresh : 1u);
                                }
                            }
                        }
                        // frees the smem stage cluster-wide when the MMAs retire
                        if (CS > 1) umma_commit_mcast(&empty_bar[stage], cmask);
                        else        umma_commit(&empty_bar[stage]);
                        if (++stage == nstages) { stage = 0; phase ^= 1; }
                    }
                    umma_commit(&acc_full[acc]);           // chunk complete -> drain
                    acc ^= 1;
                    if (acc == 0) acc_phase ^= 1;
                }
            }
        }
    } else {
        // ------------------------------------------------------------ drain + epilogue (warps 2..9)
        constexpr int NG = ACC_COLS / 64;                    // 32-column groups per thread (2 or 4)
        const int ew = warp - 2;
        const int q = warp & 3;                              // TMEM lane quarter this warp may read
        const int half = ew >> 2;                            // which half of the accumulator columns
        const int l = q * 32 + lane;                         // accumulator lane == stacked component row
        int acc = 0;
        uint32_t acc_phase = 0;
        for (int item = cid; item < nitems; item += ncl) {
            const int split = item / args.ngroups;
            const int group = item - split * args.ngroups;
            const int panel = group * CS + static_cast<int>(crank);
            const int kb0 = split * kb_per;
            const int kb1 = min(kb0 + kb_per, args.kblocks_total);
            float sums[NG * 32];
#pragma unroll
            for (int e = 0; e < NG * 32; ++e) sums[e] = 0.f;
            for (int cb = kb0; cb < kb1; cb += chunk_kb) {
                mbar_wait(&acc_full[acc], acc_phase);
                tc_fence_after();
                const uint32_t taddr = tmem_base + (static_cast<uint32_t>(q * 32) << 16) + acc * ACC_COLS +
                                       half * (ACC_COLS / 2);
#pragma unroll
                for (int g = 0; g < NG; ++g) {
                    uint32_t r[32];
                    tmem_ld32(taddr + g * 32, r);
                    tmem_ld_wait();
#pragma unroll
                    for (int e = 0; e < 32; ++e) sums[g * 32 + e] += __uint_as_float(r[e]);
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
                if (KP == 64) { c = l & 63; slot = split * 2 + (l >> 6); }
                else          { c = rb * 128 + l; slot = split; }
                float* dst = args.out + (static_cast<size_t>(slot) * KP + c) * args.Lp +
                             static_cast<size_t>(panel) * BN + r0;
#pragma unroll
                for (int g = 0; g < NG; ++g) {
                    float4* d4 = reinterpret_cast<float4*>(dst + g * 32);
#pragma unroll
                    for (int v = 0; v < 8; ++v)
                        d4[v] = make_float4(sums[g * 32 + 4 * v], sums[g * 32 + 4 * v + 1],
                                            sums[g * 32 + 4 * v + 2], sums[g * 32 + 4 * v + 3]);
                }
            }
        }
    }

    tc_fence_before();
    __syncthreads();
    if (CS > 1) cluster_sync_all();           // nobody exits while a peer may still multicast to it
    if (warp == 1) {
        tc_fence_after();
        tmem_dealloc(tmem_base, 512);
    }
}

// Every supported <KP, NB, NA, BN>: MUR (NB = 2), ADMM / AO-ADMM (NB = 3), Grams (BN = 128, data = the
// factor's own planes).
using KernelFn = void (*)(const PanelGemmArgs);
struct KernelEntry { int KP, NB, NA, BN; KernelFn fn; };
#define PG_ENTRY(kp, nb, na, bn) {kp, nb, na, bn, panel_gemm_kernel<kp, nb, na, bn>}
static const KernelEntry kKernels[] = {
    PG_ENTRY(64, 2, 1, 256),  PG_ENTRY(64, 2, 2, 256),  PG_ENTRY(128, 2, 1, 256), PG_ENTRY(128, 2, 2, 256),
    PG_ENTRY(256, 2, 1, 128), PG_ENTRY(256, 2, 2, 128),
    PG_ENTRY(64, 3, 1, 256),  PG_ENTRY(64, 3, 2, 256),  PG_ENTRY(128, 3, 1, 256), PG_ENTRY(128, 3, 2, 256),
    PG_ENTRY(256, 3, 1, 128), PG_ENTRY(256, 3, 2, 64),
    PG_ENTRY(64, 2, 2, 128),  PG_ENTRY(128, 2, 2, 128), PG_ENTRY(64, 3, 3, 128),  PG_ENTRY(128, 3, 3, 128),
    PG_ENTRY(64, 2, 1, 128),  PG_ENTRY(128, 2, 1, 128), PG_ENTRY(64, 3, 1, 128),  PG_ENTRY(128, 3, 1, 128),
};
static KernelFn find_kernel(int KP, int NB, int NA, int BN) {
    for (const KernelEntry& e : kKernels)
        if (e.KP == KP && e.NB == NB && e.NA == NA && e.BN == BN) return e.fn;
    return nullptr;
}

// ---------------------------------------------------------------------------------------------
// host side
// ---------------------------------------------------------------------------------------------
// Clusters of `cs` CTAs (one CTA per SM: the kernel uses all shared memory) that can be resident at
// once.  GPC boundaries make this less than num_sms / cs; a persistent grid must not exceed it.
static int max_resident_clusters(int cs, int num_sms) {
    static int cache[9] = {0, 0, 0, 0, 0, 0, 0, 0, 0};
    if (cs == 1) return num_sms;
    if (cache[cs] > 0) return cache[cs];
    KernelFn probe = panel_gemm_kernel<128, 2, 1, 256>;
    cudaFuncSetAttribute(probe, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024);
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(num_sms / cs * cs);
    cfg.blockDim = dim3(kThreads);
    cfg.dynamicSmemBytes = 200 * 1024;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = cs;
    attr[0].val.clusterDim.y = 1;
    attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    int n = 0;
    if (cudaOccupancyMaxActiveClusters(&n, probe, &cfg) != cudaSuccess || n <= 0) {
        cudaGetLastError();
        n = num_sms / cs * 3 / 4;                       // conservative guess
    }
    if (n > num_sms / cs) n = num_sms / cs;
    cache[cs] = n;
    return n;
}

int panel_gemm_plan(int KP, int NB, int NA, int L, int R, int force_bn, int force_split, int num_sms,
                    PanelGemmPlan* plan) {
    if (!(KP == 64 || KP == 128 || KP == 256) || NB < 1 || NB > 3 || NA < 1 || NA > 3) return -1;
    const int smem_budget = 227 * 1024 - 1024 /*align slack*/ - 256 /*barriers*/;
    // TMEM: two accumulators of BN (x2 row blocks at KP = 256) columns must fit 512 columns.
    int BN = 0, nstages = 0;
    StageLayout sl = {};
    const int cand[3] = {256, 128, 64};
    for (int ci = 0; ci < 3; ++ci) {
        const int bn = cand[ci];
        if (force_bn && bn != force_bn) continue;
        if (2 * bn * (KP == 256 ? 2 : 1) > 512) continue;          // two TMEM accumulators
        if (!find_kernel(KP, NB, NA, bn)) continue;
        sl = stage_layout(KP, NB, NA, bn);
        if (smem_budget / sl.stage_bytes < 2) continue;
        BN = bn;
        nstages = smem_budget / sl.stage_bytes;
        break;
    }
    if (BN == 0) return -2;
    if (nstages > 8) nstages = 8;
    const int npanels = (L + BN - 1) / BN;
    const int kblocks = (R + kBK - 1) / kBK;
    // cluster size: the factor tiles are multicast across the cluster; needs enough panels.
    int cluster = (KP == 64) ? 4 : 2;
    while (cluster > 1 && npanels < cluster) cluster >>= 1;
    const int ngroups = (npanels + cluster - 1) / cluster;
    const int ncl = max_resident_clusters(cluster, num_sms);
    // K-split: fill the machine (items ~ multiple of #clusters) while keeping >= 16 k-blocks per item.
    int nsplit = 1;
    if (force_split > 0) {
        nsplit = force_split;
    } else {
        double best = 1e30;
        const int max_split = 64;
        for (int s = 1; s <= max_split; ++s) {
            if (s > 1 && kblocks / s < 16) break;
            const long items = static_cast<long>(ngroups) * s;
            const long waves = (items + ncl - 1) / ncl;
            // cost ~ waves * k-blocks per item  (+ epilogue/partial traffic penalty per split)
            const double cost = static_cast<double>(waves) * ((kblocks + s - 1) / s + 8);
            if (cost < best * 0.97) { best = cost; nsplit = s; }
        }
    }
    if (nsplit > kblocks) nsplit = kblocks;
    int kb_per = (kblocks + nsplit - 1) / nsplit;
    nsplit = (kblocks + kb_per - 1) / kb_per;            // no empty splits
    const int chunk_kb = 8;
    plan->BN = BN;
    plan->nstages = nstages;
    plan->npanels = npanels;
    plan->nsplit = nsplit;
    plan->kb_per_split = kb_per;
    plan->kblocks_total = kblocks;
    plan->slots = nsplit * (KP == 64 ? 2 : 1);
    plan->smem_bytes = nstages * sl.stage_bytes + 1024 + 256;
    plan->cluster = cluster;
    plan->ngroups = ngroups;
    plan->chunk_kb = chunk_kb;
    plan->dat_box_rows = BN;
    plan->prefetch_kb = 0;
    plan->max_clusters = ncl;
    return 0;
}

cudaError_t panel_gemm_launch(const PanelGemmArgs& args, const PanelGemmPlan& plan, int num_sms,
                              cudaStream_t stream) {
    KernelFn fn = find_kernel(args.KP, args.NB, args.NA, args.BN);
    if (!fn) return cudaErrorInvalidValue;
    cudaError_t e0 = cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024);
    if (e0 != cudaSuccess) return e0;
    const int CS = plan.cluster;
    const int nitems = plan.ngroups * plan.nsplit;
    int ncl = plan.max_clusters;
    if (nitems < ncl) ncl = nitems;
    (void)num_sms;
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(ncl * CS);
    cfg.blockDim = dim3(kThreads);
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
