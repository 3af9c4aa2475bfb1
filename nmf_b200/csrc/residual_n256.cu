// Residual-form objective, wide-tile variant: 128 x 256 tiles, i.e. tcgen05.mma with N = 256.
//
// STATUS: the objective kernel of the product for KP <= 128 with one V plane (MUR-eu / MUR-KL on bf16-exact data);
// everything else takes residual_ts.cu.  Round-2 A/B at 65536^2 k=128: 2.86 ms against 3.10 ms for the 128 x 128
// SS-form kernel of round 1 (deleted) and 2.95 ms for the TS form; ncu: tensor pipe 84.6 % active at the 1.36 GHz
// the power cap leaves, i.e. the kernel sits at the power-limited MMA rate (profiles/r02a_ncu_full_variants.md).
//
// Why: the ncu capture of the round-1 kernel residual_kernel<2,1> (profiles/r01b_all_paths.md) shows l1tex throughput 48.6 % at
// 49.3 % tensor-pipe activity, and the panel GEMM (N = 256 MMAs) 62-66 % at 85 %.  Both ratios are what the
// operand fetch of an SS-mode MMA predicts: an M = 128, N = 128, K = 16 instruction reads 4 KB of A and 4 KB of B
// from shared memory per 64 cycles = 128 B/cycle, the whole shared-memory bandwidth of an SM, so that kernel can
// never run its tensor pipe much above ~75 % (TMA fills and the epilogue's V reads need the same port); with
// N = 256 it is 4 + 8 KB per 128 cycles = 96 B/cycle.  The wide tile also halves the number of MMA instructions
// per FLOP (128-cycle MMAs against ~50 issue cycles each), so ONE issuing warp is enough and the dispatcher /
// second-issuer hand-off of the 128 x 128 kernels (residual_ts.cu) is not needed.
//
// Shared memory (KP = 128): W planes of the tile row 64 KB + two H^T stages of 64 KB ([256 rows x 64] x 2 planes)
// + two 16 KB V boxes = 224 KB.  V therefore streams through ONE [128 x 64] box per column half (the four
// epilogue warps of a half consume box after box); the refill of a box is an L2 hit (prefetched six tiles
// ahead) and overlaps the TMEM read-out of the next 64 columns.  Two 256-column accumulators fill TMEM (512).
//
// warp 0 = TMA producer of W / H^T planes, warp 1 = MMA issuer, warps 2..9 = epilogue (two per TMEM lane quarter,
// 128 columns each, as two 64-column blocks), warp 10 = TMA producer of the V boxes.
#include "kernels.cuh"
#include "ptx.cuh"

namespace nmfb {

namespace {
constexpr int kThreadsW = 352;
constexpr int kBoxW = 128 * 128;         // bytes of one [128 rows x 64 bf16] box
constexpr int kStageW = 4 * kBoxW;       // one H^T stage: 2 planes x [256 rows x 64 bf16]
constexpr int kStagesW = 2;
constexpr int kPrefetchTilesW = 3;       // L2 prefetch distance of the V stream, in 128 x 256 tiles

__host__ __device__ constexpr int wide_smem_payload(int KCH) {
    return 2 * KCH * kBoxW + kStagesW * kStageW + 2 * kBoxW;
}
}  // namespace

int residual_n256_smem_bytes(int KP, int NV) {
    if (NV != 1 || (KP != 64 && KP != 128)) return -1;
    return wide_smem_payload(KP / 64) + 1024 + 512;
}

template <int KCH>
__global__ void __launch_bounds__(kThreadsW, 1)
residual_n256_kernel(const __grid_constant__ ResidualArgs args) {
    extern __shared__ __align__(1024) uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) &
                                               ~static_cast<uintptr_t>(1023));
    constexpr int a_bytes = 2 * KCH * kBoxW;
    uint8_t* a_smem = smem;
    uint8_t* b_smem = a_smem + a_bytes;
    uint8_t* v_smem = b_smem + kStagesW * kStageW;
    uint64_t* bars = reinterpret_cast<uint64_t*>(v_smem + 2 * kBoxW);
    uint64_t* b_full = bars;             // [2]
    uint64_t* b_empty = bars + 2;        // [2]
    uint64_t* v_full = bars + 4;         // [2]  one per column half
    uint64_t* v_empty = bars + 6;        // [2]
    uint64_t* acc_full = bars + 8;       // [2]
    uint64_t* acc_empty = bars + 10;     // [2]
    uint64_t* a_full = bars + 12;
    uint64_t* a_empty = bars + 13;
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 14);
    double* red = reinterpret_cast<double*>(bars + 16);      // [8]

    const int warp = warp_idx_uniform(), lane = threadIdx.x & 31;
    // Clusters of CS CTAs walk CS adjacent row tiles in lockstep over the same column tiles; every H^T stage is
    // fetched once per cluster (each CTA loads 1/CS of its rows and multicasts), as in residual.cu.
    const int CS = args.cluster;
    const uint32_t crank = (CS > 1) ? cluster_ctarank() : 0;
    const uint16_t cmask = static_cast<uint16_t>((1u << CS) - 1);
    const int ncl = gridDim.x / CS, cid = blockIdx.x / CS;
    const int msuper = (args.mtiles + CS - 1) / CS;
    const long ntot = static_cast<long>(msuper) * args.ntiles;        // ntiles counts 256-column tiles here
    const long t_begin = ntot * cid / ncl;
    const long t_end = ntot * (cid + 1) / ncl;

    if (warp == 0 && lane == 0) {
        for (int p = 0; p < 2; ++p) { tma_prefetch_desc(&args.wr[p]); tma_prefetch_desc(&args.hr[p]); }
        tma_prefetch_desc(&args.v[0]);
        for (int s = 0; s < 2; ++s) {
            mbar_init(&b_full[s], 1);   mbar_init(&b_empty[s], CS);
            mbar_init(&v_full[s], 1);   mbar_init(&v_empty[s], 4);
            mbar_init(&acc_full[s], 1); mbar_init(&acc_empty[s], 8);
        }
        mbar_init(a_full, 1);
        mbar_init(a_empty, 1);
        fence_mbar_init();
    }
    if (warp == 1) { tmem_alloc(tmem_slot, 512); tmem_relinquish(); }
    tc_fence_before();
    __syncthreads();
    if (CS > 1) cluster_sync_all();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    if (warp == 0) {
        // ------------------------------------------------------------ producer of the factor planes
        if (lane == 0) {
            const int srow = 256 / CS;                      // H^T rows this CTA fetches of every stage
            int bs = 0; uint32_t bph = 0;
            uint32_t aph = 0;
            int cur_i = -1;
            for (long t = t_begin; t < t_end; ++t) {
                const int sr = static_cast<int>(t / args.ntiles);
                const int tj = static_cast<int>(t - static_cast<long>(sr) * args.ntiles);
                const int ti = min(sr * CS + static_cast<int>(crank), args.mtiles - 1);   // clamp: padded rows are masked
                if (sr != cur_i) {
                    mbar_wait(a_empty, aph ^ 1);
                    mbar_expect_tx(a_full, a_bytes);
                    for (int p = 0; p < 2; ++p)
                        for (int kc = 0; kc < KCH; ++kc)
                            tma_load_2d_hint(a_smem + (p * KCH + kc) * kBoxW, &args.wr[p], a_full, kc * 64,
                                             ti * 128, kPolicyEvictLast);
                    aph ^= 1;
                    cur_i = sr;
                }
                for (int kc = 0; kc < KCH; ++kc) {
                    mbar_wait(&b_empty[bs], bph ^ 1);
                    mbar_expect_tx(&b_full[bs], kStageW);
                    for (int p = 0; p < 2; ++p) {
                        uint8_t* dst = b_smem + bs * kStageW + p * 2 * kBoxW + crank * srow * 128;
                        if (CS > 1)
                            tma_load_2d_mcast(dst, &args.hr[p], &b_full[bs], kc * 64, tj * 256 + crank * srow, cmask,
                                              kPolicyEvictLast);
                        else
                            tma_load_2d_hint(dst, &args.hr[p], &b_full[bs], kc * 64, tj * 256, kPolicyEvictLast);
                    }
                    if (++bs == kStagesW) { bs = 0; bph ^= 1; }
                }
            }
        }
    } else if (warp == 1) {
        // ------------------------------------------------------------ MMA issuer (whole warp, converged)
        constexpr uint32_t idesc = umma_idesc_bf16(128, 256);
        const uint64_t desc0 = umma_desc_k_sw128(0);
        int bs = 0; uint32_t bph = 0;
        uint32_t accph = 0;                     // bit b = parity of acc_empty[b]
        uint32_t aph = 0;
        const uint64_t a_desc = desc0 + (smem_u32(a_smem) >> 4);
        const uint32_t b_addr = smem_u32(b_smem);
        int tj = static_cast<int>(t_begin % args.ntiles);
        bool new_row = true;
        int par = 0;
        for (long t = t_begin; t < t_end; ++t, par ^= 1) {
            if (new_row) {
                mbar_wait(a_full, aph);
                aph ^= 1;
                new_row = false;
            }
            mbar_wait(&acc_empty[par], ((accph >> par) & 1u) ^ 1u);
            accph ^= 1u << par;
            tc_fence_after();
            const uint32_t d = tmem_base + par * 256;
#pragma unroll
            for (int kc = 0; kc < KCH; ++kc) {
                mbar_wait(&b_full[bs], bph);
                tc_fence_after();
                const uint64_t b_desc = desc0 + ((b_addr + bs * kStageW) >> 4);
#pragma unroll
                for (int ks = 0; ks < 4; ++ks) {
                    const uint64_t a_hi = a_desc + (((0 * KCH + kc) * kBoxW + ks * 32) >> 4);
                    const uint64_t a_lo = a_desc + (((1 * KCH + kc) * kBoxW + ks * 32) >> 4);
                    const uint64_t b_hi = b_desc + ((ks * 32) >> 4);
                    const uint64_t b_lo = b_desc + ((2 * kBoxW + ks * 32) >> 4);
                    umma_bf16_warp(d, a_hi, b_hi, idesc, (kc == 0 && ks == 0) ? 0u : 1u);
                    umma_bf16_warp(d, a_hi, b_lo, idesc, 1);
                    umma_bf16_warp(d, a_lo, b_hi, idesc, 1);
                }
                if (CS > 1) umma_commit_mcast_warp(&b_empty[bs], cmask);
                else        umma_commit_warp(&b_empty[bs]);
                if (++bs == kStagesW) { bs = 0; bph ^= 1; }
            }
            umma_commit_warp(&acc_full[par]);
            if (++tj == args.ntiles) { tj = 0; new_row = true; }
            if (t + 1 >= t_end || new_row) umma_commit_warp(a_empty);     // W planes of the row are free again
        }
    } else if (warp == 10) {
        // ------------------------------------------------------------ producer of the V boxes
        if (lane == 0) {
            uint32_t vph = 0;                   // bit hb = parity of v_empty[hb]
            for (long t = t_begin; t < t_end; ++t) {
                const int sr = static_cast<int>(t / args.ntiles);
                const int tj = static_cast<int>(t - static_cast<long>(sr) * args.ntiles);
                const int ti = min(sr * CS + static_cast<int>(crank), args.mtiles - 1);
                {
                    const long tp = t + kPrefetchTilesW;
                    if (tp < t_end) {
                        const int srp = static_cast<int>(tp / args.ntiles);
                        const int tjp = static_cast<int>(tp - static_cast<long>(srp) * args.ntiles);
                        const int tip = min(srp * CS + static_cast<int>(crank), args.mtiles - 1);
                        for (int cb = 0; cb < 4; ++cb)
                            tma_prefetch_2d(&args.v[0], 0, (tjp * 4 + cb) * args.v_rows_p + tip * 128);
                    }
                }
                for (int cb = 0; cb < 2; ++cb)
                    for (int hb = 0; hb < 2; ++hb) {
                        mbar_wait(&v_empty[hb], ((vph >> hb) & 1u) ^ 1u);
                        vph ^= 1u << hb;
                        mbar_expect_tx(&v_full[hb], kBoxW);
                        tma_load_2d_hint(v_smem + hb * kBoxW, &args.v[0], &v_full[hb], 0,
                                         (tj * 4 + hb * 2 + cb) * args.v_rows_p + ti * 128, kPolicyEvictFirst);
                    }
            }
        }
    } else {
        // ------------------------------------------------------------ epilogue
        const int ew = warp - 2;
        const int q = warp & 3;                   // TMEM lane quarter
        const int hb = ew >> 2;                   // which 128-column half of the tile
        const int row = q * 32 + lane;
        int acc = 0; uint32_t accph = 0;
        uint32_t vph = 0;
        double total = 0.0;
        const uint8_t* vrow = v_smem + hb * kBoxW + row * 128;
        for (long t = t_begin; t < t_end; ++t) {
            const bool live = static_cast<int>(t / args.ntiles) * CS + static_cast<int>(crank) < args.mtiles;
            mbar_wait(&acc_full[acc], accph);
            tc_fence_after();
            float ts[4] = {0.f, 0.f, 0.f, 0.f};          // independent chains
#pragma unroll
            for (int cb = 0; cb < 2; ++cb) {
                const uint32_t taddr = tmem_base + (static_cast<uint32_t>(q * 32) << 16) + acc * 256 + hb * 128 + cb * 64;
                uint32_t s[2][32];
                tmem_ld32(taddr, s[0]);
                tmem_ld32(taddr + 32, s[1]);
                tmem_ld_wait();
                if (cb == 1) {                           // last read of this accumulator: hand it back
                    tc_fence_before();
                    __syncwarp();
                    if (lane == 0) mbar_arrive(&acc_empty[acc]);
                }
                mbar_wait(&v_full[hb], vph);
                vph ^= 1;
#pragma unroll
                for (int g = 0; g < 2; ++g) {
#pragma unroll
                    for (int c8 = 0; c8 < 4; ++c8) {
                        const int chunk = (g * 4 + c8) ^ (row & 7);          // SWIZZLE_128B
                        const uint4 raw = *reinterpret_cast<const uint4*>(vrow + chunk * 16);
                        const uint32_t w4[4] = {raw.x, raw.y, raw.z, raw.w};
                        float vv[8];
#pragma unroll
                        for (int e = 0; e < 4; ++e) {
                            vv[2 * e] = __uint_as_float(w4[e] << 16);
                            vv[2 * e + 1] = __uint_as_float(w4[e] & 0xFFFF0000u);
                        }
#pragma unroll
                        for (int e = 0; e < 8; ++e) {
                            const float sv = __uint_as_float(s[g][c8 * 8 + e]);
                            if (args.kl == 0) {
                                const float dlt = vv[e] - sv;
                                ts[e & 3] = fmaf(dlt, dlt, ts[e & 3]);
                            } else {
                                ts[e & 3] += kl_term(vv[e], sv);
                            }
                        }
                    }
                }
                __syncwarp();
                if (lane == 0) mbar_arrive(&v_empty[hb]);
            }
            if (live) total += static_cast<double>((ts[0] + ts[1]) + (ts[2] + ts[3]));
            if (++acc == 2) { acc = 0; accph ^= 1; }
        }
        for (int o = 16; o; o >>= 1) total += __shfl_xor_sync(0xffffffffu, total, o);
        if (lane == 0) red[ew] = total;
    }

    tc_fence_before();
    __syncthreads();
    if (CS > 1) cluster_sync_all();
    if (warp == 1) {
        tc_fence_after();
        tmem_dealloc(tmem_base, 512);
    }
    if (threadIdx.x == 0) {
        double t = 0.0;
        for (int w = 0; w < 8; ++w) t += red[w];
        args.partial[blockIdx.x] = t;
    }
}

cudaError_t residual_n256_launch(const ResidualArgs& args, int grid, cudaStream_t stream) {
    const int smem = residual_n256_smem_bytes(args.KP, args.NV);
    if (smem < 0 || args.BN != 256) return cudaErrorInvalidValue;
    void (*fn)(const ResidualArgs) = args.KP == 64 ? residual_n256_kernel<1> : residual_n256_kernel<2>;
    cudaError_t e = cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024);
    if (e != cudaSuccess) return e;
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(grid);                       // a multiple of args.cluster
    cfg.blockDim = dim3(kThreadsW);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = args.cluster; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr; cfg.numAttrs = 1;
    e = cudaLaunchKernelEx(&cfg, fn, args);
    if (e != cudaSuccess) return e;
    count_launch();
    return cudaGetLastError();
}

}  // namespace nmfb
