// Residual-form objective on tcgen05:  sum_{i,j} f(V_ij, S_ij),  S = W.H formed tile by tile in TMEM and never
// written to HBM.
//   mode 0 (eu):  f = (v - s)^2                         -> 0.5 * sum = utils.py:29
//   mode 1 (kl):  f = v*log(v/s) - v + s  (+inf/NaN of the log term -> 0, utils.py:23-26)
// S needs fp32-grade accuracy because the residual is ~1e-4 of |V| on low-rank data (the trace form
// ||V||^2 - 2<W^T V,H> + <W^T W,H H^T> cancels catastrophically), so each 128x128 tile is three bf16 MMA chains:
// Whi.Hhi + Whi.Hlo + Wlo.Hhi.
//
// "TS" form: the resident operand (the W planes of the tile row) lives in TENSOR MEMORY and is the A operand of
// tcgen05.mma; only the streamed H^T stages are read from shared memory.  This kernel serves every shape the
// wide-tile kernel (residual_n256.cu) does not: KP = 256, two V planes (hi + lo, or the fixed-point planes of
// ADMM / AO-ADMM).  Round-2 A/B at 65536^2 k=128 (profiles/r02a_*): SS-form 128x128 kernel 3.10 ms (deleted),
// this kernel 2.95 ms, wide tiles 2.86 ms; ADMM objective (two V planes) 1.07 -> 0.97 ms.
// One CTA walks a contiguous run of 128x128 tiles in row-major tile order; warp 0 = TMA producer of the H^T
// stages, warps 1 and 11 = MMA issuers (even / odd tiles, one TMEM accumulator each), warps 2..9 = epilogue
// warps, warp 10 = TMA producer of the V tiles.  Relative to a plain SS-form tile loop:
//   1. TMEM holds two 128-column S accumulators (columns 0..255) AND the W planes (columns 256..): 128 lanes =
//      the 128 rows of the tile, two bf16 K-values per 32-bit column, hi plane then lo plane, KP/2 columns each
//      (KP = 256: 2 x 128 columns, 512 in all).  The eight epilogue warps write them with tcgen05.st at the first
//      tile of every tile row (warp = lane quarter x plane; a row of a plane is KP contiguous bf16 of the
//      row-major plane, i.e. exactly the KP/2 words of its TMEM lane), after the MMAs of the previous row have
//      retired (a_empty), and hand them to the issuer through a_full.
//   2. The MMAs are issued in the [d], [a_tmem], b_desc form.  Per 64-cycle instruction the tensor core then
//      fetches 4 KB from shared memory instead of 8 KB: 64 instead of 128 B/cycle, the port limit that
//      profiles/r01e_smem_operand_model.md derives from the r01b captures.
//   3. The 64 KB of shared memory the W planes used to take go to the H^T ring (4 stages at every KP).
//
#include "kernels.cuh"
#include "ptx.cuh"

namespace nmfb {

namespace {
constexpr int kThreadsT = 384;
constexpr int kBox = 128 * 128;      // bytes of one [128 rows x 64 bf16] box
constexpr int kPrefetchTiles = 6;    // L2 prefetch distance of the V stream, in 128 x 128 tiles

struct TsLayout {
    int a_bytes, b_stage_bytes, v_buf_bytes;
    int nbs, nvs;
    int total;
};

__host__ __device__ inline TsLayout ts_layout(int KP, int NV) {
    TsLayout l;
    l.a_bytes = 0;                       // W planes live in tensor memory
    (void)KP;
    l.b_stage_bytes = 2 * kBox;
    l.v_buf_bytes = NV * 2 * kBox;
    const int budget = 227 * 1024 - 1024 - 512;
    int rest = budget - l.a_bytes;
    l.nvs = (rest - 2 * l.v_buf_bytes >= 2 * l.b_stage_bytes) ? 2 : 1;
    rest -= l.nvs * l.v_buf_bytes;
    l.nbs = rest / l.b_stage_bytes;
    if (l.nbs > 4) l.nbs = 4;
    l.total = l.a_bytes + l.nbs * l.b_stage_bytes + l.nvs * l.v_buf_bytes;
    return l;
}
}  // namespace

int residual_ts_smem_bytes(int KP, int NV) {
    const TsLayout l = ts_layout(KP, NV);
    if (l.nbs < 2) return -1;
    return l.total + 1024 + 512;
}

// Compile-time <KCH = KP / 64, NV>: the single MMA-issuing thread must not spend its time computing
// loop bounds and descriptor fields (see panel_gemm.cu).
template <int KCH, int NV>
__global__ void __launch_bounds__(kThreadsT, 1)
residual_ts_kernel(const __grid_constant__ ResidualArgs args) {
    extern __shared__ __align__(1024) uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) &
                                               ~static_cast<uintptr_t>(1023));
    constexpr int KP = KCH * 64;
    constexpr int kchunks = KCH;
    const TsLayout L = ts_layout(KP, NV);
    uint8_t* b_smem = smem;
    uint8_t* v_smem = b_smem + L.nbs * L.b_stage_bytes;
    uint64_t* bars = reinterpret_cast<uint64_t*>(v_smem + L.nvs * L.v_buf_bytes);
    uint64_t* b_full = bars;             // [4]
    uint64_t* b_empty = bars + 4;        // [4]
    uint64_t* v_full = bars + 8;         // [2]
    uint64_t* v_empty = bars + 10;       // [2]
    uint64_t* acc_full = bars + 12;      // [2]
    uint64_t* acc_empty = bars + 14;     // [2]
    uint64_t* a_full = bars + 16;
    uint64_t* a_empty = bars + 17;
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 18);
    uint64_t* go_b = bars + 19;          // issuer A -> issuer B: "tile ready"
    double* red = reinterpret_cast<double*>(bars + 20);      // [8]

    const int warp = warp_idx_uniform(), lane = threadIdx.x & 31;
    // Clusters of CS CTAs walk "super rows" of CS adjacent row tiles in lockstep over the same column
    // tiles, so the H^T stage every one of them needs is fetched once per cluster (each CTA loads
    // 1/CS of it and multicasts): the kernel is bound by L2 output otherwise (3 x |V| of L2 reads).
    const int CS = args.cluster;
    const uint32_t crank = (CS > 1) ? cluster_ctarank() : 0;
    const uint16_t cmask = static_cast<uint16_t>((1u << CS) - 1);
    const int ncl = gridDim.x / CS, cid = blockIdx.x / CS;
    const int msuper = (args.mtiles + CS - 1) / CS;
    const long ntot = static_cast<long>(msuper) * args.ntiles;
    const long t_begin = ntot * cid / ncl;
    const long t_end = ntot * (cid + 1) / ncl;

    if (warp == 0 && lane == 0) {
        for (int p = 0; p < 2; ++p) { tma_prefetch_desc(&args.wr[p]); tma_prefetch_desc(&args.hr[p]); }
        for (int p = 0; p < NV; ++p) tma_prefetch_desc(&args.v[p]);
        for (int s = 0; s < 4; ++s) { mbar_init(&b_full[s], 1); mbar_init(&b_empty[s], CS); }
        for (int s = 0; s < 2; ++s) {
            mbar_init(&v_full[s], 1);   mbar_init(&v_empty[s], 8);
            mbar_init(&acc_full[s], 1); mbar_init(&acc_empty[s], 8);
        }
        mbar_init(a_full, 8);                             // one arrival per epilogue warp (W planes stored)
        mbar_init(a_empty, args.ntiles < 2 ? 1 : 2);      // one column tile per row: issuer A works alone
        mbar_init(go_b, 1);
        fence_mbar_init();
    }
    if (warp == 1) { tmem_alloc(tmem_slot, 512); tmem_relinquish(); }
    tc_fence_before();
    __syncthreads();
    if (CS > 1) cluster_sync_all();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    if (warp == 0) {
        if (lane == 0) {
            const int srow = 128 / CS;
            int bs = 0; uint32_t bph = 0;
            for (long t = t_begin; t < t_end; ++t) {
                const int sr = static_cast<int>(t / args.ntiles);
                const int tj = static_cast<int>(t - static_cast<long>(sr) * args.ntiles);
                // H^T chunks (the W planes of the row are stored to tensor memory by the epilogue warps)
                for (int kc = 0; kc < kchunks; ++kc) {
                    mbar_wait(&b_empty[bs], bph ^ 1);
                    mbar_expect_tx(&b_full[bs], L.b_stage_bytes);
                    for (int p = 0; p < 2; ++p) {
                        uint8_t* dst = b_smem + bs * L.b_stage_bytes + p * kBox + crank * srow * 128;
                        if (CS > 1)
                            tma_load_2d_mcast(dst, &args.hr[p], &b_full[bs], kc * 64, tj * 128 + crank * srow, cmask,
                                              kPolicyEvictLast);
                        else
                            tma_load_2d_hint(dst, &args.hr[p], &b_full[bs], kc * 64, tj * 128, kPolicyEvictLast);
                    }
                    if (++bs == L.nbs) { bs = 0; bph ^= 1; }
                }
            }
        }
    } else if (warp == 1) {
        // ---------------------------------------------------- MMA issuer A (even tiles) + dispatcher
        // Whole warp, converged: every operand is warp-uniform, one elected lane issues (ptx.cuh).
        // This warp performs ALL pipeline waits (W planes, accumulators, H^T stages) in ring order, for the
        // odd tiles too, and then hands those to issuer B through go_b.  Every mbarrier therefore has
        // exactly one waiter that sees every phase: with two independent waiters one of them skips the
        // other's phases, and mbarrier.try_wait.parity cannot tell phase p from phase p + 2 (observed as
        // MMAs on half-filled stages and, with both warps waiting on everything, as multi-tile stalls).
        constexpr uint32_t idesc = umma_idesc_bf16(128, 128);
        const uint64_t desc0 = umma_desc_k_sw128(0);
        int bs = 0; uint32_t bph = 0;
        uint32_t accph = 0;                     // bit b = parity of acc_empty[b]
        uint32_t aph = 0;
        const uint32_t a_tm = tmem_base + 256;   // W planes: hi at +0, lo at +KP/2 columns; 8 columns per K = 16
        const uint32_t b_addr = smem_u32(b_smem);
        const int b_stage = L.b_stage_bytes, nbs = L.nbs;
        int tj = static_cast<int>(t_begin % args.ntiles);
        bool new_row = true;
        int par = 0;
        const bool solo = args.ntiles < 2;      // rows of one tile would let B release a row before A is done
        for (long t = t_begin; t < t_end; ++t, par ^= 1) {
            if (new_row) {
                mbar_wait(a_full, aph);
                tc_fence_after();                   // the epilogue warps' tcgen05.st of the W planes
                aph ^= 1;
                new_row = false;
            }
            mbar_wait(&acc_empty[par], ((accph >> par) & 1u) ^ 1u);
            accph ^= 1u << par;
            if (par == 0 || solo) {
                tc_fence_after();
                const uint32_t d = tmem_base + par * 128;
#pragma unroll
                for (int kc = 0; kc < kchunks; ++kc) {
                    mbar_wait(&b_full[bs], bph);
                    tc_fence_after();
                    const uint64_t b_desc = desc0 + ((b_addr + bs * b_stage) >> 4);
#pragma unroll
                    for (int ks = 0; ks < 4; ++ks) {
                        const uint32_t a_hi = a_tm + kc * 32 + ks * 8;
                        const uint32_t a_lo = a_hi + KP / 2;
                        const uint64_t b_hi = b_desc + ((ks * 32) >> 4);
                        const uint64_t b_lo = b_desc + ((kBox + ks * 32) >> 4);
                        umma_bf16_ts_warp(d, a_hi, b_hi, idesc, (kc == 0 && ks == 0) ? 0u : 1u);
                        umma_bf16_ts_warp(d, a_hi, b_lo, idesc, 1);
                        umma_bf16_ts_warp(d, a_lo, b_hi, idesc, 1);
                    }
                    if (CS > 1) umma_commit_mcast_warp(&b_empty[bs], cmask);
                    else        umma_commit_warp(&b_empty[bs]);
                    if (++bs == nbs) { bs = 0; bph ^= 1; }
                }
                umma_commit_warp(&acc_full[par]);
            } else {
#pragma unroll
                for (int kc = 0; kc < kchunks; ++kc) {
                    mbar_wait(&b_full[bs], bph);
                    if (++bs == nbs) { bs = 0; bph ^= 1; }
                }
                __syncwarp();
                if (lane == 0) mbar_arrive(go_b);       // tile t: operands landed, accumulator 1 free
            }
            if (++tj == args.ntiles) { tj = 0; new_row = true; }
            // end of this CTA's tiles of the row: both issuers release the W planes (count 2); a warp
            // without MMAs in flight arrives at once
            if (t + 1 >= t_end || new_row) umma_commit_warp(a_empty);
        }
    } else if (warp == 11) {
        // ---------------------------------------------------- MMA issuer B (odd tiles): waits on go_b only
        constexpr uint32_t idesc = umma_idesc_bf16(128, 128);
        const uint64_t desc0 = umma_desc_k_sw128(0);
        int bs = 0;
        uint32_t goph = 0;
        const uint32_t a_tm = tmem_base + 256;   // W planes: hi at +0, lo at +KP/2 columns; 8 columns per K = 16
        const uint32_t b_addr = smem_u32(b_smem);
        const int b_stage = L.b_stage_bytes, nbs = L.nbs;
        const uint32_t d = tmem_base + 128;
        int tj = static_cast<int>(t_begin % args.ntiles);
        int par = 0;
        for (long t = (args.ntiles < 2 ? t_end : t_begin); t < t_end; ++t, par ^= 1) {
            if (par == 1) {
                mbar_wait(go_b, goph);
                goph ^= 1;
                tc_fence_after();
            }
#pragma unroll
            for (int kc = 0; kc < kchunks; ++kc) {
                if (par == 1) {
                    const uint64_t b_desc = desc0 + ((b_addr + bs * b_stage) >> 4);
#pragma unroll
                    for (int ks = 0; ks < 4; ++ks) {
                        const uint32_t a_hi = a_tm + kc * 32 + ks * 8;
                        const uint32_t a_lo = a_hi + KP / 2;
                        const uint64_t b_hi = b_desc + ((ks * 32) >> 4);
                        const uint64_t b_lo = b_desc + ((kBox + ks * 32) >> 4);
                        umma_bf16_ts_warp(d, a_hi, b_hi, idesc, (kc == 0 && ks == 0) ? 0u : 1u);
                        umma_bf16_ts_warp(d, a_hi, b_lo, idesc, 1);
                        umma_bf16_ts_warp(d, a_lo, b_hi, idesc, 1);
                    }
                    if (CS > 1) umma_commit_mcast_warp(&b_empty[bs], cmask);
                    else        umma_commit_warp(&b_empty[bs]);
                }
                if (++bs == nbs) bs = 0;
            }
            if (par == 1) umma_commit_warp(&acc_full[1]);
            bool row_end = false;
            if (++tj == args.ntiles) { tj = 0; row_end = true; }
            if (t + 1 >= t_end || row_end) umma_commit_warp(a_empty);
        }
    } else if (warp == 10) {
        if (lane == 0) {
            int vs = 0; uint32_t vph = 0;
            for (long t = t_begin; t < t_end; ++t) {
                const int sr = static_cast<int>(t / args.ntiles);
                const int tj = static_cast<int>(t - static_cast<long>(sr) * args.ntiles);
                const int ti = min(sr * CS + static_cast<int>(crank), args.mtiles - 1);
                // pull the V tile of a later step into L2: only two V buffers fit in shared memory, far
                // too few bytes in flight to cover HBM latency on their own
                {
                    const long tp = t + kPrefetchTiles;
                    if (tp < t_end) {
                        const int srp = static_cast<int>(tp / args.ntiles);
                        const int tjp = static_cast<int>(tp - static_cast<long>(srp) * args.ntiles);
                        const int tip = min(srp * CS + static_cast<int>(crank), args.mtiles - 1);
                        for (int p = 0; p < NV; ++p)
                            for (int hb = 0; hb < 2; ++hb)
                                tma_prefetch_2d(&args.v[p], 0, (tjp * 2 + hb) * args.v_rows_p + tip * 128);
                    }
                }
                mbar_wait(&v_empty[vs], vph ^ 1);
                mbar_expect_tx(&v_full[vs], L.v_buf_bytes);
                for (int p = 0; p < NV; ++p)
                    for (int hb = 0; hb < 2; ++hb)
                        tma_load_2d_hint(v_smem + vs * L.v_buf_bytes + (p * 2 + hb) * kBox, &args.v[p],
                                         &v_full[vs], 0, (tj * 2 + hb) * args.v_rows_p + ti * 128, kPolicyEvictFirst);
                if (++vs == L.nvs) { vs = 0; vph ^= 1; }
            }
        }
    } else {
        // ------------------------------------------------------------ epilogue
        const int ew = warp - 2;
        const int q = warp & 3;                   // TMEM lane quarter
        const int hb = ew >> 2;                   // which 64-column half of the tile
        const int row = q * 32 + lane;
        int acc = 0; uint32_t accph = 0;
        int vs = 0; uint32_t vph = 0;
        double total = 0.0;
        int cur_row = -1; uint32_t a_eph = 0;
        for (long t = t_begin; t < t_end; ++t) {
            const bool live = static_cast<int>(t / args.ntiles) * CS + static_cast<int>(crank) < args.mtiles;
            {
                // First tile of a tile row (or of this CTA's run): this warp stores its 32 rows of one W plane
                // (hb = plane) to tensor memory.  It has to happen BEFORE the wait for this tile's accumulator --
                // the MMAs of the tile need the planes.  a_empty: the MMAs of the previous row have retired
                // (both issuers commit to it at a row's end).  No warp can fall two phases behind on a_empty:
                // the next row's MMAs, hence its a_empty phase, need all eight arrivals on a_full.
                const int sr = static_cast<int>(t / args.ntiles);
                if (sr != cur_row) {
                    cur_row = sr;
                    const int ti = min(sr * CS + static_cast<int>(crank), args.mtiles - 1);
                    mbar_wait(a_empty, a_eph ^ 1);
                    a_eph ^= 1;
                    tc_fence_after();
                    const uint4* src = reinterpret_cast<const uint4*>(
                        args.w_rpl[hb] + (static_cast<long>(ti) * 128 + row) * KP);
                    const uint32_t abase = tmem_base + (static_cast<uint32_t>(q * 32) << 16) + 256 + hb * (KP / 2);
#pragma unroll
                    for (int c = 0; c < KP / 64; ++c) {       // 32 columns = 64 bf16 = 128 B of the row per store
                        uint32_t w[32];
#pragma unroll
                        for (int j = 0; j < 8; ++j) {
                            const uint4 x = __ldg(src + c * 8 + j);
                            w[4 * j] = x.x; w[4 * j + 1] = x.y; w[4 * j + 2] = x.z; w[4 * j + 3] = x.w;
                        }
                        tmem_st32(abase + c * 32, w);
                    }
                    tmem_st_wait();
                    tc_fence_before();
                    __syncwarp();
                    if (lane == 0) mbar_arrive(a_full);
                }
            }
            mbar_wait(&acc_full[acc], accph);
            tc_fence_after();
            const uint32_t taddr = tmem_base + (static_cast<uint32_t>(q * 32) << 16) + acc * 128 + hb * 64;
            // Both halves of this warp's 64 columns go to registers first and the accumulator is handed
            // back at once: the issuer of tile t + 2 waits for the TMEM read-out only (64 B/cycle/SM, i.e.
            // ~1000 cycles per tile), not for the arithmetic below.
            uint32_t s[2][32];
            tmem_ld32(taddr, s[0]);
            tmem_ld32(taddr + 32, s[1]);
            tmem_ld_wait();
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(&acc_empty[acc]);
            mbar_wait(&v_full[vs], vph);
            const uint8_t* vrow = v_smem + vs * L.v_buf_bytes + hb * kBox + row * 128;
            float ts[4] = {0.f, 0.f, 0.f, 0.f};          // independent chains
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
                    if (NV > 1) {
                        const uint4 rl = *reinterpret_cast<const uint4*>(vrow + 2 * kBox + chunk * 16);
                        const uint32_t l4[4] = {rl.x, rl.y, rl.z, rl.w};
#pragma unroll
                        for (int e = 0; e < 4; ++e) {
                            vv[2 * e] += __uint_as_float(l4[e] << 16);
                            vv[2 * e + 1] += __uint_as_float(l4[e] & 0xFFFF0000u);
                        }
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
            if (live) total += static_cast<double>((ts[0] + ts[1]) + (ts[2] + ts[3]));
            __syncwarp();
            if (lane == 0) mbar_arrive(&v_empty[vs]);
            if (++acc == 2) { acc = 0; accph ^= 1; }
            if (++vs == L.nvs) { vs = 0; vph ^= 1; }
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

// Resident clusters of `cs` CTAs for this kernel (one CTA per SM); cached per cluster size.
int residual_max_clusters(int cs, int num_sms) {
    static int cache[9] = {0, 0, 0, 0, 0, 0, 0, 0, 0};
    if (cs == 1) return num_sms;
    if (cache[cs] > 0) return cache[cs];
    void (*fn)(const ResidualArgs) = residual_ts_kernel<2, 1>;
    cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024);
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(num_sms / cs * cs);
    cfg.blockDim = dim3(kThreadsT);
    cfg.dynamicSmemBytes = 200 * 1024;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = cs; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr; cfg.numAttrs = 1;
    int n = 0;
    if (cudaOccupancyMaxActiveClusters(&n, fn, &cfg) != cudaSuccess || n <= 0) {
        cudaGetLastError();
        n = num_sms / cs * 3 / 4;
    }
    if (n > num_sms / cs) n = num_sms / cs;
    cache[cs] = n;
    return n;
}

cudaError_t residual_ts_launch(const ResidualArgs& args, int grid, cudaStream_t stream) {
    const int smem = residual_ts_smem_bytes(args.KP, args.NV);
    if (smem < 0 || args.BN != 128 || !args.w_rpl[0] || !args.w_rpl[1]) return cudaErrorInvalidValue;
    const int kch = args.KP / 64;
    void (*fn)(const ResidualArgs) = nullptr;
    if (args.NV == 1) fn = kch == 1 ? residual_ts_kernel<1, 1> : (kch == 2 ? residual_ts_kernel<2, 1> : residual_ts_kernel<4, 1>);
    else              fn = kch == 1 ? residual_ts_kernel<1, 2> : (kch == 2 ? residual_ts_kernel<2, 2> : residual_ts_kernel<4, 2>);
    cudaError_t e = cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024);
    if (e != cudaSuccess) return e;
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(grid);                       // a multiple of args.cluster
    cfg.blockDim = dim3(kThreadsT);
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
