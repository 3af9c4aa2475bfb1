// Fused GEMM -> quotient -> GEMM kernel of the Kullback-Leibler updates, "TS" variant: the factor being updated
// (X, the resident 128-row panel) lives in TENSOR MEMORY and is the A operand of MMA1 (S = X . Y^T).
//
// STATUS: the KP = 64 kernel of the product (KP = 128 takes kl_ts2.cu).  Parity on B200: GPUTEST_r01, round-2 A/B in
// profiles/r02a_*.  Relative to the SS-form kernel of round 1 (deleted):
//   1. TMEM: S buffers 2 x 64 columns, Out buffers 2 x 128, and -- new -- the X planes in the last 128 columns
//      (128 lanes = rows of the panel, two bf16 K-values per column, hi then lo, KP / 2 columns each).  The eight
//      epilogue warps store them with tcgen05.st at the start of every work item, gated by x_empty / x_full.
//   2. MMA1 in the [d], [a_tmem], b_desc form.  Its N = 64 instructions fetched 4 KB of A + 2 KB of B per 32
//      cycles from shared memory (192 B/cycle against a port of 128 B/cycle); now 2 KB (64 B/cycle).  Per 128 x 64
//      tile the kernel moved 336 KB through the shared-memory port (MMA1 144, MMA2 96, Q stores 32, Y / D fills 48,
//      D reads 16) = 2690 cycles at best -- the 1.39 ms of the H-side pass at 20000 x 50000 (827 tiles per SM at
//      1965 MHz) are ~3300 cycles per tile against 1536 cycles of MMA work: the default kernel runs at ~80 % of
//      the port limit, not of the tensor pipe (profiles/r01e_smem_operand_model.md); this variant moves 240 KB.
//   3. The 64 KB of shared memory X took go to the Y ring (4 stages).
//
#include "kernels.cuh"
#include "kl.cuh"
#include "ptx.cuh"

namespace nmfb {

namespace {
constexpr int kThreadsK = 384;
constexpr int kBoxK = 128 * 128;       // bytes of one [128 rows x 64 bf16] box
constexpr int kOutChunk = 8;           // t-tiles accumulated in TMEM before a drain
constexpr int kPrefetchK = 6;          // L2 prefetch distance of the data stream, in t-tiles
constexpr int kNSD = 2;                // D ring depth

// Y ring depth: what is left of 224 KB after the D ring and one Q buffer (hi + lo); X lives in tensor memory.
__host__ __device__ constexpr int kl_ts_nsy(int NV) {
    const int left = (224 * 1024 - kNSD * NV * kBoxK - 2 * kBoxK) / (2 * kBoxK);
    return left > 4 ? 4 : left;
}

// MN-major SWIZZLE_128B descriptor template (N = 64 contiguous elements per row, 8 rows per 1024 B atom).
__device__ __forceinline__ uint64_t desc_mn_sw128_base() { return umma_desc_mn_sw128(0, 16); }
}  // namespace

template <int KCH, int NV, bool OBJ>
__global__ void __launch_bounds__(kThreadsK, 1)
kl_panel_ts_kernel(const __grid_constant__ KlArgs args) {
    extern __shared__ __align__(1024) uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) &
                                               ~static_cast<uintptr_t>(1023));
    constexpr int KP = KCH * 64;
    constexpr int Y_STAGE = 2 * kBoxK;               // Y hi, Y lo
    constexpr int D_STAGE = NV * kBoxK;
    constexpr int NSY = kl_ts_nsy(NV);
    constexpr int Q_BYTES = 2 * kBoxK;
    static_assert(NSY >= 2, "Y ring too shallow");
    uint8_t* y_smem = smem;
    uint8_t* d_smem = y_smem + NSY * Y_STAGE;
    uint8_t* q_smem = d_smem + kNSD * D_STAGE;
    uint64_t* bars = reinterpret_cast<uint64_t*>(q_smem + Q_BYTES);
    uint64_t* x_full = bars + 0;
    uint64_t* x_empty = bars + 1;
    uint64_t* y_full = bars + 2;       // [4]
    uint64_t* y_empty = bars + 6;      // [4]
    uint64_t* d_full = bars + 10;      // [2]
    uint64_t* d_empty = bars + 12;     // [2]
    uint64_t* s_full = bars + 14;      // [2]  S tile in TMEM ready
    uint64_t* s_empty = bars + 16;     // [2]
    uint64_t* q_full = bars + 18;
    uint64_t* q_empty = bars + 19;
    uint64_t* o_full = bars + 20;      // [2]  Out chunk ready
    uint64_t* o_empty = bars + 22;     // [2]
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 24);
    double* red = reinterpret_cast<double*>(bars + 26);      // [8]

    const int warp = warp_idx_uniform(), lane = threadIdx.x & 31;
    const int nitems = args.npanels * args.nsplit;

    if (warp == 0 && lane == 0) {
        for (int p = 0; p < 2; ++p) { tma_prefetch_desc(&args.xr[p]); tma_prefetch_desc(&args.yc[p]); }
        for (int p = 0; p < NV; ++p) tma_prefetch_desc(&args.d[p]);
        mbar_init(x_full, 8);          // one arrival per epilogue warp (its rows of one X plane stored to TMEM)
        mbar_init(x_empty, 1);
        for (int s = 0; s < 4; ++s) { mbar_init(&y_full[s], 1); mbar_init(&y_empty[s], 1); }
        for (int s = 0; s < 2; ++s) {
            mbar_init(&d_full[s], 1);   mbar_init(&d_empty[s], 8);
            mbar_init(&s_full[s], 1);   mbar_init(&s_empty[s], 8);
            mbar_init(&o_full[s], 1);   mbar_init(&o_empty[s], 8);
        }
        mbar_init(q_full, 8);
        mbar_init(q_empty, 1);
        fence_mbar_init();
    }
    if (warp == 1) { tmem_alloc(tmem_slot, 512); tmem_relinquish(); }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    const uint32_t tm_s = tmem_base;            // S buffers: 2 x 64 columns
    const uint32_t tm_o = tmem_base + 128;      // Out buffers: 2 x 128 columns
    const uint32_t tm_x = tmem_base + 384;      // X planes (A operand of MMA1): hi, then lo, KP / 2 columns each

    if (warp == 0) {
        // ------------------------------------------------------------ TMA producer: X and the Y ring
        if (lane == 0) {
            int stg = 0; uint32_t sph = 0;
            for (int item = blockIdx.x; item < nitems; item += gridDim.x) {
                const int split = item / args.npanels;
                const int t0 = split * args.tiles_per_split;
                const int t1 = min(t0 + args.tiles_per_split, args.ntiles);
                for (int t = t0; t < t1; ++t) {
                    mbar_wait(&y_empty[stg], sph ^ 1);
                    uint8_t* st = y_smem + stg * Y_STAGE;
                    mbar_expect_tx(&y_full[stg], Y_STAGE);
#pragma unroll
                    for (int p = 0; p < 2; ++p)
                        tma_load_2d_hint(st + p * kBoxK, &args.yc[p], &y_full[stg], t * 64, 0, kPolicyEvictLast);
                    if (++stg == NSY) { stg = 0; sph ^= 1; }
                }
            }
        }
    } else if (warp == 10) {
        // ------------------------------------------------------------ TMA producer: the D ring
        if (lane == 0) {
            int ds = 0; uint32_t dph = 0;
            for (int item = blockIdx.x; item < nitems; item += gridDim.x) {
                const int split = item / args.npanels;
                const int panel = item - split * args.npanels;
                const int t0 = split * args.tiles_per_split;
                const int t1 = min(t0 + args.tiles_per_split, args.ntiles);
                for (int t = t0; t < t1; ++t) {
                    const int tp = t + kPrefetchK;
                    if (tp < t1)
                        for (int p = 0; p < NV; ++p)
                            tma_prefetch_2d(&args.d[p], 0, tp * args.d_rows_p + panel * 128);
                    mbar_wait(&d_empty[ds], dph ^ 1);
                    uint8_t* st = d_smem + ds * D_STAGE;
                    mbar_expect_tx(&d_full[ds], D_STAGE);
#pragma unroll
                    for (int p = 0; p < NV; ++p)
                        tma_load_2d_hint(st + p * kBoxK, &args.d[p], &d_full[ds], 0,
                                         t * args.d_rows_p + panel * 128, kPolicyEvictFirst);
                    if (++ds == kNSD) { ds = 0; dph ^= 1; }
                }
            }
        }
    } else if (warp == 1) {
        // ------------------------------------------------------------ MMA1 issuer: S = X . Y^T
        constexpr uint32_t idesc1 = umma_idesc_bf16(128, 64, 0, 1);     // B (Y tile) is MN-major
        const uint64_t dmn0 = desc_mn_sw128_base();
        const uint32_t y_addr = smem_u32(y_smem);
        uint32_t xph = 0;
        int stg = 0; uint32_t yph = 0;
        int sb = 0; uint32_t seph = 0;
        for (int item = blockIdx.x; item < nitems; item += gridDim.x) {
            const int split = item / args.npanels;
            const int t0 = split * args.tiles_per_split;
            const int t1 = min(t0 + args.tiles_per_split, args.ntiles);
            mbar_wait(x_full, xph); xph ^= 1;
            tc_fence_after();                      // the epilogue warps' tcgen05.st of the X planes
            for (int t = t0; t < t1; ++t) {
                mbar_wait(&y_full[stg], yph);
                mbar_wait(&s_empty[sb], seph ^ 1);
                tc_fence_after();
                const uint32_t yb = y_addr + stg * Y_STAGE;
                const uint64_t y_hi = dmn0 + (yb >> 4), y_lo = dmn0 + ((yb + kBoxK) >> 4);
                const uint32_t d = tm_s + sb * 64;
#pragma unroll
                for (int kc = 0; kc < KCH; ++kc)
#pragma unroll
                    for (int ks = 0; ks < 4; ++ks) {
                        const uint32_t a_hi = tm_x + kc * 32 + ks * 8;        // 8 columns per K = 16 step
                        const uint32_t a_lo = a_hi + KP / 2;
                        const uint32_t roff = ((kc * 64 + ks * 16) * 128) >> 4;     // 16 component rows down
                        umma_bf16_ts_warp(d, a_hi, y_hi + roff, idesc1, (kc == 0 && ks == 0) ? 0u : 1u);
                        umma_bf16_ts_warp(d, a_hi, y_lo + roff, idesc1, 1u);
                        umma_bf16_ts_warp(d, a_lo, y_hi + roff, idesc1, 1u);
                    }
                umma_commit_warp(&s_full[sb]);
                if (++stg == NSY) { stg = 0; yph ^= 1; }
                if (++sb == 2) { sb = 0; seph ^= 1; }
            }
            umma_commit_warp(x_empty);
        }
    } else if (warp == 11) {
        // ------------------------------------------------------------ MMA2 issuer: Out += Y^T . Q^T
        constexpr uint32_t idesc2 = umma_idesc_bf16(128, 128);
        const uint64_t dk0 = umma_desc_k_sw128(0);
        const uint32_t y_addr = smem_u32(y_smem);
        const uint64_t q_hi = dk0 + (smem_u32(q_smem) >> 4);
        const uint64_t q_lo = q_hi + (kBoxK >> 4);
        int stg = 0; uint32_t yph = 0;
        uint32_t qph = 0;
        uint32_t oe_ph = 0;                        // bit b = parity of o_empty[b]
        int ob = 0;
        for (int item = blockIdx.x; item < nitems; item += gridDim.x) {
            const int split = item / args.npanels;
            const int t0 = split * args.tiles_per_split;
            const int t1 = min(t0 + args.tiles_per_split, args.ntiles);
            const int nt = t1 - t0;
            int jc = 0;
            for (int j = 0; j < nt; ++j) {
                mbar_wait(&y_full[stg], yph);               // long complete; orders this warp after the TMA writes
                mbar_wait(q_full, qph); qph ^= 1;
                if (jc == 0) { mbar_wait(&o_empty[ob], ((oe_ph >> ob) & 1u) ^ 1u); oe_ph ^= 1u << ob; }
                tc_fence_after();
                const uint32_t yb = y_addr + stg * Y_STAGE;
                const uint64_t y_hi = dk0 + (yb >> 4), y_lo = dk0 + ((yb + kBoxK) >> 4);
                const uint32_t d = tm_o + ob * 128;
#pragma unroll
                for (int ks = 0; ks < 4; ++ks) {
                    const uint32_t o = (ks * 32) >> 4;
                    umma_bf16_warp(d, y_hi + o, q_hi + o, idesc2, (jc == 0 && ks == 0) ? 0u : 1u);
                    umma_bf16_warp(d, y_hi + o, q_lo + o, idesc2, 1u);
                    umma_bf16_warp(d, y_lo + o, q_hi + o, idesc2, 1u);
                }
                umma_commit_warp(&y_empty[stg]);
                umma_commit_warp(q_empty);
                if (++jc == kOutChunk || j == nt - 1) { umma_commit_warp(&o_full[ob]); ob ^= 1; jc = 0; }
                if (++stg == NSY) { stg = 0; yph ^= 1; }
            }
        }
    } else {
        // ------------------------------------------------------------ quotient epilogue + Out drain
        const int ew = warp - 2;
        const int q = warp & 3;                      // TMEM lane quarter
        const int half = ew >> 2;
        const int row = q * 32 + lane;               // S lane = data row r; Out lane = component c
        int sb = 0; uint32_t sfph = 0;               // S buffer / D stage walk together (both 2 deep)
        uint32_t qeph = 1;                           // first wait on the fresh q_empty passes
        uint32_t of_ph = 0;                          // bit b = parity of o_full[b]
        int ob = 0;
        double total = 0.0;
        uint32_t x_eph = 0;
        for (int item = blockIdx.x; item < nitems; item += gridDim.x) {
            const int split = item / args.npanels;
            const int panel = item - split * args.npanels;
            const int t0 = split * args.tiles_per_split;
            const int t1 = min(t0 + args.tiles_per_split, args.ntiles);
            const int nt = t1 - t0;
            {
                // This warp's 32 rows of one X plane (half = plane) go to tensor memory before anything of the
                // item is waited for (MMA1 of the first tile needs them).  x_empty: the MMA1s of the previous
                // item have retired.  No warp can fall two phases behind on x_empty: the next item's MMA1s,
                // hence its x_empty phase, need all eight arrivals on x_full.
                mbar_wait(x_empty, x_eph ^ 1);
                x_eph ^= 1;
                tc_fence_after();
                const uint4* src = reinterpret_cast<const uint4*>(
                    args.x_rpl[half] + (static_cast<long>(panel) * 128 + row) * KP);
                const uint32_t xbase = tm_x + (static_cast<uint32_t>(q * 32) << 16) + half * (KP / 2);
#pragma unroll
                for (int c = 0; c < KP / 64; ++c) {           // 32 columns = 64 bf16 = 128 B of the row per store
                    uint32_t w[32];
#pragma unroll
                    for (int j = 0; j < 8; ++j) {
                        const uint4 x = __ldg(src + c * 8 + j);
                        w[4 * j] = x.x; w[4 * j + 1] = x.y; w[4 * j + 2] = x.z; w[4 * j + 3] = x.w;
                    }
                    tmem_st32(xbase + c * 32, w);
                }
                tmem_st_wait();
                tc_fence_before();
                __syncwarp();
                if (lane == 0) mbar_arrive(x_full);
            }
            float sums[64];
#pragma unroll
            for (int e = 0; e < 64; ++e) sums[e] = 0.f;
            int jc = 0;
            for (int j = 0; j < nt; ++j) {
                mbar_wait(&s_full[sb], sfph);
                mbar_wait(&d_full[sb], sfph);
                tc_fence_after();
                uint32_t s[32];
                tmem_ld32(tm_s + (static_cast<uint32_t>(q * 32) << 16) + sb * 64 + half * 32, s);
                tmem_ld_wait();
                const uint8_t* drow = d_smem + sb * D_STAGE + row * 128;
                uint32_t hi4[4][4], lo4[4][4];
                float tsum[2] = {0.f, 0.f};                 // objective terms (OBJ), two chains
#pragma unroll
                for (int c8 = 0; c8 < 4; ++c8) {
                    const int chunk = (half * 4 + c8) ^ (row & 7);          // SWIZZLE_128B
                    const uint4 raw = *reinterpret_cast<const uint4*>(drow + chunk * 16);
                    const uint32_t w4[4] = {raw.x, raw.y, raw.z, raw.w};
                    float dv[8];
#pragma unroll
                    for (int e = 0; e < 4; ++e) {
                        dv[2 * e] = __uint_as_float(w4[e] << 16);
                        dv[2 * e + 1] = __uint_as_float(w4[e] & 0xFFFF0000u);
                    }
                    if (NV > 1) {
                        const uint4 rl = *reinterpret_cast<const uint4*>(drow + kBoxK + chunk * 16);
                        const uint32_t l4[4] = {rl.x, rl.y, rl.z, rl.w};
#pragma unroll
                        for (int e = 0; e < 4; ++e) {
                            dv[2 * e] += __uint_as_float(l4[e] << 16);
                            dv[2 * e + 1] += __uint_as_float(l4[e] & 0xFFFF0000u);
                        }
                    }
#pragma unroll
                    for (int e = 0; e < 4; ++e) {
                        float qq[2];
#pragma unroll
                        for (int u = 0; u < 2; ++u) {
                            const float v = dv[2 * e + u];
                            const float sv = __uint_as_float(s[c8 * 8 + 2 * e + u]);
                            const float r = rcp_approx(sv + 1e-9f);          // mur.py:25 / :41  x / (wh + 1e-9)
                            qq[u] = v * r;
                            if (OBJ) tsum[u] += kl_term_r(v, sv, r);
                        }
                        __nv_bfloat16 h0, l0, h1, l1;
                        split2(qq[0], h0, l0);
                        split2(qq[1], h1, l1);
                        hi4[c8][e] = static_cast<uint32_t>(__bfloat16_as_ushort(h0)) |
                                     (static_cast<uint32_t>(__bfloat16_as_ushort(h1)) << 16);
                        lo4[c8][e] = static_cast<uint32_t>(__bfloat16_as_ushort(l0)) |
                                     (static_cast<uint32_t>(__bfloat16_as_ushort(l1)) << 16);
                    }
                }
                if (OBJ) total += static_cast<double>(tsum[0] + tsum[1]);
                // S and D are consumed: hand both back before waiting for the Q buffer
                tc_fence_before();
                __syncwarp();
                if (lane == 0) { mbar_arrive(&s_empty[sb]); mbar_arrive(&d_empty[sb]); }
                mbar_wait(q_empty, qeph); qeph ^= 1;            // MMA2 of the previous tile has read Q
                uint8_t* qrow = q_smem + row * 128;
#pragma unroll
                for (int c8 = 0; c8 < 4; ++c8) {
                    const int chunk = (half * 4 + c8) ^ (row & 7);
                    *reinterpret_cast<uint4*>(qrow + chunk * 16) =
                        make_uint4(hi4[c8][0], hi4[c8][1], hi4[c8][2], hi4[c8][3]);
                    *reinterpret_cast<uint4*>(qrow + kBoxK + chunk * 16) =
                        make_uint4(lo4[c8][0], lo4[c8][1], lo4[c8][2], lo4[c8][3]);
                }
                fence_async_smem();                  // generic-proxy writes of Q -> visible to the MMA
                __syncwarp();
                if (lane == 0) mbar_arrive(q_full);
                if (++sb == 2) { sb = 0; sfph ^= 1; }
                if (++jc == kOutChunk || j == nt - 1) {
                    jc = 0;
                    mbar_wait(&o_full[ob], (of_ph >> ob) & 1u); of_ph ^= 1u << ob;
                    tc_fence_after();
                    const uint32_t taddr = tm_o + (static_cast<uint32_t>(q * 32) << 16) + ob * 128 + half * 64;
#pragma unroll
                    for (int gq = 0; gq < 2; ++gq) {
                        uint32_t r[32];
                        tmem_ld32(taddr + gq * 32, r);
                        tmem_ld_wait();
#pragma unroll
                        for (int e = 0; e < 32; ++e) sums[gq * 32 + e] += __uint_as_float(r[e]);
                    }
                    tc_fence_before();
                    __syncwarp();
                    if (lane == 0) mbar_arrive(&o_empty[ob]);
                    ob ^= 1;
                }
            }
            if (row < KP) {
                float* dst = args.out + (static_cast<size_t>(split) * KP + row) * args.Lp +
                             static_cast<size_t>(panel) * 128 + half * 64;
#pragma unroll
                for (int v = 0; v < 16; ++v)
                    reinterpret_cast<float4*>(dst)[v] =
                        make_float4(sums[4 * v], sums[4 * v + 1], sums[4 * v + 2], sums[4 * v + 3]);
            }
        }
        if (OBJ) {
            for (int o = 16; o; o >>= 1) total += __shfl_xor_sync(0xffffffffu, total, o);
            if (lane == 0) red[ew] = total;
        }
    }

    tc_fence_before();
    __syncthreads();
    if (warp == 1) {
        tc_fence_after();
        tmem_dealloc(tmem_base, 512);
    }
    if (OBJ && threadIdx.x == 0) {
        double t = 0.0;
        for (int w = 0; w < 8; ++w) t += red[w];
        args.obj_partial[blockIdx.x] = t;
    }
}

int kl_ts_smem_bytes(int NV) {
    return kl_ts_nsy(NV) * 2 * kBoxK + kNSD * NV * kBoxK + 2 * kBoxK + 1024 + 512;
}

cudaError_t kl_panel_ts_launch(const KlArgs& args, const KlPlan& plan, int NV, cudaStream_t stream) {
    if (!args.x_rpl[0] || !args.x_rpl[1]) return cudaErrorInvalidValue;
    void (*fn)(const KlArgs) = nullptr;
    const bool obj = args.obj_partial != nullptr;
    if (args.KP == 64)
        fn = (NV == 1) ? (obj ? kl_panel_ts_kernel<1, 1, true> : kl_panel_ts_kernel<1, 1, false>)
                       : (obj ? kl_panel_ts_kernel<1, 2, true> : kl_panel_ts_kernel<1, 2, false>);
    else if (args.KP == 128)
        fn = (NV == 1) ? (obj ? kl_panel_ts_kernel<2, 1, true> : kl_panel_ts_kernel<2, 1, false>)
                       : (obj ? kl_panel_ts_kernel<2, 2, true> : kl_panel_ts_kernel<2, 2, false>);
    else return cudaErrorInvalidValue;
    cudaError_t e = cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024);
    if (e != cudaSuccess) return e;
    fn<<<plan.grid, kThreadsK, kl_ts_smem_bytes(NV), stream>>>(args);
    count_launch();
    return cudaGetLastError();
}

}  // namespace nmfb
