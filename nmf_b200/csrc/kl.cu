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
// One shared-memory tile of Y's component-major planes ([c][t], 128 c x 64 t) serves both MMAs: as
// the MN-major B operand of MMA1 (K = c runs over the rows, N = t is contiguous) and as the K-major
// A operand of MMA2.  fp32-grade arithmetic comes from bf16 hi/lo planes: 3 MMAs per product.
//
// Warp roles (384 threads):
//   warp 0       TMA producer of X (once per work item) and of the Y ring (NSY stages)
//   warp 10      TMA producer of the D ring (2 stages) + L2 prefetch of the data stream
//   warp 1       issues MMA1 (whole warp, one elected lane per instruction: ptx.cuh)
//   warp 11      issues MMA2
//   warps 2..9   quotient epilogue, objective terms, Out drain
// Rings and issuers are separate on purpose.  The first version had one producer, one issuer and two
// combined (Y, D) stages: a stage could be refilled only after MMA2 of its tile had retired, the lone
// issuer sat in the wait for that refill while the next Q was already there, and every tcgen05.mma
// issued from an `if (lane == 0)` region cost ~130 cycles of R2UR waterfall against 32 cycles of N = 64
// tensor work: 32 % tensor-pipe activity (profiles/r01b_all_paths.md).
// Out is accumulated in chunks of kOutChunk tiles that are drained into register sums (see
// panel_gemm.cu for why), and written once per work item into a partial slot.
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

// Y ring depth: what is left of 224 KB after X (2 KCH boxes), the D ring and one Q buffer (hi + lo).
__host__ __device__ constexpr int kl_nsy(int KCH, int NV) {
    const int left = (224 * 1024 - 2 * KCH * kBoxK - kNSD * NV * kBoxK - 2 * kBoxK) / (2 * kBoxK);
    return left > 4 ? 4 : left;
}

// MN-major SWIZZLE_128B descriptor template (N = 64 contiguous elements per row, 8 rows per 1024 B atom).
__device__ __forceinline__ uint64_t desc_mn_sw128_base() { return umma_desc_mn_sw128(0, 16); }
}  // namespace

template <int KCH, int NV, bool OBJ>
__global__ void __launch_bounds__(kThreadsK, 1)
kl_panel_kernel(const __grid_constant__ KlArgs args) {
    extern __shared__ __align__(1024) uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) &
                                               ~static_cast<uintptr_t>(1023));
    constexpr int KP = KCH * 64;
    constexpr int X_BYTES = 2 * KCH * kBoxK;
    constexpr int Y_STAGE = 2 * kBoxK;               // Y hi, Y lo
    constexpr int D_STAGE = NV * kBoxK;
    constexpr int NSY = kl_nsy(KCH, NV);
    constexpr int Q_BYTES = 2 * kBoxK;
    static_assert(NSY >= 2, "Y ring too shallow");
    uint8_t* x_smem = smem;
    uint8_t* y_smem = x_smem + X_BYTES;
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
        mbar_init(x_full, 1);
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

    if (warp == 0) {
        // ------------------------------------------------------------ TMA producer: X and the Y ring
        if (lane == 0) {
            uint32_t xph = 0;
            int stg = 0; uint32_t sph = 0;
            for (int item = blockIdx.x; item < nitems; item += gridDim.x) {
                const int split = item / args.npanels;
                const int panel = item - split * args.npanels;
                const int t0 = split * args.tiles_per_split;
                const int t1 = min(t0 + args.tiles_per_split, args.ntiles);
                mbar_wait(x_empty, xph ^ 1);
                mbar_expect_tx(x_full, X_BYTES);
#pragma unroll
                for (int p = 0; p < 2; ++p)
#pragma unroll
                    for (int kc = 0; kc < KCH; ++kc)
                        tma_load_2d_hint(x_smem + (p * KCH + kc) * kBoxK, &args.xr[p], x_full, kc * 64,
                                         panel * 128, kPolicyEvictLast);
                xph ^= 1;
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
        const uint64_t dk0 = umma_desc_k_sw128(0);
        const uint64_t dmn0 = desc_mn_sw128_base();
        const uint64_t x_desc = dk0 + (smem_u32(x_smem) >> 4);
        const uint32_t y_addr = smem_u32(y_smem);
        uint32_t xph = 0;
        int stg = 0; uint32_t yph = 0;
        int sb = 0; uint32_t seph = 0;
        for (int item = blockIdx.x; item < nitems; item += gridDim.x) {
            const int split = item / args.npanels;
            const int t0 = split * args.tiles_per_split;
            const int t1 = min(t0 + args.tiles_per_split, args.ntiles);
            mbar_wait(x_full, xph); xph ^= 1;
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
                        const uint64_t a_hi = x_desc + (((0 * KCH + kc) * kBoxK + ks * 32) >> 4);
                        const uint64_t a_lo = x_desc + (((1 * KCH + kc) * kBoxK + ks * 32) >> 4);
                        const uint32_t roff = ((kc * 64 + ks * 16) * 128) >> 4;     // 16 component rows down
                        umma_bf16_warp(d, a_hi, y_hi + roff, idesc1, (kc == 0 && ks == 0) ? 0u : 1u);
                        umma_bf16_warp(d, a_hi, y_lo + roff, idesc1, 1u);
                        umma_bf16_warp(d, a_lo, y_hi + roff, idesc1, 1u);
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
        for (int item = blockIdx.x; item < nitems; item += gridDim.x) {
            const int split = item / args.npanels;
            const int panel = item - split * args.npanels;
            const int t0 = split * args.tiles_per_split;
            const int t1 = min(t0 + args.tiles_per_split, args.ntiles);
            const int nt = t1 - t0;
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

int kl_smem_bytes(int KP, int NV) {
    const int kch = KP / 64;
    return 2 * kch * kBoxK + kl_nsy(kch, NV) * 2 * kBoxK + kNSD * NV * kBoxK + 2 * kBoxK + 1024 + 512;
}

int kl_plan(int KP, int NV, int R, int T, int num_sms, KlPlan* plan) {
    if (KP != 64 && KP != 128) return -1;            // k <= 128 in this build
    if (NV < 1 || NV > 2) return -1;
    if (kl_smem_bytes(KP, NV) > 227 * 1024) return -2;
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
    plan->smem_bytes = kl_smem_bytes(KP, NV);
    const int nitems = npanels * nsplit;
    plan->grid = nitems < num_sms ? nitems : num_sms;
    return 0;
}

cudaError_t kl_panel_launch(const KlArgs& args, const KlPlan& plan, int NV, int num_sms, cudaStream_t stream) {
    (void)num_sms;
    if (const char* e = getenv("NMFB_KL_TS")) {      // opt-in experimental variants (kl_ts.cu, kl_ts2.cu)
        const int v = atoi(e);
        if (v == 2 && args.KP == 128) return kl_panel_ts2_launch(args, plan, NV, stream);
        if (v == 1 || v == 2) return kl_panel_ts_launch(args, plan, NV, stream);
    }
    void (*fn)(const KlArgs) = nullptr;
    const bool obj = args.obj_partial != nullptr;
    if (args.KP == 64)
        fn = (NV == 1) ? (obj ? kl_panel_kernel<1, 1, true> : kl_panel_kernel<1, 1, false>)
                       : (obj ? kl_panel_kernel<1, 2, true> : kl_panel_kernel<1, 2, false>);
    else if (args.KP == 128)
        fn = (NV == 1) ? (obj ? kl_panel_kernel<2, 1, true> : kl_panel_kernel<2, 1, false>)
                       : (obj ? kl_panel_kernel<2, 2, true> : kl_panel_kernel<2, 2, false>);
    else return cudaErrorInvalidValue;
    cudaError_t e = cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024);
    if (e != cudaSuccess) return e;
    fn<<<plan.grid, kThreadsK, plan.smem_bytes, stream>>>(args);
    count_launch();
    return cudaGetLastError();
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
