// Data step of the Kullback-Leibler branch of ADMM / AO-ADMM (admm.py:309-313, ao_admm.py:88-94), fused:
//     P      = A . B            (w_aux @ h_aux, or w @ h_aux / w_aux @ h for AO-ADMM)   tile by tile in TMEM
//     v_bar  = P - dual_v
//     v_aux  = ((v_bar - 1) + sqrt((v_bar - 1)^2 + 4 V)) / 2
//     dual_v = dual_v + v_aux - P
//     D      = v_aux + dual_v   (the data term of the next right-hand sides), and |max| of D per row and column
// P never reaches HBM (round 1: cuBLAS SGEMM into an m x n fp32 buffer, then an element-wise kernel, then a |max|
// pass: ~30 B per element of V before the re-ingest; here 14: V 2, dual_v 4 + 4, D 4).
//
// The right-hand sides A^T D are sensitive to ~1e-8 relative noise, so P must be fp32-grade:
// both operands enter as THREE bf16 planes (24-bit significands, nmfb_row_planes3) and every 128 x 128 tile is six
// MMA chains (p0.p0 + p0.p1 + p1.p0 + p0.p2 + p2.p0 + p1.p1; the dropped terms are below 2^-24).  The tile loop is
// residual_ts.cu's: the A planes of the tile row live in tensor memory (tcgen05.mma with a TMEM A operand), the B
// stages stream through shared memory (multicast across a CTA pair), V tiles come from their own producer warp, two
// issuer warps alternate tiles, eight epilogue warps own 32 rows x 64 columns each.  The fixed-point planes of D
// need the row / column maxima of ALL of D first, so the plane stores are a second pass (nmfb_ingest_data_fixed).
#include "kernels.cuh"
#include "ptx.cuh"

namespace nmfb {

namespace {
constexpr int kThreadsD = 384;
constexpr int kBox = 128 * 128;      // bytes of one [128 rows x 64 bf16] box
constexpr int kPrefetchD = 6;        // L2 prefetch distance of the V stream, in 128 x 128 tiles

// 16-byte global load issued where it is written (the epilogue reads dual_v and writes it back in place).
__device__ __forceinline__ float4 ld_global_f4(const float* p) {
    float4 v;
    asm volatile("ld.global.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p) : "memory");
    return v;
}

// sqrt and quotient in fp64 from an fp32 MUFU seed and Newton steps (relative error ~1e-14): the correctly-rounded
// library routines cost ~25 instructions each and made the epilogue -- 58 instructions per element at 0.24 IPC, ncu
// r02f -- the pace-setter of a kernel that should be bound by its 14 B per element of HBM traffic.  Branch-free on
// purpose: the eight elements of a group are independent chains that the scheduler interleaves; a first version
// with early-out branches lost that and ran 30 % slower than the library calls.  Arguments are data-scale
// (|x| within fp32 range); x <= 0 gives 0.
__device__ __forceinline__ double sqrt_newton(double x) {
    const double y = static_cast<double>(rsqrtf(fmaxf(static_cast<float>(x), 1e-37f)));     // ~2^-22
    double r = x * y;
    r = fma(0.5 * y, fma(-r, r, x), r);                                                      // ~1e-14
    return x > 0.0 ? r : 0.0;
}
__device__ __forceinline__ double div_newton(double a, double b) {      // a / b for b > 0 (callers select otherwise)
    double y = static_cast<double>(__frcp_rn(fmaxf(static_cast<float>(b), 1e-37f)));
    y = fma(y, fma(-b, y, 1.0), y);
    y = fma(y, fma(-b, y, 1.0), y);                                                          // ~1e-15
    return a * y;
}

struct KdLayout {
    int b_stage_bytes, v_buf_bytes;
    int nbs, nvs;
    int total;
};

__host__ __device__ inline KdLayout kd_layout(int KP, int NV) {
    KdLayout l;
    (void)KP;                            // the A planes live in tensor memory
    l.b_stage_bytes = 3 * kBox;
    l.v_buf_bytes = NV * 2 * kBox;
    const int budget = 227 * 1024 - 1024 - 512;
    int rest = budget;
    l.nvs = (rest - 2 * l.v_buf_bytes >= 2 * l.b_stage_bytes) ? 2 : 1;
    rest -= l.nvs * l.v_buf_bytes;
    l.nbs = rest / l.b_stage_bytes;
    if (l.nbs > 4) l.nbs = 4;
    l.total = l.nbs * l.b_stage_bytes + l.nvs * l.v_buf_bytes;
    return l;
}
}  // namespace

int kl_data_smem_bytes(int KP, int NV) {
    if (KP != 64 && KP != 128) return -1;            // 3 x KP/2 plane columns + two accumulators must fit 512
    const KdLayout l = kd_layout(KP, NV);
    if (l.nbs < 2) return -1;
    return l.total + 1024 + 512;
}

template <int KCH, int NV>
__global__ void __launch_bounds__(kThreadsD, 1)
kl_data_kernel(const __grid_constant__ KlDataArgs args) {
    if (args.done_flag != nullptr && *args.done_flag != 0) return;    // inner loop has broken: D and dual_v stay
    extern __shared__ __align__(1024) uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) &
                                               ~static_cast<uintptr_t>(1023));
    constexpr int KP = KCH * 64;
    constexpr int kchunks = KCH;
    const KdLayout L = kd_layout(KP, NV);
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
        for (int p = 0; p < 3; ++p) tma_prefetch_desc(&args.br[p]);
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
                    for (int p = 0; p < 3; ++p) {
                        uint8_t* dst = b_smem + bs * L.b_stage_bytes + p * kBox + crank * srow * 128;
                        if (CS > 1)
                            tma_load_2d_mcast(dst, &args.br[p], &b_full[bs], kc * 64, tj * 128 + crank * srow, cmask,
                                              kPolicyEvictLast);
                        else
                            tma_load_2d_hint(dst, &args.br[p], &b_full[bs], kc * 64, tj * 128, kPolicyEvictLast);
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
        const uint32_t a_tm = tmem_base + 256;   // A planes: plane p at +p*KP/2 columns; 8 columns per K = 16
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
                        const uint32_t a0 = a_tm + kc * 32 + ks * 8;
                        const uint32_t a1 = a0 + KP / 2, a2 = a0 + KP;
                        const uint64_t b0 = b_desc + ((ks * 32) >> 4);
                        const uint64_t b1 = b_desc + ((kBox + ks * 32) >> 4);
                        const uint64_t b2 = b_desc + ((2 * kBox + ks * 32) >> 4);
                        umma_bf16_ts_warp(d, a0, b0, idesc, (kc == 0 && ks == 0) ? 0u : 1u);
                        umma_bf16_ts_warp(d, a0, b1, idesc, 1);
                        umma_bf16_ts_warp(d, a1, b0, idesc, 1);
                        umma_bf16_ts_warp(d, a0, b2, idesc, 1);
                        umma_bf16_ts_warp(d, a2, b0, idesc, 1);
                        umma_bf16_ts_warp(d, a1, b1, idesc, 1);
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
        const uint32_t a_tm = tmem_base + 256;   // A planes: plane p at +p*KP/2 columns; 8 columns per K = 16
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
                        const uint32_t a0 = a_tm + kc * 32 + ks * 8;
                        const uint32_t a1 = a0 + KP / 2, a2 = a0 + KP;
                        const uint64_t b0 = b_desc + ((ks * 32) >> 4);
                        const uint64_t b1 = b_desc + ((kBox + ks * 32) >> 4);
                        const uint64_t b2 = b_desc + ((2 * kBox + ks * 32) >> 4);
                        umma_bf16_ts_warp(d, a0, b0, idesc, (kc == 0 && ks == 0) ? 0u : 1u);
                        umma_bf16_ts_warp(d, a0, b1, idesc, 1);
                        umma_bf16_ts_warp(d, a1, b0, idesc, 1);
                        umma_bf16_ts_warp(d, a0, b2, idesc, 1);
                        umma_bf16_ts_warp(d, a2, b0, idesc, 1);
                        umma_bf16_ts_warp(d, a1, b1, idesc, 1);
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
                    const long tp = t + kPrefetchD;
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
                    for (int pl = hb; pl < 3; pl += 2) {       // hb = 0: planes 0 and 2, hb = 1: plane 1
                        const uint4* src = reinterpret_cast<const uint4*>(
                            args.a_rpl[pl] + (static_cast<long>(ti) * 128 + row) * KP);
                        const uint32_t abase = tmem_base + (static_cast<uint32_t>(q * 32) << 16) + 256 + pl * (KP / 2);
#pragma unroll
                        for (int c = 0; c < KP / 64; ++c) {   // 32 columns = 64 bf16 = 128 B of the row per store
                            uint32_t w[32];
#pragma unroll
                            for (int j = 0; j < 8; ++j) {
                                const uint4 x = __ldg(src + c * 8 + j);
                                w[4 * j] = x.x; w[4 * j + 1] = x.y; w[4 * j + 2] = x.z; w[4 * j + 3] = x.w;
                            }
                            tmem_st32(abase + c * 32, w);
                        }
                    }
                    tmem_st_wait();
                    tc_fence_before();
                    __syncwarp();
                    if (lane == 0) mbar_arrive(a_full);
                }
            }
            // v_aux / dual_v step (admm.py:309-313, ao_admm.py:88-94) on this thread's row, 64 columns:
            //   v_bar = P - dual_v;  v_aux = ((v_bar - 1) + sqrt((v_bar - 1)^2 + 4 v)) / 2  (cancellation-free form
            //   for v_bar < 1, fp64);  dual_v += v_aux - P;  D = v_aux + dual_v  -> fp32, plus |max| of D per row / column
            // The kernel is bound by the 12 B per element of dual_v / D traffic, not by the MMAs: the 32 dual_v values of
            // a column half are requested BEFORE the wait for the accumulator / the TMEM read-out, so that 128 B per
            // thread (32 KB per SM) are in flight.
            const int ti_e = min(static_cast<int>(t / args.ntiles) * CS + static_cast<int>(crank), args.mtiles - 1);
            const int tj_e = static_cast<int>(t - (t / args.ntiles) * args.ntiles);
            const long grow = static_cast<long>(ti_e) * 128 + row;                  // row of this rank's shard
            const int col0 = tj_e * 128 + hb * 64;
            // dual_v and D are K-blocked like V ([col / 64][rows_p][64] fp32): this thread's 64 columns are 256
            // contiguous bytes and the warp's 32 rows one 8 KB run -- with a row-major pitch of 64 KB every warp
            // instruction touched 32 DRAM pages (2 TB/s, profiles/r02d)
            const long boff = (static_cast<long>(col0 >> 6) * args.rows_p + grow) * 64;
            float* dvp = args.dual_v + boff;
            float* dop = args.d_out + boff;
            const uint32_t taddr = tmem_base + (static_cast<uint32_t>(q * 32) << 16) + acc * 128 + hb * 64;
            unsigned int rmax = 0, cmine = 0, cmine2 = 0;
            const uint8_t* vrow = nullptr;
#pragma unroll
            for (int g = 0; g < 2; ++g) {
                float4 dq[8];
#pragma unroll
                for (int j = 0; j < 8; ++j)
                    dq[j] = live ? ld_global_f4(dvp + g * 32 + j * 4) : make_float4(0.f, 0.f, 0.f, 0.f);
                if (g == 0) {
                    mbar_wait(&acc_full[acc], accph);
                    tc_fence_after();
                }
                uint32_t s[32];
                tmem_ld32(taddr + g * 32, s);
                tmem_ld_wait();
                if (g == 1) {                                  // both halves read: the accumulator goes back
                    tc_fence_before();
                    __syncwarp();
                    if (lane == 0) mbar_arrive(&acc_empty[acc]);
                } else {
                    mbar_wait(&v_full[vs], vph);
                    vrow = v_smem + vs * L.v_buf_bytes + hb * kBox + row * 128;
                }
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
                    const int cofs = g * 32 + c8 * 8;
                    const float dv[8] = {dq[2 * c8].x, dq[2 * c8].y, dq[2 * c8].z, dq[2 * c8].w,
                                         dq[2 * c8 + 1].x, dq[2 * c8 + 1].y, dq[2 * c8 + 1].z, dq[2 * c8 + 1].w};
                    float dn[8], dd[8];
#pragma unroll
                    for (int e = 0; e < 8; ++e) {
                        const double P = static_cast<double>(__uint_as_float(s[c8 * 8 + e]));
                        const double v = static_cast<double>(vv[e]);
                        const double tq = P - static_cast<double>(dv[e]) - 1.0;
                        const double root = sqrt_newton(fma(tq, tq, 4.0 * v));
                        const double den = root - tq;
                        const double va_neg = den > 0.0 ? div_newton(2.0 * v, den) : 0.0;
                        const double va = tq >= 0.0 ? 0.5 * (tq + root) : va_neg;
                        dn[e] = static_cast<float>(static_cast<double>(dv[e]) + va - P);
                        dd[e] = static_cast<float>(va + static_cast<double>(dn[e]));
                        const unsigned int bits = live ? __float_as_uint(fabsf(dd[e])) : 0u;
                        rmax = max(rmax, bits);
                        // column maximum over the warp's 32 rows; lane j keeps column (cofs + e) when j == that column % 32
                        const unsigned int cm = __reduce_max_sync(0xffffffffu, bits);
                        if (((cofs + e) & 31) == lane) { if (g == 0) cmine = cm; else cmine2 = cm; }
                    }
                    if (live) {
                        *reinterpret_cast<float4*>(dvp + cofs) = make_float4(dn[0], dn[1], dn[2], dn[3]);
                        *reinterpret_cast<float4*>(dvp + cofs + 4) = make_float4(dn[4], dn[5], dn[6], dn[7]);
                        *reinterpret_cast<float4*>(dop + cofs) = make_float4(dd[0], dd[1], dd[2], dd[3]);
                        *reinterpret_cast<float4*>(dop + cofs + 4) = make_float4(dd[4], dd[5], dd[6], dd[7]);
                    }
                }
            }
            if (live) {
                if (rmax) atomicMax(args.rowmax + args.row_origin + grow, rmax);
                if (cmine) atomicMax(args.colmax + col0 + lane, cmine);
                if (cmine2) atomicMax(args.colmax + col0 + 32 + lane, cmine2);
            }
            __syncwarp();
            if (lane == 0) mbar_arrive(&v_empty[vs]);
            if (++acc == 2) { acc = 0; accph ^= 1; }
            if (++vs == L.nvs) { vs = 0; vph ^= 1; }
        }
        (void)ew;
    }

    tc_fence_before();
    __syncthreads();
    if (CS > 1) cluster_sync_all();
    if (warp == 1) {
        tc_fence_after();
        tmem_dealloc(tmem_base, 512);
    }
}

cudaError_t kl_data_launch(const KlDataArgs& args, int grid, cudaStream_t stream) {
    const int smem = kl_data_smem_bytes(args.KP, args.NV);
    if (smem < 0) return cudaErrorInvalidValue;
    void (*fn)(const KlDataArgs) = nullptr;
    if (args.NV == 1) fn = args.KP == 64 ? kl_data_kernel<1, 1> : kl_data_kernel<2, 1>;
    else              fn = args.KP == 64 ? kl_data_kernel<1, 2> : kl_data_kernel<2, 2>;
    cudaError_t e = cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024);
    if (e != cudaSuccess) return e;
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(grid);                       // a multiple of args.cluster
    cfg.blockDim = dim3(kThreadsD);
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
