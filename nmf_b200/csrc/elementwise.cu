// CUDA-core kernels around the tensor-core contractions: data/factor ingest (precision
// splitting + layout), Gram reduction, the fused MUR-Euclidean update, a CUDA-core residual
// objective (small problems / cross-check) and factor export.
//
// Layout vocabulary:
//   master  : fp32  [KP][Lp]        component-major factor (W^T or H), the solver state
//   cplanes : bf16  [3][256][Lp]    component-major hi/lo/lo2 planes  (GEMM A operand, Gram)
//   rplanes : bf16  [2][Lp][KP]     row-major hi/lo planes            (residual kernel operands)
//   data    : bf16  [NV][rows_p][ld] row-major planes of V or V^T
#include "elementwise.cuh"
#include "ptx.cuh"

namespace nmfb {

// ------------------------------------------------------------------ ingest of V
// One 64x64 tile per block.  Reads the caller's chunk (fp32 or fp64, row-major), writes the
// row-major planes directly and the transposed planes through shared memory, and folds
// min / "needs a lo plane" / non-finite flags into `stats`.
template <typename T>
__global__ void __launch_bounds__(256)
ingest_data_kernel(const T* __restrict__ src, long src_ld, int row0, int nrows, int ncols,
                   __nv_bfloat16* __restrict__ rowp, long rowp_plane, int rowp_rows_p,
                   __nv_bfloat16* __restrict__ colp, long colp_plane, int colp_rows_p, int nv,
                   unsigned int* __restrict__ stats, const unsigned int* __restrict__ rowmax,
                   const unsigned int* __restrict__ colmax, int row_lo, int row_n, int col_lo, int col_n) {
    // Sharded runs: this rank keeps rows [row_lo, row_lo + row_n) of V in the row copy and columns
    // [col_lo, col_lo + col_n) in the transposed copy (both stored from local index 0); the flags are
    // folded over everything it reads, so that every rank reaches the same decisions.
    __shared__ float tile[64][65];
    const int tx = threadIdx.x & 63, ty = threadIdx.x >> 6;      // 64 x 4
    const int c0 = blockIdx.x * 64, r0 = blockIdx.y * 64;
    float vmin = INFINITY;
    unsigned int need_lo = 0, bad = 0;
    for (int rr = ty; rr < 64; rr += 4) {
        const int r = r0 + rr, c = c0 + tx;
        float x = 0.f;
        if (r < nrows && c < ncols) {
            // src_ld < 0: K-blocked source [cols / 64][-src_ld][64] (the data term of the KL branch, kl_data.cu)
            const long si = src_ld >= 0 ? static_cast<long>(r) * src_ld + c
                                        : (static_cast<long>(c >> 6) * (-src_ld) + r) * 64 + (c & 63);
            x = static_cast<float>(src[si]);
            vmin = fminf(vmin, x);
            if (!isfinite(x)) bad = 1;
            __nv_bfloat16 hi, lo, lo2 = __float2bfloat16_rn(0.f);
            if (rowmax != nullptr) {                       // fixed-point plane 0 + floating remainder(s)
                float u, iu;
                fixed_unit(rowmax[row0 + r], u, iu);
                const float v0 = fixed_slice(x, u, iu), rem = x - v0;
                hi = __float2bfloat16_rn(v0);
                lo = __float2bfloat16_rn(rem);
                float left = rem - __bfloat162float(lo);
                if (nv > 2) {                              // third plane (KL data term of ADMM: arbitrary fp32 values)
                    lo2 = __float2bfloat16_rn(left);
                    left -= __bfloat162float(lo2);
                }
                if (left != 0.f) need_lo = 1;              // the planes do not hold x exactly
            } else {
                split2(x, hi, lo);
                if (__bfloat162float(lo) != 0.f) need_lo = 1;
            }
            const int lr = row0 + r - row_lo;
            if (lr >= 0 && lr < row_n) {
                const long o = blocked_index(lr, c, rowp_rows_p);
                rowp[o] = hi;
                if (nv > 1) rowp[rowp_plane + o] = lo;
                if (nv > 2) rowp[2 * rowp_plane + o] = lo2;
            }
        }
        tile[rr][tx] = x;
    }
    __syncthreads();
    for (int cc = ty; cc < 64; cc += 4) {
        const int c = c0 + cc, r = r0 + tx;
        if (r < nrows && c < ncols && c >= col_lo && c < col_lo + col_n) {
            __nv_bfloat16 hi, lo, lo2 = __float2bfloat16_rn(0.f);
            const float x = tile[tx][cc];
            if (colmax != nullptr) {
                float u, iu;
                fixed_unit(colmax[c], u, iu);
                const float v0 = fixed_slice(x, u, iu), rem = x - v0;
                hi = __float2bfloat16_rn(v0);
                lo = __float2bfloat16_rn(rem);
                float left = rem - __bfloat162float(lo);
                if (nv > 2) {
                    lo2 = __float2bfloat16_rn(left);
                    left -= __bfloat162float(lo2);
                }
                if (left != 0.f) need_lo = 1;
            } else {
                split2(x, hi, lo);
            }
            const long o = blocked_index(c - col_lo, row0 + r, colp_rows_p);
            colp[o] = hi;
            if (nv > 1) colp[colp_plane + o] = lo;
            if (nv > 2) colp[2 * colp_plane + o] = lo2;
        }
    }
    // block-level fold
    for (int o = 16; o; o >>= 1) {
        vmin = fminf(vmin, __shfl_xor_sync(0xffffffffu, vmin, o));
        need_lo |= __shfl_xor_sync(0xffffffffu, need_lo, o);
        bad |= __shfl_xor_sync(0xffffffffu, bad, o);
    }
    if ((threadIdx.x & 31) == 0) {
        const unsigned int bits = __float_as_uint(vmin);
        const unsigned int key = (bits & 0x80000000u) ? ~bits : (bits | 0x80000000u);
        atomicMin(&stats[0], key);
        if (need_lo) atomicOr(&stats[1], 1u);
        if (bad) atomicOr(&stats[2], 1u);
    }
}

cudaError_t ingest_data_launch(const void* src, int is_f64, long src_ld, int row0, int nrows, int ncols,
                               __nv_bfloat16* rowp, long rowp_plane, int rowp_rows_p, __nv_bfloat16* colp,
                               long colp_plane, int colp_rows_p, int nv, unsigned int* stats,
                               const unsigned int* rowmax, const unsigned int* colmax, int row_lo, int row_n,
                               int col_lo, int col_n, cudaStream_t stream) {
    dim3 grid((ncols + 63) / 64, (nrows + 63) / 64);
    if (grid.x == 0 || grid.y == 0) return cudaSuccess;
    if (is_f64)
        ingest_data_kernel<double><<<grid, 256, 0, stream>>>(static_cast<const double*>(src), src_ld, row0,
                                                            nrows, ncols, rowp, rowp_plane, rowp_rows_p, colp,
                                                            colp_plane, colp_rows_p, nv, stats, rowmax, colmax,
                                                            row_lo, row_n, col_lo, col_n);
    else
        ingest_data_kernel<float><<<grid, 256, 0, stream>>>(static_cast<const float*>(src), src_ld, row0,
                                                           nrows, ncols, rowp, rowp_plane, rowp_rows_p, colp,
                                                           colp_plane, colp_rows_p, nv, stats, rowmax, colmax,
                                                            row_lo, row_n, col_lo, col_n);
    count_launch();
    return cudaGetLastError();
}

// ------------------------------------------------------------------ factor planes from master
// Re-derives every bf16 plane from the fp32 master (used after ingest of an initial factor and
// by solvers whose update kernel does not write planes itself).
// adopt != nullptr: the values come from `master` (an exchange buffer of the same [KP][Lp] layout that the ranks'
// windows were stored into, peer.cu) and are also written to `adopt`, the factor's own master.
// Rows [skip_lo, skip_hi) are left alone (a rank's own window, whose update kernel wrote master and planes itself).
// grid = (Lp / 256, KP / 16): a thread owns one r and 16 components -- 32-byte runs of the row-major planes, and
// enough CTAs in flight to cover the latency of the strided master reads (one thread per r over all components was
// the pace-setter of the sharded exchange: 443 threads per SM, each a serial chain of 128 loads).
__global__ void __launch_bounds__(256)
refresh_planes_kernel(const float* __restrict__ master, float* __restrict__ adopt, int KP, int Lp,
                      __nv_bfloat16* __restrict__ cpl, long cpl_plane, __nv_bfloat16* __restrict__ rpl,
                      long rpl_plane, int skip_lo, int skip_hi) {
    const int r = blockIdx.x * 256 + threadIdx.x;
    if (r >= Lp || (r >= skip_lo && r < skip_hi)) return;
    const int c0 = blockIdx.y * 16;
    float x[16];
#pragma unroll
    for (int q = 0; q < 16; ++q) x[q] = master[static_cast<long>(c0 + q) * Lp + r];
    __align__(16) __nv_bfloat16 h16[16], l16[16];
#pragma unroll
    for (int q = 0; q < 16; ++q) {
        const long o = static_cast<long>(c0 + q) * Lp + r;
        if (adopt) adopt[o] = x[q];
        __nv_bfloat16 p0, p1, p2;
        split3(x[q], p0, p1, p2);
        cpl[o] = p0;
        cpl[cpl_plane + o] = p1;
        cpl[2 * cpl_plane + o] = p2;
        h16[q] = p0;
        l16[q] = p1;
    }
    const long ro = static_cast<long>(r) * KP + c0;
    *reinterpret_cast<uint4*>(rpl + ro) = *reinterpret_cast<const uint4*>(h16);
    *reinterpret_cast<uint4*>(rpl + ro + 8) = *reinterpret_cast<const uint4*>(h16 + 8);
    *reinterpret_cast<uint4*>(rpl + rpl_plane + ro) = *reinterpret_cast<const uint4*>(l16);
    *reinterpret_cast<uint4*>(rpl + rpl_plane + ro + 8) = *reinterpret_cast<const uint4*>(l16 + 8);
}

cudaError_t refresh_planes_launch(const float* master, float* adopt, int KP, int Lp, __nv_bfloat16* cpl,
                                  long cpl_plane, __nv_bfloat16* rpl, long rpl_plane, int skip_lo, int skip_hi,
                                  cudaStream_t stream) {
    refresh_planes_kernel<<<dim3((Lp + 255) / 256, KP / 16), 256, 0, stream>>>(master, adopt, KP, Lp, cpl, cpl_plane,
                                                                             rpl, rpl_plane, skip_lo, skip_hi);
    count_launch();
    return cudaGetLastError();
}

// Three row-major bf16 planes [3][Lp][KP] of a master (24-bit significands): operands of the KL data step of
// ADMM / AO-ADMM (kl_data.cu), whose product has to be fp32-grade.
__global__ void __launch_bounds__(256)
row_planes3_kernel(const float* __restrict__ master, int KP, int Lp, __nv_bfloat16* __restrict__ out) {
    const int r = blockIdx.x * 256 + threadIdx.x;
    if (r >= Lp) return;
    const long plane = static_cast<long>(Lp) * KP;
    for (int c0 = 0; c0 < KP; c0 += 8) {
        __align__(16) __nv_bfloat16 p0[8], p1[8], p2[8];
#pragma unroll
        for (int q = 0; q < 8; ++q) split3(master[static_cast<long>(c0 + q) * Lp + r], p0[q], p1[q], p2[q]);
        const long ro = static_cast<long>(r) * KP + c0;
        *reinterpret_cast<uint4*>(out + ro) = *reinterpret_cast<const uint4*>(p0);
        *reinterpret_cast<uint4*>(out + plane + ro) = *reinterpret_cast<const uint4*>(p1);
        *reinterpret_cast<uint4*>(out + 2 * plane + ro) = *reinterpret_cast<const uint4*>(p2);
    }
}

cudaError_t row_planes3_launch(const float* master, int KP, int Lp, __nv_bfloat16* out, cudaStream_t stream) {
    row_planes3_kernel<<<(Lp + 255) / 256, 256, 0, stream>>>(master, KP, Lp, out);
    count_launch();
    return cudaGetLastError();
}

// words[0..n) = 0 unless *done_flag != 0 (the |max| tables of the KL data step, whose kernel honours the same flag)
__global__ void __launch_bounds__(256)
zero_words_guarded_kernel(unsigned int* __restrict__ a, int na, unsigned int* __restrict__ b, int nb,
                          const int* __restrict__ done_flag) {
    if (done_flag != nullptr && *done_flag != 0) return;
    const int i = blockIdx.x * 256 + threadIdx.x;
    if (i < na) a[i] = 0u;
    if (i < nb) b[i] = 0u;
}

cudaError_t zero_words_guarded_launch(unsigned int* a, int na, unsigned int* b, int nb, const int* done_flag,
                                      cudaStream_t stream) {
    const int n = na > nb ? na : nb;
    zero_words_guarded_kernel<<<(n + 255) / 256, 256, 0, stream>>>(a, na, b, nb, done_flag);
    count_launch();
    return cudaGetLastError();
}

// Strided import of a caller factor (W [m][k] or H [k][n], fp32/fp64) into the fp32 master.
template <typename T>
__global__ void __launch_bounds__(256)
import_factor_kernel(const T* __restrict__ src, long stride_l, long stride_c, int L, int k,
                     float* __restrict__ master, int KP, int Lp) {
    const long idx = static_cast<long>(blockIdx.x) * 256 + threadIdx.x;
    if (idx >= static_cast<long>(KP) * Lp) return;
    const int c = static_cast<int>(idx / Lp), r = static_cast<int>(idx - static_cast<long>(c) * Lp);
    float x = 0.f;
    if (c < k && r < L) x = static_cast<float>(src[r * stride_l + c * stride_c]);
    master[idx] = x;
}

cudaError_t import_factor_launch(const void* src, int is_f64, long stride_l, long stride_c, int L, int k,
                                 float* master, int KP, int Lp, cudaStream_t stream) {
    const long total = static_cast<long>(KP) * Lp;
    const int grid = static_cast<int>((total + 255) / 256);
    if (is_f64)
        import_factor_kernel<double><<<grid, 256, 0, stream>>>(static_cast<const double*>(src), stride_l,
                                                              stride_c, L, k, master, KP, Lp);
    else
        import_factor_kernel<float><<<grid, 256, 0, stream>>>(static_cast<const float*>(src), stride_l,
                                                             stride_c, L, k, master, KP, Lp);
    count_launch();
    return cudaGetLastError();
}

template <typename T>
__global__ void __launch_bounds__(256)
export_factor_kernel(const float* __restrict__ master, int KP, int Lp, T* __restrict__ dst, long stride_l,
                     long stride_c, int L, int k) {
    const long idx = static_cast<long>(blockIdx.x) * 256 + threadIdx.x;
    if (idx >= static_cast<long>(k) * L) return;
    const int c = static_cast<int>(idx / L), r = static_cast<int>(idx - static_cast<long>(c) * L);
    dst[r * stride_l + c * stride_c] = static_cast<T>(master[static_cast<long>(c) * Lp + r]);
}

cudaError_t export_factor_launch(const float* master, int KP, int Lp, void* dst, int is_f64, long stride_l,
                                 long stride_c, int L, int k, cudaStream_t stream) {
    const long total = static_cast<long>(k) * L;
    const int grid = static_cast<int>((total + 255) / 256);
    if (is_f64)
        export_factor_kernel<double><<<grid, 256, 0, stream>>>(master, KP, Lp, static_cast<double*>(dst),
                                                              stride_l, stride_c, L, k);
    else
        export_factor_kernel<float><<<grid, 256, 0, stream>>>(master, KP, Lp, static_cast<float*>(dst),
                                                             stride_l, stride_c, L, k);
    count_launch();
    return cudaGetLastError();
}

// ------------------------------------------------------------------ Gram reduction
// G[c][c'] = sum over partial slots (fixed order, fp64) of P[s][c][c'];  P pitch = Lp floats.
__global__ void __launch_bounds__(256)
gram_reduce_kernel(const float* __restrict__ part, int slots, int KP, int Lp, float* __restrict__ g32,
                   double* __restrict__ g64) {
    const int idx = blockIdx.x * 256 + threadIdx.x;
    if (idx >= KP * KP) return;
    const int c = idx / KP, d = idx - c * KP;
    double acc = 0.0;
    for (int s = 0; s < slots; ++s)
        acc += static_cast<double>(part[(static_cast<long>(s) * KP + c) * Lp + d]);
    g32[idx] = static_cast<float>(acc);
    g64[idx] = acc;
}

cudaError_t gram_reduce_launch(const float* part, int slots, int KP, int Lp, float* g32, double* g64,
                               cudaStream_t stream) {
    gram_reduce_kernel<<<(KP * KP + 255) / 256, 256, 0, stream>>>(part, slots, KP, Lp, g32, g64);
    count_launch();
    return cudaGetLastError();
}

// ------------------------------------------------------------------ fused MUR-Euclidean update
//   x_new = x * num / ( (G x)  + lambda * x + 1e-9 )            mur.py:29 / mur.py:45
// num = sum of the GEMM partial slots, G = Gram of the *other* factor (so G x == (W H) H^T resp.
// W^T (W H) without ever forming W H).  One block owns 64 consecutive r (rows of the tall factor)
// and all KP components: a register-tiled (CPT x 4) fp32 product against G staged through smem,
// then the multiplicative update and all plane stores straight from registers.
template <int CPT>
__global__ void __launch_bounds__(256, CPT <= 8 ? 2 : 1)
mur_eu_update_kernel(const float* __restrict__ part, int slots, const float* __restrict__ gram,
                     float* __restrict__ master, __nv_bfloat16* __restrict__ cpl, long cpl_plane,
                     __nv_bfloat16* __restrict__ rpl, long rpl_plane, int k, int L, int Lp, float lambda) {
    constexpr int KP = CPT * 16;
    constexpr int TR = 64;
    constexpr int CK = 32;
    extern __shared__ float smem_f[];
    float* xs = smem_f;                 // [KP][TR]
    float* gs = smem_f + KP * TR;       // [CK][KP]
    const int tid = threadIdx.x;
    const int tr = tid & 15, tc = tid >> 4;
    const int r0 = blockIdx.x * TR;

    {
        constexpr int NV4 = KP * TR / 4 / 256;             // all loads of the tile in flight at once
        float4 xr[NV4];
#pragma unroll
        for (int j = 0; j < NV4; ++j) {
            const int i = tid + j * 256;
            const int c = i / (TR / 4), q = i - c * (TR / 4);
            xr[j] = *reinterpret_cast<const float4*>(master + static_cast<long>(c) * Lp + r0 + q * 4);
        }
#pragma unroll
        for (int j = 0; j < NV4; ++j) reinterpret_cast<float4*>(xs)[tid + j * 256] = xr[j];
    }
    float acc[CPT][4];
#pragma unroll
    for (int a = 0; a < CPT; ++a)
#pragma unroll
        for (int b = 0; b < 4; ++b) acc[a][b] = 0.f;

    // rows of G: chunk ck + 1 travels from L2 to registers while chunk ck is multiplied
    constexpr int NG = CK * KP / 4 / 256;
    float4 gr[NG];
#pragma unroll
    for (int j = 0; j < NG; ++j) gr[j] = *reinterpret_cast<const float4*>(gram + (tid + j * 256) * 4);
    for (int ck = 0; ck < KP; ck += CK) {
        __syncthreads();
#pragma unroll
        for (int j = 0; j < NG; ++j) reinterpret_cast<float4*>(gs)[tid + j * 256] = gr[j];
        if (ck + CK < KP) {
#pragma unroll
            for (int j = 0; j < NG; ++j)
                gr[j] = *reinterpret_cast<const float4*>(gram + static_cast<long>(ck + CK) * KP + (tid + j * 256) * 4);
        }
        __syncthreads();
#pragma unroll 4
        for (int cc = 0; cc < CK; ++cc) {
            const float4 xv = *reinterpret_cast<const float4*>(xs + (ck + cc) * TR + tr * 4);
            const float* grow = gs + cc * KP + tc * CPT;      // G symmetric: row (ck+cc), my columns
#pragma unroll
            for (int a = 0; a < CPT; a += 4) {
                const float4 g = *reinterpret_cast<const float4*>(grow + a);
                acc[a + 0][0] += g.x * xv.x; acc[a + 0][1] += g.x * xv.y; acc[a + 0][2] += g.x * xv.z; acc[a + 0][3] += g.x * xv.w;
                acc[a + 1][0] += g.y * xv.x; acc[a + 1][1] += g.y * xv.y; acc[a + 1][2] += g.y * xv.z; acc[a + 1][3] += g.y * xv.w;
                acc[a + 2][0] += g.z * xv.x; acc[a + 2][1] += g.z * xv.y; acc[a + 2][2] += g.z * xv.z; acc[a + 2][3] += g.z * xv.w;
                acc[a + 3][0] += g.w * xv.x; acc[a + 3][1] += g.w * xv.y; acc[a + 3][2] += g.w * xv.z; acc[a + 3][3] += g.w * xv.w;
            }
        }
    }

    // epilogue: numerator from the partial slots, multiplicative update, stores
    __align__(16) __nv_bfloat16 rh[4][CPT], rl[4][CPT];
    // numerator = sum of the GEMM's partial slots: the slots of component a + 1 are requested before the stores of
    // component a (up to kPre slots ahead; the waits for these loads were the largest stall of the kernel)
    constexpr int kPre = 4;
    const long slot_stride = static_cast<long>(KP) * Lp;
    float4 pre[kPre];
    {
        const long off0 = static_cast<long>(tc * CPT) * Lp + r0 + tr * 4;
#pragma unroll
        for (int s = 0; s < kPre; ++s)
            pre[s] = s < slots ? *reinterpret_cast<const float4*>(part + s * slot_stride + off0)
                               : make_float4(0.f, 0.f, 0.f, 0.f);
    }
#pragma unroll
    for (int a = 0; a < CPT; ++a) {
        const int c = tc * CPT + a;
        const long off = static_cast<long>(c) * Lp + r0 + tr * 4;
        float4 num = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
        for (int s = 0; s < kPre; ++s) { num.x += pre[s].x; num.y += pre[s].y; num.z += pre[s].z; num.w += pre[s].w; }
        if (a + 1 < CPT) {
#pragma unroll
            for (int s = 0; s < kPre; ++s)
                pre[s] = s < slots ? *reinterpret_cast<const float4*>(part + s * slot_stride + off + Lp)
                                   : make_float4(0.f, 0.f, 0.f, 0.f);
        }
        for (int s = kPre; s < slots; ++s) {
            const float4 p = *reinterpret_cast<const float4*>(part + s * slot_stride + off);
            num.x += p.x; num.y += p.y; num.z += p.z; num.w += p.w;
        }
        const float4 x = *reinterpret_cast<const float4*>(xs + c * TR + tr * 4);
        float xn[4];
        xn[0] = x.x * num.x / (acc[a][0] + lambda * x.x + 1e-9f);
        xn[1] = x.y * num.y / (acc[a][1] + lambda * x.y + 1e-9f);
        xn[2] = x.z * num.z / (acc[a][2] + lambda * x.z + 1e-9f);
        xn[3] = x.w * num.w / (acc[a][3] + lambda * x.w + 1e-9f);
        __align__(8) __nv_bfloat16 ch[4], cl[4];
#pragma unroll
        for (int b = 0; b < 4; ++b) {
            if (c >= k || r0 + tr * 4 + b >= L) xn[b] = 0.f;
            split2(xn[b], ch[b], cl[b]);
            rh[b][a] = ch[b];
            rl[b][a] = cl[b];
        }
        *reinterpret_cast<float4*>(master + off) = make_float4(xn[0], xn[1], xn[2], xn[3]);
        *reinterpret_cast<uint2*>(cpl + off) = *reinterpret_cast<const uint2*>(ch);
        *reinterpret_cast<uint2*>(cpl + cpl_plane + off) = *reinterpret_cast<const uint2*>(cl);
    }
#pragma unroll
    for (int b = 0; b < 4; ++b) {
        const long ro = static_cast<long>(r0 + tr * 4 + b) * KP + tc * CPT;
#pragma unroll
        for (int a = 0; a < CPT; a += 4) {
            *reinterpret_cast<uint2*>(rpl + ro + a) = *reinterpret_cast<const uint2*>(&rh[b][a]);
            *reinterpret_cast<uint2*>(rpl + rpl_plane + ro + a) = *reinterpret_cast<const uint2*>(&rl[b][a]);
        }
    }
}

cudaError_t mur_eu_update_launch(const float* part, int slots, const float* gram, float* master,
                                 __nv_bfloat16* cpl, long cpl_plane, __nv_bfloat16* rpl, long rpl_plane,
                                 int KP, int k, int L, int Lp, int nrows_p, float lambda, cudaStream_t stream) {
    const int grid = nrows_p / 64;
    const size_t smem = static_cast<size_t>(KP) * 64 * 4 + 32 * static_cast<size_t>(KP) * 4;
    cudaError_t e = cudaSuccess;
    if (KP == 64) {
        mur_eu_update_kernel<4><<<grid, 256, smem, stream>>>(part, slots, gram, master, cpl, cpl_plane, rpl,
                                                             rpl_plane, k, L, Lp, lambda);
    } else if (KP == 128) {
        e = cudaFuncSetAttribute(mur_eu_update_kernel<8>, cudaFuncAttributeMaxDynamicSharedMemorySize, 64 * 1024);
        if (e != cudaSuccess) return e;
        mur_eu_update_kernel<8><<<grid, 256, smem, stream>>>(part, slots, gram, master, cpl, cpl_plane, rpl,
                                                             rpl_plane, k, L, Lp, lambda);
    } else {
        e = cudaFuncSetAttribute(mur_eu_update_kernel<16>, cudaFuncAttributeMaxDynamicSharedMemorySize, 128 * 1024);
        if (e != cudaSuccess) return e;
        mur_eu_update_kernel<16><<<grid, 256, smem, stream>>>(part, slots, gram, master, cpl, cpl_plane, rpl,
                                                              rpl_plane, k, L, Lp, lambda);
    }
    if (e != cudaSuccess) return e;
    count_launch();
    return cudaGetLastError();
}

// ------------------------------------------------------------------ CUDA-core residual objective
// 0.5 * sum (v - w.h)^2 over a 64x64 tile per block from the fp32 masters (utils.py:29).
// O(mnk) on the FP32 pipe: used for small problems and to cross-check the tcgen05 residual kernel.
__global__ void __launch_bounds__(256)
objective_simt_kernel(const __nv_bfloat16* __restrict__ vpl, long v_plane, int v_rows_p, int nv,
                      const float* __restrict__ wt, int Mp, const float* __restrict__ ht, int Np, int k,
                      int m, int n, double* __restrict__ partial) {
    __shared__ float ws[16][64], hs[16][64];
    const int tid = threadIdx.x, tx = tid & 15, ty = tid >> 4;
    const int i0 = blockIdx.y * 64, j0 = blockIdx.x * 64;
    float acc[4][4] = {};
    for (int c0 = 0; c0 < k; c0 += 16) {
        __syncthreads();
        for (int i = tid; i < 16 * 64; i += 256) {
            const int c = i >> 6, o = i & 63;
            ws[c][o] = (c0 + c < k) ? wt[static_cast<long>(c0 + c) * Mp + i0 + o] : 0.f;
            hs[c][o] = (c0 + c < k) ? ht[static_cast<long>(c0 + c) * Np + j0 + o] : 0.f;
        }
        __syncthreads();
#pragma unroll
        for (int c = 0; c < 16; ++c) {
            const float4 a = *reinterpret_cast<const float4*>(&ws[c][ty * 4]);
            const float4 b = *reinterpret_cast<const float4*>(&hs[c][tx * 4]);
            const float av[4] = {a.x, a.y, a.z, a.w}, bv[4] = {b.x, b.y, b.z, b.w};
#pragma unroll
            for (int p = 0; p < 4; ++p)
#pragma unroll
                for (int q = 0; q < 4; ++q) acc[p][q] += av[p] * bv[q];
        }
    }
    double sum = 0.0;
#pragma unroll
    for (int p = 0; p < 4; ++p) {
        const int i = i0 + ty * 4 + p;
        float rowsum = 0.f;
#pragma unroll
        for (int q = 0; q < 4; ++q) {
            const int j = j0 + tx * 4 + q;
            if (i < m && j < n) {
                const long o = blocked_index(i, j, v_rows_p);
                float v = __bfloat162float(vpl[o]);
                float d = v - acc[p][q];
                if (nv > 1) d += __bfloat162float(vpl[v_plane + o]);
                rowsum += d * d;
            }
        }
        sum += static_cast<double>(rowsum);
    }
    __shared__ double red[8];
    for (int o = 16; o; o >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, o);
    if ((tid & 31) == 0) red[tid >> 5] = sum;
    __syncthreads();
    if (tid == 0) {
        double t = 0.0;
        for (int w = 0; w < 8; ++w) t += red[w];
        partial[blockIdx.y * gridDim.x + blockIdx.x] = t;
    }
}

cudaError_t objective_simt_launch(const __nv_bfloat16* vpl, long v_plane, int v_rows_p, int nv, const float* wt,
                                  int Mp, const float* ht, int Np, int k, int m, int n, double* partial,
                                  int* nparts, cudaStream_t stream) {
    dim3 grid((n + 63) / 64, (m + 63) / 64);
    *nparts = static_cast<int>(grid.x * grid.y);
    objective_simt_kernel<<<grid, 256, 0, stream>>>(vpl, v_plane, v_rows_p, nv, wt, Mp, ht, Np, k, m, n, partial);
    count_launch();
    return cudaGetLastError();
}


// ------------------------------------------------------------------ row / column |max| of a V chunk
// rowmax[row0 + r], colmax[c] = max |x| as float bits (monotone for non-negative floats).
template <typename T>
__global__ void __launch_bounds__(256)
data_absmax_kernel(const T* __restrict__ src, long src_ld, int row0, int nrows, int ncols,
                   unsigned int* __restrict__ rowmax, unsigned int* __restrict__ colmax) {
    const int tx = threadIdx.x & 63, ty = threadIdx.x >> 6;
    const int c = blockIdx.x * 64 + tx, r0 = blockIdx.y * 64;
    unsigned int cm = 0;
    for (int rr = ty; rr < 64; rr += 4) {
        const int r = r0 + rr;
        unsigned int b = 0;
        if (r < nrows && c < ncols)
            b = __float_as_uint(fabsf(static_cast<float>(src[static_cast<long>(r) * src_ld + c])));
        cm = max(cm, b);
        unsigned int rm = b;
        for (int o = 16; o; o >>= 1) rm = max(rm, __shfl_xor_sync(0xffffffffu, rm, o));
        if ((threadIdx.x & 31) == 0 && r < nrows) atomicMax(&rowmax[row0 + r], rm);
    }
    if (c < ncols) atomicMax(&colmax[c], cm);
}

cudaError_t data_absmax_launch(const void* src, int is_f64, long src_ld, int row0, int nrows, int ncols,
                               unsigned int* rowmax, unsigned int* colmax, cudaStream_t stream) {
    dim3 grid((ncols + 63) / 64, (nrows + 63) / 64);
    if (grid.x == 0 || grid.y == 0) return cudaSuccess;
    if (is_f64)
        data_absmax_kernel<double><<<grid, 256, 0, stream>>>(static_cast<const double*>(src), src_ld, row0, nrows,
                                                            ncols, rowmax, colmax);
    else
        data_absmax_kernel<float><<<grid, 256, 0, stream>>>(static_cast<const float*>(src), src_ld, row0, nrows,
                                                           ncols, rowmax, colmax);
    count_launch();
    return cudaGetLastError();
}

// ------------------------------------------------------------------ fixed-point factor planes
// One block per component row: |max| of master[c][0..Lp) by shuffles, then cplanes = (fixed slice,
// bf16 remainder, bf16 second remainder).  Operands of the exact-mode panel GEMM / Gram.
__global__ void __launch_bounds__(256)
fixed_planes_kernel(const float* __restrict__ master, int Lp, __nv_bfloat16* __restrict__ cpl, long cpl_plane) {
    __shared__ unsigned int red[8];
    const float* rowp = master + static_cast<long>(blockIdx.x) * Lp;
    unsigned int m = 0;
    for (int i = threadIdx.x * 4; i < Lp; i += 1024) {
        const float4 v = *reinterpret_cast<const float4*>(rowp + i);
        m = max(max(__float_as_uint(fabsf(v.x)), __float_as_uint(fabsf(v.y))),
                max(max(__float_as_uint(fabsf(v.z)), __float_as_uint(fabsf(v.w))), m));
    }
    for (int o = 16; o; o >>= 1) m = max(m, __shfl_xor_sync(0xffffffffu, m, o));
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = m;
    __syncthreads();
    m = 0;
    for (int w = 0; w < 8; ++w) m = max(m, red[w]);
    float u, iu;
    fixed_unit(m, u, iu);
    __nv_bfloat16* out = cpl + static_cast<long>(blockIdx.x) * Lp;
    for (int i = threadIdx.x * 4; i < Lp; i += 1024) {
        const float4 v = *reinterpret_cast<const float4*>(rowp + i);
        const float x[4] = {v.x, v.y, v.z, v.w};
        __align__(8) __nv_bfloat16 p0[4], p1[4], p2[4];
#pragma unroll
        for (int q = 0; q < 4; ++q) {
            const float a0 = fixed_slice(x[q], u, iu);
            float rem = x[q] - a0;
            p0[q] = __float2bfloat16_rn(a0);
            p1[q] = __float2bfloat16_rn(rem);
            rem -= __bfloat162float(p1[q]);
            p2[q] = __float2bfloat16_rn(rem);
        }
        *reinterpret_cast<uint2*>(out + i) = *reinterpret_cast<const uint2*>(p0);
        *reinterpret_cast<uint2*>(out + cpl_plane + i) = *reinterpret_cast<const uint2*>(p1);
        *reinterpret_cast<uint2*>(out + 2 * cpl_plane + i) = *reinterpret_cast<const uint2*>(p2);
    }
}

cudaError_t fixed_planes_launch(const float* master, int KP, int Lp, __nv_bfloat16* cpl, long cpl_plane,
                                cudaStream_t stream) {
    fixed_planes_kernel<<<KP, 256, 0, stream>>>(master, Lp, cpl, cpl_plane);
    count_launch();
    return cudaGetLastError();
}

// out[0] = scale * sum(partial[0..n)) in a fixed order (single block, fp64).
__global__ void __launch_bounds__(256)
sum_partials_kernel(const double* __restrict__ partial, int n, double scale, double* __restrict__ out) {
    __shared__ double red[256];
    double t = 0.0;
    for (int i = threadIdx.x; i < n; i += 256) t += partial[i];
    red[threadIdx.x] = t;
    __syncthreads();
    for (int s = 128; s; s >>= 1) {
        if (threadIdx.x < s) red[threadIdx.x] += red[threadIdx.x + s];
        __syncthreads();
    }
    if (threadIdx.x == 0) out[0] = scale * red[0];
}

cudaError_t sum_partials_launch(const double* partial, int n, double scale, double* out,
                                cudaStream_t stream) {
    sum_partials_kernel<<<1, 256, 0, stream>>>(partial, n, scale, out);
    count_launch();
    return cudaGetLastError();
}

// ------------------------------------------------------------------ KL-ADMM data step
// admm.py:310-313 / ao_admm.py:89-94, element-wise over V:  with P = (W_aux H_aux)[i][j] in prod,
//   v_bar = P - dual_v;  v_aux = ((v_bar - 1) + sqrt((v_bar - 1)^2 + 4 v)) / 2;  dual_v += v_aux - P
// and d_out <- v_aux + dual_v, the data term of the next right-hand sides (admm.py:224; d_out may be prod
// itself).  The root is taken in fp64, in the cancellation-free form 2v / (sqrt(.) - t) when
// t = v_bar - 1 < 0.  A set done_flag (AO-ADMM: the inner loop has broken) leaves everything untouched.
__global__ void __launch_bounds__(256)
kl_vaux_kernel(const float* prod, float* d_out, float* __restrict__ dual_v, long ld, int rows, int cols,
               const __nv_bfloat16* __restrict__ vpl, long v_plane, int v_rows_p, int nv,
               const int* __restrict__ done_flag) {
    if (done_flag != nullptr && *done_flag != 0) return;
    const int r = blockIdx.x;                              // rows on x: gridDim.y stops at 65535
    const int c0 = blockIdx.y * 1024 + threadIdx.x;
    if (r >= rows) return;
    float pv[4], dv[4], vv[4];
#pragma unroll
    for (int u = 0; u < 4; ++u) {
        const int c = c0 + u * 256;
        pv[u] = dv[u] = vv[u] = 0.f;
        if (c < cols) {
            const long o = static_cast<long>(r) * ld + c;
            pv[u] = prod[o];
            dv[u] = dual_v[o];
            const long b = blocked_index(r, c, v_rows_p);
            vv[u] = __bfloat162float(vpl[b]);
            if (nv > 1) vv[u] += __bfloat162float(vpl[v_plane + b]);
        }
    }
#pragma unroll
    for (int u = 0; u < 4; ++u) {
        const int c = c0 + u * 256;
        if (c < cols) {
            const double P = static_cast<double>(pv[u]), v = static_cast<double>(vv[u]);
            const double t = P - static_cast<double>(dv[u]) - 1.0;
            const double root = sqrt(fma(t, t, 4.0 * v));
            const double va = (t >= 0.0) ? 0.5 * (t + root) : (root - t > 0.0 ? 2.0 * v / (root - t) : 0.0);
            const float dnew = static_cast<float>(static_cast<double>(dv[u]) + va - P);
            const long o = static_cast<long>(r) * ld + c;
            dual_v[o] = dnew;
            d_out[o] = static_cast<float>(va + static_cast<double>(dnew));
        }
    }
}

cudaError_t kl_vaux_launch(const float* prod, float* d_out, float* dual_v, long ld, int rows, int cols,
                           const __nv_bfloat16* vpl, long v_plane, int v_rows_p, int nv, const int* done_flag,
                           cudaStream_t stream) {
    if (rows <= 0 || cols <= 0) return cudaSuccess;
    if ((cols + 1023) / 1024 > 65535) return cudaErrorInvalidValue;
    dim3 grid(rows, (cols + 1023) / 1024);
    kl_vaux_kernel<<<grid, 256, 0, stream>>>(prod, d_out, dual_v, ld, rows, cols, vpl, v_plane, v_rows_p, nv, done_flag);
    count_launch();
    return cudaGetLastError();
}

}  // namespace nmfb
