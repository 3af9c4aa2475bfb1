// extern "C" surface of libnmf_b200.so (declared in include/nmf_b200.h): argument checking,
// TMA tensor-map construction and kernel sequencing.  No torch types, no allocation, no syncs.
#include <cstdarg>
#include <cstring>
#include <mutex>

#include "../../include/nmf_b200.h"
#include "admm.cuh"
#include "elementwise.cuh"
#include "kernels.cuh"
#include "kl.cuh"
#include "peer.cuh"

using namespace nmfb;

namespace nmfb {
unsigned long long g_launch_count = 0;
}

namespace {

thread_local char g_err[512] = "";

int fail(int code, const char* fmt, ...) {
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
    return code;
}

#define CUDA_TRY(expr)                                                                       \
    do {                                                                                     \
        cudaError_t e__ = (expr);                                                            \
        if (e__ != cudaSuccess)                                                              \
            return fail(NMFB_ERR_CUDA, "%s: %s (%s:%d)", #expr, cudaGetErrorString(e__), __FILE__, \
                        __LINE__);                                                           \
    } while (0)

using EncodeTiledFn = CUresult (*)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                   const cuuint64_t*, const cuuint32_t*, const cuuint32_t*,
                                   CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion,
                                   CUtensorMapFloatOOBfill);

EncodeTiledFn encode_fn() {
    static EncodeTiledFn fn = nullptr;
    if (!fn) {
        void* p = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
            q == cudaDriverEntryPointSuccess)
            fn = reinterpret_cast<EncodeTiledFn>(p);
    }
    return fn;
}

// bf16 row-major [rows][ld] matrix, box = [box_rows][64 elements], 128-byte swizzle.
int make_map(CUtensorMap* map, const void* base, long rows, long ld, int box_rows) {
    EncodeTiledFn fn = encode_fn();
    if (!fn) return fail(NMFB_ERR_CUDA, "cuTensorMapEncodeTiled not available from the driver");
    cuuint64_t gdim[2] = {static_cast<cuuint64_t>(ld), static_cast<cuuint64_t>(rows)};
    cuuint64_t gstr[1] = {static_cast<cuuint64_t>(ld) * 2};
    cuuint32_t box[2] = {64, static_cast<cuuint32_t>(box_rows)};
    cuuint32_t estr[2] = {1, 1};
    CUresult r = fn(map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(base), gdim, gstr, box, estr,
                    CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                    CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS)
        return fail(NMFB_ERR_CUDA, "cuTensorMapEncodeTiled failed (%d) rows=%ld ld=%ld box_rows=%d base=%p",
                    static_cast<int>(r), rows, ld, box_rows, base);
    return 0;
}

inline __nv_bfloat16* bf(void* p) { return static_cast<__nv_bfloat16*>(p); }
inline const __nv_bfloat16* bf(const void* p) { return static_cast<const __nv_bfloat16*>(p); }
inline cudaStream_t st(void* s) { return static_cast<cudaStream_t>(s); }

int round_up(int x, int a) { return (x + a - 1) / a * a; }

// Profiling hook: record the caller's event `idx` on the stream (no-op when none were supplied).
inline void mark(const nmfb_problem* p, int idx, void* stream) {
    if (p->phase_events && idx < p->n_phase_events && p->phase_events[idx])
        cudaEventRecord(static_cast<cudaEvent_t>(p->phase_events[idx]), st(stream));
}

}  // namespace

extern "C" {

const char* nmfb_last_error(void) { return g_err; }
int nmfb_abi_version(void) { return 1; }
unsigned long long nmfb_launch_count(void) { return g_launch_count; }

int nmfb_padded_components(int k) {
    if (k < 1 || k > 256) return fail(NMFB_ERR_ARG, "components must be in [1, 256], got %d", k);
    return k <= 64 ? 64 : (k <= 128 ? 128 : 256);
}
int nmfb_padded_length(int L) { return round_up(L, 256); }

size_t nmfb_partial_bytes(int m, int n, int k, int num_sms) {
    const int KP = nmfb_padded_components(k);
    if (KP < 0) return 0;
    size_t best = gram_fp64_scratch_bytes(KP, nmfb_padded_length(m > n ? m : n) + 256 * 16);   // fp64 Gram (ADMM)
    const int Ls[3] = {m, n, 256}, Rs[3] = {n, m, (m > n ? m : n)};
    for (int i = 0; i < 3; ++i)
        for (int nb = 2; nb <= 3; ++nb)
            for (int na = 1; na <= 3; ++na) {
                PanelGemmPlan pl;
                if (panel_gemm_plan(KP, nb, na, Ls[i], Rs[i], 0, 0, num_sms, &pl) != 0) continue;
                const size_t b = static_cast<size_t>(pl.slots) * KP * nmfb_padded_length(Ls[i]) * 4;
                if (b > best) best = b;
            }
    for (int side = 0; side < 2; ++side)             // exact-hi ADMM contractions (2 data planes; 3 for the KL data term),
        for (int na = 2; na <= 3; ++na)              // on all SMs or with one CTA pair left to the side stream
            for (int sms = num_sms - 2; sms <= num_sms; sms += 2) {
                PanelGemmPlan xp;
                const int R = side == 0 ? m : n, T = side == 0 ? n : m;
                if (sms < 2 || panel_gemm_exact_plan(KP, na, R, T, 1, sms, &xp) != 0) continue;
                const size_t b = static_cast<size_t>(xp.slots) * KP * nmfb_padded_length(R) * 4;
                if (b > best) best = b;
            }
    {
        PanelGemmPlan xp;                            // exact Gram: data = the factor's own three planes
        if (panel_gemm_exact_plan(KP, 3, KP, m > n ? m : n, 0, num_sms, &xp) == 0) {
            const size_t b = static_cast<size_t>(xp.slots) * KP * 256 * 4;
            if (b > best) best = b;
        }
    }
    if (KP <= 128)
        for (int side = 0; side < 2; ++side) {
            KlPlan kp;
            const int R = side == 0 ? m : n, T = side == 0 ? n : m;
            if (kl_plan(KP, 1, R, T, num_sms, &kp) != 0) continue;
            const size_t b = static_cast<size_t>(kp.slots) * KP * nmfb_padded_length(R) * 4;
            if (b > best) best = b;
        }
    return best;
}

size_t nmfb_gram_partial_bytes(int k, int L, int num_sms) {
    const int KP = nmfb_padded_components(k);
    if (KP < 0) return 0;
    size_t best = 0;
    for (int nb = 2; nb <= 3; ++nb) {
        PanelGemmPlan pl;
        if (panel_gemm_plan(KP, nb, nb, KP, L, 128, 0, num_sms, &pl) != 0) continue;
        const size_t b = static_cast<size_t>(pl.slots) * KP * 256 * 4;
        if (b > best) best = b;
    }
    PanelGemmPlan xp;
    if (panel_gemm_exact_plan(KP, 3, KP, L, 0, num_sms, &xp) == 0) {
        const size_t b = static_cast<size_t>(xp.slots) * KP * 256 * 4;
        if (b > best) best = b;
    }
    const size_t f64 = gram_fp64_scratch_bytes(KP, nmfb_padded_length(L) + 256 * 16);    // + sharding pad
    if (f64 > best) best = f64;
    return best;
}

int nmfb_obj_partial_len(int m, int n, int num_sms) {
    const long simt = static_cast<long>((m + 63) / 64) * ((n + 63) / 64);
    (void)simt;
    return static_cast<int>(simt > num_sms ? simt : num_sms);
}

namespace {
int ingest_common(const void* src, int src_is_f64, long src_ld, int row0, int nrows, const nmfb_data* vr,
                  const nmfb_data* vc, uint32_t* stats, const uint32_t* rowmax, const uint32_t* colmax,
                  void* stream) {
    if (!src || !vr || !vc || !stats) return fail(NMFB_ERR_ARG, "ingest_data: null argument");
    // v_rows holds rows [origin, origin + rows) of V over all n columns, v_cols columns [origin, origin + rows)
    // over all m rows (origin = 0 and rows = m / n when the run is not sharded)
    if (vr->origin + vr->rows > vc->cols || vc->origin + vc->rows > vr->cols || vr->nv != vc->nv)
        return fail(NMFB_ERR_ARG, "ingest_data: v_rows / v_cols shapes disagree");
    if (row0 < 0 || row0 + nrows > vc->cols) return fail(NMFB_ERR_ARG, "ingest_data: row range");
    if (!vr->blocked || !vc->blocked) return fail(NMFB_ERR_ARG, "ingest_data: data planes must be K-blocked");
    CUDA_TRY(ingest_data_launch(src, src_is_f64, src_ld, row0, nrows, vr->cols, bf(vr->planes),
                                static_cast<long>(vr->rows_p) * vr->ld, vr->rows_p, bf(vc->planes),
                                static_cast<long>(vc->rows_p) * vc->ld, vc->rows_p, vr->nv, stats, rowmax, colmax,
                                vr->origin, vr->rows, vc->origin, vc->rows, st(stream)));
    return 0;
}
}  // namespace

int nmfb_ingest_data(const void* src, int src_is_f64, long src_ld, int row0, int nrows, const nmfb_data* vr,
                     const nmfb_data* vc, uint32_t* stats, void* stream) {
    return ingest_common(src, src_is_f64, src_ld, row0, nrows, vr, vc, stats, nullptr, nullptr, stream);
}

int nmfb_ingest_data_fixed(const void* src, int src_is_f64, long src_ld, int row0, int nrows, const nmfb_data* vr,
                           const nmfb_data* vc, uint32_t* stats, const uint32_t* rowmax, const uint32_t* colmax,
                           void* stream) {
    if (!rowmax || !colmax) return fail(NMFB_ERR_ARG, "ingest_data_fixed: row / column maxima missing");
    if (!vr || !vc || vr->nv < 2 || vr->nv > 3 || !vr->fixed || !vc->fixed)
        return fail(NMFB_ERR_ARG, "ingest_data_fixed: needs two or three fixed-point planes");
    return ingest_common(src, src_is_f64, src_ld, row0, nrows, vr, vc, stats, rowmax, colmax, stream);
}

int nmfb_data_absmax(const void* src, int src_is_f64, long src_ld, int row0, int nrows, int ncols,
                     uint32_t* rowmax, uint32_t* colmax, void* stream) {
    if (!src || !rowmax || !colmax) return fail(NMFB_ERR_ARG, "data_absmax: null argument");
    CUDA_TRY(data_absmax_launch(src, src_is_f64, src_ld, row0, nrows, ncols, rowmax, colmax, st(stream)));
    return 0;
}

int nmfb_fixed_planes(int KP, const nmfb_factor* f, void* stream) {
    if (!f) return fail(NMFB_ERR_ARG, "fixed_planes: null factor");
    CUDA_TRY(fixed_planes_launch(f->master, KP, f->Lp, bf(f->cplanes), 256L * f->Lp, st(stream)));
    return 0;
}

int nmfb_refresh_planes(int KP, const nmfb_factor* f, void* stream) {
    if (!f) return fail(NMFB_ERR_ARG, "refresh_planes: null factor");
    CUDA_TRY(refresh_planes_launch(f->master, nullptr, KP, f->Lp, bf(f->cplanes), 256L * f->Lp, bf(f->rplanes),
                                   static_cast<long>(f->Lp) * KP, 0, 0, st(stream)));
    return 0;
}

int nmfb_adopt_exchange(int KP, const nmfb_factor* f, const float* xchg, int refresh, void* stream) {
    if (!f || !xchg) return fail(NMFB_ERR_ARG, "adopt_exchange: null argument");
    if (refresh) {
        // refresh == 2: the rank's own window already holds master and planes of its update kernel (MUR-eu)
        const int lo = refresh == 2 && f->win_p > 0 ? f->win0 : 0;
        const int hi = refresh == 2 && f->win_p > 0 ? f->win0 + f->win_p : 0;
        CUDA_TRY(refresh_planes_launch(xchg, f->master, KP, f->Lp, bf(f->cplanes), 256L * f->Lp, bf(f->rplanes),
                                       static_cast<long>(f->Lp) * KP, lo, hi, st(stream)));
    } else {
        CUDA_TRY(cudaMemcpyAsync(f->master, xchg, static_cast<size_t>(KP) * f->Lp * 4, cudaMemcpyDeviceToDevice,
                                 st(stream)));
    }
    return 0;
}

int nmfb_peer_push(const void* src, long src_pitch, void* const* peers_host, int world, long dst_off,
                   long dst_pitch, int rows, long row_bytes, void* stream) {
    if (!src || !peers_host) return fail(NMFB_ERR_ARG, "peer_push: null argument");
    if (world < 1 || world > kMaxPeers || rows < 1 || (row_bytes & 15) || (dst_off & 15) || (src_pitch & 15) ||
        (dst_pitch & 15))
        return fail(NMFB_ERR_ARG, "peer_push: world in [1, %d], 16-byte multiples (world=%d rows=%d row_bytes=%ld)",
                    kMaxPeers, world, rows, row_bytes);
    CUDA_TRY(peer_push_launch(src, src_pitch, peers_host, world, dst_off, dst_pitch, rows, row_bytes, st(stream)));
    return 0;
}

int nmfb_peer_scatter(const void* src, long src_pitch, long src_peer_stride, void* const* peers_host, int world,
                      long dst_off, long dst_pitch, int rows, const long* row_bytes_host, void* stream) {
    if (!src || !peers_host || !row_bytes_host) return fail(NMFB_ERR_ARG, "peer_scatter: null argument");
    if (world < 1 || world > kMaxPeers || rows < 1) return fail(NMFB_ERR_ARG, "peer_scatter: bad world / rows");
    for (int q = 0; q < world; ++q)
        if (row_bytes_host[q] < 0 || (row_bytes_host[q] & 3))
            return fail(NMFB_ERR_ARG, "peer_scatter: row widths must be multiples of 4 bytes");
    CUDA_TRY(peer_scatter_launch(src, src_pitch, src_peer_stride, peers_host, world, dst_off, dst_pitch, rows,
                                 row_bytes_host, st(stream)));
    return 0;
}

int nmfb_peer_push_mc(const void* src, long src_pitch, void* mc_base, long dst_off, long dst_pitch, int rows,
                      long row_bytes, void* stream) {
    if (!src || !mc_base) return fail(NMFB_ERR_ARG, "peer_push_mc: null argument");
    if (rows < 1 || (row_bytes & 15) || (dst_off & 15) || (src_pitch & 15) || (dst_pitch & 15))
        return fail(NMFB_ERR_ARG, "peer_push_mc: 16-byte multiples (rows=%d row_bytes=%ld)", rows, row_bytes);
    CUDA_TRY(peer_push_mc_launch(src, src_pitch, mc_base, dst_off, dst_pitch, rows, row_bytes, st(stream)));
    return 0;
}

int nmfb_peer_barrier(void* const* flags_host, int rank, int world, unsigned int seq, void* stream) {
    if (!flags_host || world < 1 || world > kMaxPeers || rank < 0 || rank >= world)
        return fail(NMFB_ERR_ARG, "peer_barrier: bad argument");
    CUDA_TRY(peer_barrier_launch(flags_host, rank, world, seq, st(stream)));
    return 0;
}

int nmfb_sum_rank_slots(const double* slots, int world, long stride, int n, double* out64, float* out32,
                        void* stream) {
    if (!slots || world < 1 || n < 1 || (!out64 && !out32)) return fail(NMFB_ERR_ARG, "sum_rank_slots: bad argument");
    CUDA_TRY(sum_rank_slots_launch(slots, world, stride, n, out64, out32, st(stream)));
    return 0;
}

int nmfb_upload_2d(void* dst, long dst_pitch, const void* src_host, long src_pitch, long width_bytes, long height,
                   void* stream) {
    if (!dst || !src_host || width_bytes < 1 || height < 1) return fail(NMFB_ERR_ARG, "upload_2d: bad argument");
    CUDA_TRY(cudaMemcpy2DAsync(dst, static_cast<size_t>(dst_pitch), src_host, static_cast<size_t>(src_pitch),
                               static_cast<size_t>(width_bytes), static_cast<size_t>(height),
                               cudaMemcpyHostToDevice, st(stream)));
    return 0;
}

int nmfb_import_factor(const void* src, int src_is_f64, long stride_l, long stride_c, int k, int KP,
                       const nmfb_factor* f, void* stream) {
    if (!src || !f) return fail(NMFB_ERR_ARG, "import_factor: null argument");
    CUDA_TRY(import_factor_launch(src, src_is_f64, stride_l, stride_c, f->L, k, f->master, KP, f->Lp,
                                  st(stream)));
    return nmfb_refresh_planes(KP, f, stream);
}

int nmfb_export_factor(const nmfb_factor* f, int k, int KP, void* dst, int dst_is_f64, long stride_l,
                       long stride_c, void* stream) {
    if (!dst || !f) return fail(NMFB_ERR_ARG, "export_factor: null argument");
    CUDA_TRY(export_factor_launch(f->master, KP, f->Lp, dst, dst_is_f64, stride_l, stride_c, f->L, k,
                                  st(stream)));
    return 0;
}

int nmfb_panel_gemm(const nmfb_data* data, const void* factor_cplanes, int factor_Lp, int KP, int nb,
                    float* partial, size_t partial_bytes, int out_Lp, int num_sms, int force_bn,
                    int force_split, void* stream) {
    if (!data || !factor_cplanes || !partial) return fail(NMFB_ERR_ARG, "panel_gemm: null argument");
    if (factor_Lp < data->ld) return fail(NMFB_ERR_ARG, "panel_gemm: factor pitch %d < data pitch %d", factor_Lp, data->ld);
    PanelGemmPlan pl;
    if (panel_gemm_plan(KP, nb, data->nv, data->rows, data->cols, force_bn, force_split, num_sms, &pl) != 0)
        return fail(NMFB_ERR_ARG, "panel_gemm: unsupported configuration KP=%d nb=%d nv=%d", KP, nb, data->nv);
    if (pl.npanels * pl.BN > out_Lp || pl.npanels * pl.BN > data->rows_p)
        return fail(NMFB_ERR_ARG, "panel_gemm: padded extents too small (panels=%d BN=%d out_Lp=%d rows_p=%d)",
                    pl.npanels, pl.BN, out_Lp, data->rows_p);
    const size_t need = static_cast<size_t>(pl.slots) * KP * out_Lp * 4;
    if (need > partial_bytes)
        return fail(NMFB_ERR_WORKSPACE, "panel_gemm: partial workspace %zu < %zu bytes", partial_bytes, need);
    PanelGemmArgs a;
    memset(&a, 0, sizeof(a));
    const int a_box_rows = ((KP == 64) ? 64 : 128) / pl.cluster;     // each CTA fetches 1/cluster of a tile
    for (int j = 0; j < nb; ++j) {
        const __nv_bfloat16* base = bf(factor_cplanes) + static_cast<long>(j) * 256 * factor_Lp;
        int rc = make_map(&a.fac[j], base, 256, factor_Lp, a_box_rows);
        if (rc) return rc;
    }
    for (int i = 0; i < data->nv; ++i) {
        const __nv_bfloat16* base = bf(data->planes) + static_cast<long>(i) * data->rows_p * data->ld;
        int rc = data->blocked
                     ? make_map(&a.dat[i], base, static_cast<long>(data->ld / 64) * data->rows_p, 64, pl.dat_box_rows)
                     : make_map(&a.dat[i], base, data->rows_p, data->ld, pl.dat_box_rows);
        if (rc) return rc;
    }
    a.out = partial;
    a.KP = KP; a.NB = nb; a.NA = data->nv; a.BN = pl.BN;
    a.npanels = pl.npanels; a.nsplit = pl.nsplit; a.kb_per_split = pl.kb_per_split;
    a.kblocks_total = pl.kblocks_total;
    a.Lp = out_Lp;
    a.nstages = pl.nstages;
    a.dat_rows_p = data->blocked ? data->rows_p : 0;
    a.cluster = pl.cluster; a.ngroups = pl.ngroups; a.chunk_kb = pl.chunk_kb;
    a.dat_box_rows = pl.dat_box_rows; a.prefetch_kb = pl.prefetch_kb;
    CUDA_TRY(panel_gemm_launch(a, pl, num_sms, st(stream)));
    return pl.slots;
}

int nmfb_panel_gemm_exact(const nmfb_data* data, const void* factor_cplanes, int factor_Lp, int KP,
                          float* partial, size_t partial_bytes, int out_Lp, int num_sms, void* stream) {
    if (!data || !factor_cplanes || !partial) return fail(NMFB_ERR_ARG, "panel_gemm_exact: null argument");
    if (factor_Lp < data->ld) return fail(NMFB_ERR_ARG, "panel_gemm_exact: factor pitch %d < data pitch %d", factor_Lp, data->ld);
    PanelGemmPlan pl;
    if (panel_gemm_exact_plan(KP, data->nv, data->rows, data->cols, data->blocked, num_sms, &pl) != 0)
        return fail(NMFB_ERR_ARG, "panel_gemm_exact: unsupported configuration KP=%d nv=%d", KP, data->nv);
    if (pl.npanels * pl.BN > out_Lp || pl.npanels * pl.BN > data->rows_p)
        return fail(NMFB_ERR_ARG, "panel_gemm_exact: padded extents too small");
    const size_t need = static_cast<size_t>(pl.slots) * KP * out_Lp * 4;
    if (need > partial_bytes)
        return fail(NMFB_ERR_WORKSPACE, "panel_gemm_exact: partial workspace %zu < %zu bytes", partial_bytes, need);
    PanelGemmArgs a;
    memset(&a, 0, sizeof(a));
    const int a_box_rows = ((KP == 64) ? 64 : 128) / pl.cluster;
    for (int j = 0; j < 3; ++j) {
        int rc = make_map(&a.fac[j], bf(factor_cplanes) + static_cast<long>(j) * 256 * factor_Lp, 256, factor_Lp,
                          a_box_rows);
        if (rc) return rc;
    }
    for (int i = 0; i < data->nv; ++i) {
        const __nv_bfloat16* base = bf(data->planes) + static_cast<long>(i) * data->rows_p * data->ld;
        int rc = data->blocked
                     ? make_map(&a.dat[i], base, static_cast<long>(data->ld / 64) * data->rows_p, 64, pl.BN)
                     : make_map(&a.dat[i], base, data->rows_p, data->ld, pl.BN);
        if (rc) return rc;
    }
    a.out = partial;
    a.KP = KP; a.NB = 3; a.NA = data->nv; a.BN = pl.BN;
    a.npanels = pl.npanels; a.nsplit = pl.nsplit; a.kb_per_split = pl.kb_per_split;
    a.kblocks_total = pl.kblocks_total;
    a.Lp = out_Lp;
    a.nstages = pl.nstages;
    a.dat_rows_p = data->blocked ? data->rows_p : 0;
    a.cluster = pl.cluster; a.ngroups = pl.ngroups; a.chunk_kb = pl.chunk_kb;
    a.dat_box_rows = pl.dat_box_rows; a.prefetch_kb = pl.prefetch_kb;
    CUDA_TRY(panel_gemm_exact_launch(a, pl, st(stream)));
    return pl.slots;
}

int nmfb_gram_exact(const nmfb_factor* f, int KP, float* partial, size_t partial_bytes, float* gram32,
                    double* gram64, int num_sms, void* stream) {
    if (KP == 256)      // 3 x 3 planes do not fit two stages: floating 2-plane Gram
        return nmfb_gram(f, KP, 2, partial, partial_bytes, gram32, gram64, num_sms, stream);
    nmfb_data d;
    d.planes = f->cplanes;
    d.rows = KP; d.cols = f->L; d.rows_p = 256; d.ld = f->Lp; d.nv = 3; d.blocked = 0; d.fixed = 1; d.origin = 0;
    const int slots = nmfb_panel_gemm_exact(&d, f->cplanes, f->Lp, KP, partial, partial_bytes, 256, num_sms, stream);
    if (slots < 0) return slots;
    CUDA_TRY(gram_reduce_launch(partial, slots, KP, 256, gram32, gram64, st(stream)));
    return 0;
}

int nmfb_gram(const nmfb_factor* f, int KP, int nplanes, float* partial, size_t partial_bytes, float* gram32,
              double* gram64, int num_sms, void* stream) {
    // The factor's own bf16 planes act as the "data": rows = components, contraction over L.
    if (nplanes != 2 && nplanes != 3) return fail(NMFB_ERR_ARG, "gram: nplanes must be 2 or 3");
    if (KP == 256) nplanes = 2;          // 3 x 3 planes do not fit two smem stages at KP = 256
    nmfb_data d;
    d.planes = f->cplanes;
    d.rows = KP; d.cols = f->L; d.rows_p = 256; d.ld = f->Lp; d.nv = nplanes; d.blocked = 0; d.fixed = 0; d.origin = 0;
    const int slots = nmfb_panel_gemm(&d, f->cplanes, f->Lp, KP, nplanes, partial, partial_bytes, 256, num_sms,
                                      128, 0, stream);
    if (slots < 0) return slots;
    CUDA_TRY(gram_reduce_launch(partial, slots, KP, 256, gram32, gram64, st(stream)));
    return 0;
}

int nmfb_gram_fp64(const nmfb_factor* f, int KP, double* scratch, size_t scratch_bytes, float* gram32, double* gram64,
                   void* stream) {
    if (!f || !scratch || !gram32 || !gram64) return fail(NMFB_ERR_ARG, "gram_fp64: null argument");
    if (KP != 64 && KP != 128 && KP != 256) return fail(NMFB_ERR_ARG, "gram_fp64: KP must be 64 / 128 / 256");
    if (gram_fp64_scratch_bytes(KP, f->Lp) > scratch_bytes)
        return fail(NMFB_ERR_WORKSPACE, "gram_fp64: scratch %zu < %zu bytes", scratch_bytes,
                    gram_fp64_scratch_bytes(KP, f->Lp));
    CUDA_TRY(gram_fp64_launch(bf(f->cplanes), 256L * f->Lp, KP, f->Lp, scratch, gram32, gram64, st(stream)));
    return 0;
}

// ---- Gram matrix without caller scratch ------------------------------------------------------------------------
// Same kernels as nmfb_gram (panel_gemm_kernel into K-split partial slots, gram_reduce_kernel over them), but the
// slots live in a file-scope device array instead of a caller buffer: for callers that cannot allocate (CUDA graph
// capture, serving loops, harnesses that forbid cudaMalloc).  Every GPU that loads the library holds the array until
// the process exits.  No counterpart in the reference.
#define NMFB_GRAM_STATIC_BYTES (8u << 20)
static __device__ __align__(256) float g_gram_static_scratch[NMFB_GRAM_STATIC_BYTES / 4];

namespace {
// `g_gram_static_scratch` is a host-side stand-in in host code, not a device address: resolve the address once per
// device, with that device current, and hand THAT to the kernels.
constexpr int kMaxDevices = 64;
float* g_gram_static_ptr[kMaxDevices] = {};
std::mutex g_gram_static_mu;

int gram_static_scratch(float** out) {
    int dev = -1;
    CUDA_TRY(cudaGetDevice(&dev));
    if (dev < 0 || dev >= kMaxDevices) return fail(NMFB_ERR_ARG, "gram_static: device ordinal %d out of range", dev);
    std::lock_guard<std::mutex> lock(g_gram_static_mu);
    if (!g_gram_static_ptr[dev]) {
        float* ptr = nullptr;
        CUDA_TRY(cudaGetSymbolAddress(reinterpret_cast<void**>(&ptr), g_gram_static_scratch));
        if (!ptr) return fail(NMFB_ERR_CUDA, "gram_static: cudaGetSymbolAddress returned a null address");
        g_gram_static_ptr[dev] = ptr;
    }
    *out = g_gram_static_ptr[dev];
    return 0;
}
}  // namespace

size_t nmfb_gram_static_max_bytes(void) { return NMFB_GRAM_STATIC_BYTES; }

int nmfb_gram_static(const nmfb_factor* f, int KP, int nplanes, float* gram32, double* gram64, int num_sms,
                     void* stream) {
    if (!f || !gram32 || !gram64) return fail(NMFB_ERR_ARG, "gram_static: null argument");
    if (nplanes != 2 && nplanes != 3) return fail(NMFB_ERR_ARG, "gram_static: nplanes must be 2 or 3");
    if (KP == 256) nplanes = 2;                       // as nmfb_gram
    PanelGemmPlan pl;
    if (panel_gemm_plan(KP, nplanes, nplanes, KP, f->L, 128, 0, num_sms, &pl) != 0)
        return fail(NMFB_ERR_ARG, "gram_static: unsupported configuration KP=%d nplanes=%d", KP, nplanes);
    const size_t need = static_cast<size_t>(pl.slots) * KP * 256 * 4;
    if (need > NMFB_GRAM_STATIC_BYTES)
        return fail(NMFB_ERR_ARG,
                    "gram_static: KP=%d L=%d needs %zu bytes of K-split slots, the built-in scratch holds %u; "
                    "use nmfb_gram with a caller buffer of nmfb_gram_partial_bytes()",
                    KP, f->L, need, static_cast<unsigned>(NMFB_GRAM_STATIC_BYTES));
    float* scratch = nullptr;
    const int rc = gram_static_scratch(&scratch);
    if (rc < 0) return rc;
    // nothing to clear: every word gram_reduce_kernel reads below is written by the contraction of this same call
    return nmfb_gram(f, KP, nplanes, scratch, NMFB_GRAM_STATIC_BYTES, gram32, gram64, num_sms, stream);
}

namespace {
// Window of a factor this rank updates: rows [w0, w0 + wp) with `live` true rows inside it.
struct Window { int w0, wp, live; };
Window window_of(const nmfb_factor* f) {
    Window w;
    w.w0 = f->win_p > 0 ? f->win0 : 0;
    w.wp = f->win_p > 0 ? f->win_p : f->Lp;
    int live = f->L - w.w0;
    if (live > w.wp) live = w.wp;
    w.live = live < 0 ? 0 : live;
    return w;
}
}  // namespace

int nmfb_gram_window(const nmfb_factor* f, int KP, int nplanes, float* partial, size_t partial_bytes, float* gram32,
                     double* gram64, int num_sms, void* stream) {
    // This rank's term of the Gram matrix: the contraction runs over the rows of its window only.  The padded
    // rows of a window ([L, Lp)) hold zeros, so whole 64-row contraction blocks may be read.
    if (!f) return fail(NMFB_ERR_ARG, "gram_window: null factor");
    if (nplanes != 2 && nplanes != 3) return fail(NMFB_ERR_ARG, "gram_window: nplanes must be 2 or 3");
    if (KP == 256) nplanes = 2;
    const Window w = window_of(f);
    if (w.live <= 0) return fail(NMFB_ERR_ARG, "gram_window: empty window");
    nmfb_data d;
    d.planes = bf(f->cplanes) + w.w0;
    d.rows = KP; d.cols = w.live; d.rows_p = 256; d.ld = f->Lp; d.nv = nplanes; d.blocked = 0; d.fixed = 0; d.origin = 0;
    const int slots = nmfb_panel_gemm(&d, bf(f->cplanes) + w.w0, f->Lp, KP, nplanes, partial, partial_bytes, 256,
                                      num_sms, 128, 0, stream);
    if (slots < 0) return slots;
    CUDA_TRY(gram_reduce_launch(partial, slots, KP, 256, gram32, gram64, st(stream)));
    return 0;
}

int nmfb_mur_eu_update(const float* partial, int slots, const float* gram_other, const nmfb_factor* f,
                       int k, int KP, float lambda, void* stream) {
    const Window w = window_of(f);
    CUDA_TRY(mur_eu_update_launch(partial + w.w0, slots, gram_other, f->master + w.w0, bf(f->cplanes) + w.w0,
                                  256L * f->Lp, bf(f->rplanes) + static_cast<long>(w.w0) * KP,
                                  static_cast<long>(f->Lp) * KP, KP, k, w.live, f->Lp, w.wp, lambda, st(stream)));
    return 0;
}

namespace {
int objective_common(const nmfb_problem* p, int impl, int kl, double* obj_out, void* stream) {
    if (!p || !obj_out) return fail(NMFB_ERR_ARG, "objective: null argument");
    const nmfb_data* v = &p->v_rows;            // the rows of V this rank holds == the window of W
    const Window ww = window_of(&p->w);
    int nparts = 0;
    // impl 0 = the product's choice: wide tiles (residual_n256.cu) for KP <= 128 with one V plane, the TS form
    // (residual_ts.cu) for everything else it fits, the CUDA-core kernel as the last resort.  impl 1 / 2 / 3 force
    // the CUDA-core / wide / TS kernel (cross-checks in the tests).
    bool wide = impl == 2, ts = impl == 3;
    if (wide && residual_n256_smem_bytes(p->KP, v->nv) < 0)
        return fail(NMFB_ERR_ARG, "objective: the wide-tile kernel needs KP = 64 or 128 and one V plane (KP=%d nv=%d)",
                    p->KP, v->nv);
    if (ts && residual_ts_smem_bytes(p->KP, v->nv) < 0)
        return fail(NMFB_ERR_ARG, "objective: the TS kernel does not fit KP=%d nv=%d", p->KP, v->nv);
    if (impl == 0) {
        wide = residual_n256_smem_bytes(p->KP, v->nv) > 0;
        ts = !wide && residual_ts_smem_bytes(p->KP, v->nv) > 0;
        if (!wide && !ts) impl = 1;
    }
    const int BN = wide ? 256 : 128;
    if (impl == 1) {
        const long need = static_cast<long>((v->rows + 63) / 64) * ((p->n + 63) / 64);
        if (need > p->obj_partial_len) return fail(NMFB_ERR_WORKSPACE, "objective: obj_partial too small");
        CUDA_TRY(objective_simt_launch(bf(v->planes), static_cast<long>(v->rows_p) * v->ld, v->rows_p, v->nv,
                                       p->w.master + ww.w0, p->w.Lp, p->h.master, p->h.Lp, p->k, v->rows, p->n,
                                       p->obj_partial, &nparts, st(stream)));
    } else {
        ResidualArgs a;
        memset(&a, 0, sizeof(a));
        a.mtiles = (v->rows + 127) / 128;
        a.ntiles = (p->n + BN - 1) / BN;
        int cs = 2;                                   // H^T stages are multicast across a CTA pair
        while (cs > 1 && a.mtiles < 2 * cs) cs >>= 1;
        a.cluster = cs;
        for (int q = 0; q < 2; ++q) {
            int rc = make_map(&a.wr[q], bf(p->w.rplanes) + (static_cast<long>(q) * p->w.Lp + ww.w0) * p->KP, ww.wp,
                              p->KP, 128);
            if (rc) return rc;
            a.w_rpl[q] = bf(p->w.rplanes) + (static_cast<long>(q) * p->w.Lp + ww.w0) * p->KP;
            rc = make_map(&a.hr[q], bf(p->h.rplanes) + static_cast<long>(q) * p->h.Lp * p->KP, p->h.Lp, p->KP,
                          BN / cs);
            if (rc) return rc;
        }
        for (int q = 0; q < v->nv; ++q) {
            int rc = make_map(&a.v[q], bf(v->planes) + static_cast<long>(q) * v->rows_p * v->ld,
                              static_cast<long>(v->ld / 64) * v->rows_p, 64, 128);
            if (rc) return rc;
        }
        a.partial = p->obj_partial;
        a.KP = p->KP; a.NV = v->nv; a.BN = BN;
        a.m = p->m; a.n = p->n;
        a.kl = kl;
        a.v_rows_p = v->rows_p;
        const long ntot = static_cast<long>((a.mtiles + cs - 1) / cs) * a.ntiles;
        int ncl = residual_max_clusters(cs, p->num_sms);
        if (ntot < ncl) ncl = static_cast<int>(ntot);
        const int grid = ncl * cs;
        if (grid > p->obj_partial_len) return fail(NMFB_ERR_WORKSPACE, "objective: obj_partial too small");
        nparts = grid;
        if (wide) { CUDA_TRY(residual_n256_launch(a, grid, st(stream))); }
        else      { CUDA_TRY(residual_ts_launch(a, grid, st(stream))); }
    }
    CUDA_TRY(sum_partials_launch(p->obj_partial, nparts, kl ? 1.0 : 0.5, obj_out, st(stream)));
    return 0;
}
}  // namespace

int nmfb_objective_eu(const nmfb_problem* p, int impl, double* obj_out, void* stream) {
    return objective_common(p, impl, 0, obj_out, stream);
}

int nmfb_objective_kl(const nmfb_problem* p, double* obj_out, void* stream) {
    if (p && residual_n256_smem_bytes(p->KP, p->v_rows.nv) < 0 && residual_ts_smem_bytes(p->KP, p->v_rows.nv) < 0)
        return fail(NMFB_ERR_ARG, "objective_kl: unsupported shape");
    return objective_common(p, 0, 1, obj_out, stream);
}

namespace {
int push_to_all(const nmfb_peer_ctx* x, const void* src, long src_pitch, long dst_off, long dst_pitch, int rows,
                long row_bytes, void* stream) {
    if (x->mc_base) return nmfb_peer_push_mc(src, src_pitch, x->mc_base, dst_off, dst_pitch, rows, row_bytes, stream);
    return nmfb_peer_push(src, src_pitch, x->peers, x->world, dst_off, dst_pitch, rows, row_bytes, stream);
}
inline void probe(void* const* ev, int i, void* stream) {
    if (ev && ev[i]) cudaEventRecord(static_cast<cudaEvent_t>(ev[i]), st(stream));
}
}  // namespace

int nmfb_exchange_factor(const nmfb_peer_ctx* x, int KP, const nmfb_factor* f, long xchg_off, int refresh, int fixed,
                         int gram_nplanes, long gram_off, double* gram64, float* gram32, unsigned int seq,
                         void* const* probe_events, void* stream) {
    if (!x || !f || !x->peers) return fail(NMFB_ERR_ARG, "exchange_factor: null argument");
    if (x->world < 2 || x->world > kMaxPeers || x->rank < 0 || x->rank >= x->world)
        return fail(NMFB_ERR_ARG, "exchange_factor: rank %d of %d", x->rank, x->world);
    if (f->win_p <= 0) return fail(NMFB_ERR_ARG, "exchange_factor: the factor has no window");
    const bool gram = gram_nplanes != 0;
    if (gram && (!x->gram_scratch || !x->gram_term64 || !x->gram_term32 || (!gram64 && !gram32)))
        return fail(NMFB_ERR_ARG, "exchange_factor: Gram workspaces missing");
    const bool forked = gram && x->side_stream && x->ev_fork && x->ev_join;
    const long g = static_cast<long>(KP) * KP * 8;
    int rc;
    probe(probe_events, 0, stream);
    if (forked) CUDA_TRY(cudaEventRecord(static_cast<cudaEvent_t>(x->ev_fork), st(stream)));
    // 1. my rows of the master into every rank's exchange buffer
    if ((rc = push_to_all(x, f->master + f->win0, 4L * f->Lp, xchg_off + 4L * f->win0, 4L * f->Lp, KP, 4L * f->win_p,
                          stream)) < 0)
        return rc;
    // 2. my term of the Gram matrix, under the NVLink-bound store above
    if (gram) {
        void* gs = forked ? x->side_stream : stream;
        if (forked) CUDA_TRY(cudaStreamWaitEvent(st(gs), static_cast<cudaEvent_t>(x->ev_fork), 0));
        if ((rc = nmfb_gram_window(f, KP, gram_nplanes, x->gram_scratch, x->gram_scratch_bytes, x->gram_term32,
                                   x->gram_term64, x->num_sms, gs)) < 0)
            return rc;
        if ((rc = push_to_all(x, x->gram_term64, g, gram_off + x->rank * g, g, 1, g, gs)) < 0) return rc;
        if (forked) {
            CUDA_TRY(cudaEventRecord(static_cast<cudaEvent_t>(x->ev_join), st(gs)));
            CUDA_TRY(cudaStreamWaitEvent(st(stream), static_cast<cudaEvent_t>(x->ev_join), 0));
        }
    }
    probe(probe_events, 1, stream);
    // 3. one barrier covers the window and the Gram term
    if ((rc = nmfb_peer_barrier(x->peers, x->rank, x->world, seq, stream)) < 0) return rc;
    probe(probe_events, 2, stream);
    // 4. all windows into master + planes
    const char* local = static_cast<const char*>(x->peers[x->rank]);
    if ((rc = nmfb_adopt_exchange(KP, f, reinterpret_cast<const float*>(local + xchg_off), refresh, stream)) < 0)
        return rc;
    if (fixed && (rc = nmfb_fixed_planes(KP, f, stream)) < 0) return rc;
    probe(probe_events, 3, stream);
    // 5. k x k all-reduce, second half
    if (gram && (rc = nmfb_sum_rank_slots(reinterpret_cast<const double*>(local + gram_off), x->world,
                                          static_cast<long>(KP) * KP, KP * KP, gram64, gram32, stream)) < 0)
        return rc;
    probe(probe_events, 4, stream);
    return 0;
}

int nmfb_exchange_scalar(const nmfb_peer_ctx* x, const double* src, long table_off, unsigned int seq, double* out,
                         void* stream) {
    if (!x || !src || !out || !x->peers) return fail(NMFB_ERR_ARG, "exchange_scalar: null argument");
    int rc;
    if ((rc = push_to_all(x, src, 16, table_off + 64L * x->rank, 64, 1, 16, stream)) < 0) return rc;
    if ((rc = nmfb_peer_barrier(x->peers, x->rank, x->world, seq, stream)) < 0) return rc;
    const char* local = static_cast<const char*>(x->peers[x->rank]);
    return nmfb_sum_rank_slots(reinterpret_cast<const double*>(local + table_off), x->world, 8, 1, out, nullptr,
                               stream);
}

namespace {
int mur_eu_half(const nmfb_problem* p, int side, float lambda, bool with_gram, void* stream) {
    if (!p || (side != 0 && side != 1)) return fail(NMFB_ERR_ARG, "mur_eu_half_step: bad argument");
    const int KP = p->KP;
    const size_t gsz = static_cast<size_t>(KP) * KP;
    // side 0: W <- W * (V H^T) / (W (H H^T) + lambda W + eps)              mur.py:122 -> :29
    // side 1: H <- H * (W^T V) / ((W^T W) H + lambda H + eps)              mur.py:123 -> :45
    const nmfb_factor* self = side == 0 ? &p->w : &p->h;
    const nmfb_factor* other = side == 0 ? &p->h : &p->w;
    const nmfb_data* data = side == 0 ? &p->v_rows : &p->v_cols;
    float* g32 = p->gram32 + (side == 0 ? gsz : 0);
    double* g64 = p->gram64 + (side == 0 ? gsz : 0);
    int rc;
    // with_gram = false: the Gram of the other factor is already in gram32 (sharded runs reduce the ranks' window
    // terms during the exchange of that factor, nmfb_gram_window + nmfb_sum_rank_slots)
    if (with_gram && (rc = nmfb_gram(other, KP, 2, p->partial, p->partial_bytes, g32, g64, p->num_sms, stream)) < 0)
        return rc;
    mark(p, 1 + 3 * side, stream);
    // sharded runs: this rank's rows of the product land at its window of the full-pitch partial
    const int slots = nmfb_panel_gemm(data, other->cplanes, other->Lp, KP, 2, p->partial + window_of(self).w0,
                                      p->partial_bytes, self->Lp, p->num_sms, 0, 0, stream);
    if (slots < 0) return slots;
    mark(p, 2 + 3 * side, stream);
    if ((rc = nmfb_mur_eu_update(p->partial, slots, g32, self, p->k, KP, lambda, stream)) < 0) return rc;
    mark(p, 3 + 3 * side, stream);
    return 0;
}
}  // namespace

int nmfb_mur_eu_half_step(const nmfb_problem* p, int side, float lambda, void* stream) {
    return mur_eu_half(p, side, lambda, true, stream);
}

int nmfb_mur_eu_window_step(const nmfb_problem* p, int side, float lambda, void* stream) {
    return mur_eu_half(p, side, lambda, false, stream);
}

int nmfb_mur_eu_step(const nmfb_problem* p, float lambda_w, float lambda_h, double* obj_out, void* stream) {
    if (!p) return fail(NMFB_ERR_ARG, "mur_eu_step: null problem");
    int rc;
    // phase events: 0 start | 1 gram(H) | 2 V.H^T | 3 W update | 4 gram(W) | 5 W^T.V | 6 H update | 7 objective
    mark(p, 0, stream);
    if ((rc = nmfb_mur_eu_half_step(p, 0, lambda_w, stream)) < 0) return rc;
    if ((rc = nmfb_mur_eu_half_step(p, 1, lambda_h, stream)) < 0) return rc;   // with the NEW W
    // objective of the new pair                                            mur.py:124,127
    if (obj_out && (rc = nmfb_objective_eu(p, 0, obj_out, stream)) < 0) return rc;
    mark(p, 7, stream);
    return 0;
}

namespace {
// One side of the KL update: X (rows of `data`) <- KL rule with the other factor Y.  In a sharded run
// `data` holds this rank's rows only and the window of X is updated.
int kl_side(const nmfb_problem* p, const nmfb_data* data, const nmfb_factor* x, const nmfb_factor* y,
            double* ysum, float lambda, double* obj_start_out, void* stream) {
    const int KP = p->KP;
    const Window w = window_of(x);
    KlPlan pl;
    if (kl_plan(KP, data->nv, data->rows, data->cols, p->num_sms, &pl) != 0)
        return fail(NMFB_ERR_ARG, "mur_kl: unsupported configuration KP=%d nv=%d (k <= 128 in this build)", KP, data->nv);
    const size_t need = static_cast<size_t>(pl.slots) * KP * x->Lp * 4;
    if (need > p->partial_bytes)
        return fail(NMFB_ERR_WORKSPACE, "mur_kl: partial workspace %zu < %zu bytes", p->partial_bytes, need);
    if (pl.npanels * 128 > w.wp || pl.ntiles * 64 > y->Lp || pl.npanels * 128 > data->rows_p)
        return fail(NMFB_ERR_ARG, "mur_kl: padded extents too small");
    KlArgs a;
    memset(&a, 0, sizeof(a));
    int rc;
    for (int q = 0; q < 2; ++q) {
        if ((rc = make_map(&a.xr[q], bf(x->rplanes) + (static_cast<long>(q) * x->Lp + w.w0) * KP, w.wp, KP, 128))) return rc;
        a.x_rpl[q] = bf(x->rplanes) + (static_cast<long>(q) * x->Lp + w.w0) * KP;
        if ((rc = make_map(&a.yc[q], bf(y->cplanes) + static_cast<long>(q) * 256 * y->Lp, 256, y->Lp, 128))) return rc;
    }
    for (int q = 0; q < data->nv; ++q)
        if ((rc = make_map(&a.d[q], bf(data->planes) + static_cast<long>(q) * data->rows_p * data->ld,
                           static_cast<long>(data->ld / 64) * data->rows_p, 64, 128))) return rc;
    a.out = p->partial + w.w0;
    if (obj_start_out) {
        if (pl.grid > p->obj_partial_len) return fail(NMFB_ERR_WORKSPACE, "mur_kl: obj_partial too small");
        a.obj_partial = p->obj_partial;
    }
    a.KP = KP; a.Lp = x->Lp;
    a.npanels = pl.npanels; a.ntiles = pl.ntiles; a.nsplit = pl.nsplit; a.tiles_per_split = pl.tiles_per_split;
    a.d_rows_p = data->rows_p;
    CUDA_TRY(factor_rowsum_launch(y->master, KP, y->L, y->Lp, ysum, st(stream)));
    CUDA_TRY(kl_panel_launch(a, pl, data->nv, p->num_sms, st(stream)));
    CUDA_TRY(mur_kl_update_launch(p->partial + w.w0, pl.slots, ysum, x->master + w.w0, bf(x->cplanes) + w.w0,
                                  256L * x->Lp, bf(x->rplanes) + static_cast<long>(w.w0) * KP,
                                  static_cast<long>(x->Lp) * KP, KP, p->k, w.live, x->Lp, w.wp, lambda, st(stream)));
    if (obj_start_out) CUDA_TRY(sum_partials_launch(p->obj_partial, pl.grid, 1.0, obj_start_out, st(stream)));
    return 0;
}
}  // namespace

int nmfb_mur_kl_half_step(const nmfb_problem* p, int side, float lambda, double* obj_start_out, void* stream) {
    if (!p || (side != 0 && side != 1)) return fail(NMFB_ERR_ARG, "mur_kl_half_step: bad argument");
    if (!p->ksum) return fail(NMFB_ERR_ARG, "mur_kl_half_step: problem.ksum workspace missing");
    // phase events: 0 | W update (+ objective of the incoming pair) | 1, 2 | H update | 3
    mark(p, 2 * side, stream);
    const int rc = side == 0 ? kl_side(p, &p->v_rows, &p->w, &p->h, p->ksum, lambda, obj_start_out, stream)             // mur.py:122
                             : kl_side(p, &p->v_cols, &p->h, &p->w, p->ksum + p->KP, lambda, obj_start_out, stream);    // mur.py:123
    mark(p, 2 * side + 1, stream);
    return rc;
}

int nmfb_mur_kl_step_lagged(const nmfb_problem* p, float lambda_w, float lambda_h, double* obj_start_out,
                            void* stream) {
    if (!p) return fail(NMFB_ERR_ARG, "mur_kl_step_lagged: null problem");
    int rc;
    if ((rc = nmfb_mur_kl_half_step(p, 0, lambda_w, obj_start_out, stream)) < 0) return rc;
    if ((rc = nmfb_mur_kl_half_step(p, 1, lambda_h, nullptr, stream)) < 0) return rc;
    return 0;
}

int nmfb_mur_kl_step(const nmfb_problem* p, float lambda_w, float lambda_h, double* obj_out, void* stream) {
    if (!p) return fail(NMFB_ERR_ARG, "mur_kl_step: null problem");
    if (!p->ksum) return fail(NMFB_ERR_ARG, "mur_kl_step: problem.ksum workspace missing");
    int rc;
    // phase events: 0 | W update | 1, 2 | H update | 3 | objective | 4
    if ((rc = nmfb_mur_kl_half_step(p, 0, lambda_w, nullptr, stream)) < 0) return rc;
    if ((rc = nmfb_mur_kl_half_step(p, 1, lambda_h, nullptr, stream)) < 0) return rc;
    if (obj_out && (rc = nmfb_objective_kl(p, obj_out, stream)) < 0) return rc;                       // mur.py:127
    mark(p, 4, stream);
    return 0;
}

int nmfb_chol_inverse(const double* gram64, int k, int KP, int rho_mode, double rho_in, double* scratch,
                      double* ainv, double* rho_out, int* flags, void* stream) {
    if (!gram64 || !ainv || !rho_out || !flags) return fail(NMFB_ERR_ARG, "chol_inverse: null argument");
    if (KP == 256 && !scratch) return fail(NMFB_ERR_ARG, "chol_inverse: KP = 256 needs the scratch matrix");
    CUDA_TRY(chol_inverse_launch(gram64, KP, k, rho_mode, rho_in, scratch, ainv, rho_out, flags, st(stream)));
    return 0;
}

namespace {

// aux / prox / dual of one factor (x = `f`, dual `u`) from the numerator in p->partial (or s->c64).
int admm_apply(const nmfb_problem* p, const nmfb_admm_state* s, const float* part, int slots, bool use_c64,
               int side /*0 = H, 1 = W*/, const nmfb_factor* f, float* u, const nmfb_factor* aux, float lambda,
               int prox, bool want_cpl, bool norms, void* stream) {
    AdmmApplyArgs a;
    memset(&a, 0, sizeof(a));
    const Window w = window_of(f);                        // sharded runs: this rank's rows of the factor only
    a.part = part ? part + w.w0 : nullptr; a.slots = slots;
    a.c64 = use_c64 ? s->c64 + w.w0 : nullptr;
    a.ainv = s->ainv + static_cast<size_t>(side) * p->KP * p->KP;
    a.rho = s->scal + side;
    a.lambda = lambda; a.prox = prox;
    if (prox == NMFB_PROX_L2N && lambda != 0.f) {
        if (!s->binv) return fail(NMFB_ERR_ARG, "prox l2n: admm_state.binv workspace missing");
        if (p->KP > 128) return fail(NMFB_ERR_ARG, "prox l2n with lambda > 0 supports k <= 128 in this build");
        a.binv = s->binv + static_cast<size_t>(side) * p->KP * p->KP;
    }
    a.x = f->master + w.w0; a.u = u + w.w0;
    if (aux) {
        a.aux_master = aux->master + w.w0;
        if (!s->exact) { a.aux_cpl = bf(aux->cplanes) + w.w0; a.aux_cpl_plane = 256L * aux->Lp; }
    }
    if (want_cpl && !s->exact) { a.x_cpl = bf(f->cplanes) + w.w0; a.x_cpl_plane = 256L * f->Lp; }
    a.x_rpl = bf(f->rplanes) + static_cast<long>(w.w0) * p->KP; a.x_rpl_plane = static_cast<long>(f->Lp) * p->KP;
    if (norms) { a.norm_partial = s->norm_partial + static_cast<long>(w.w0 / 64) * 4; a.done_flag = s->flags + 1; }
    a.KP = p->KP; a.k = p->k; a.L = w.live; a.Lp = f->Lp; a.nrows_p = w.wp;
    CUDA_TRY(admm_apply_launch(a, st(stream)));
    return 0;
}

// Gram and numerator contraction of one ADMM side, in the floating 3-plane or the exact-hi arithmetic.
// The k x k work of a half step may run on the caller's side stream next to the numerator GEMM (nmfb_admm_state).
struct KxK {
    void* stream;            // where Gram / Cholesky / Tikhonov go
    float* partial;          // scratch of the Gram contraction
    size_t partial_bytes;
    bool forked;
};
KxK kxk_begin(const nmfb_problem* p, const nmfb_admm_state* s, void* stream) {
    KxK k{stream, p->partial, p->partial_bytes, false};
    if (s->side_stream && s->ev_fork && s->ev_join && s->gram_partial) {
        if (cudaEventRecord(static_cast<cudaEvent_t>(s->ev_fork), st(stream)) == cudaSuccess &&
            cudaStreamWaitEvent(st(s->side_stream), static_cast<cudaEvent_t>(s->ev_fork), 0) == cudaSuccess) {
            k.stream = s->side_stream; k.partial = s->gram_partial; k.partial_bytes = s->gram_partial_bytes;
            k.forked = true;
        }
    }
    return k;
}
// The fp64 Gram fills the GPU (tiles x row splits) and runs on the caller's stream BEFORE the fork: next to the
// persistent numerator GEMM it only got the two SMs that GEMM leaves free (measured: the "hidden" Gram stretched the
// GEMM phase from 1.10 to 1.43 - 1.65 ms).  What the side stream hides is the single-CTA work: Cholesky inverse,
// Tikhonov system.
KxK kxk_main(const nmfb_problem* p, const nmfb_admm_state* s, void* stream) {
    if (s->gram_partial) return KxK{stream, s->gram_partial, s->gram_partial_bytes, false};
    return KxK{stream, p->partial, p->partial_bytes, false};
}
int kxk_join(const nmfb_admm_state* s, const KxK& k, void* stream) {
    if (!k.forked) return 0;
    CUDA_TRY(cudaEventRecord(static_cast<cudaEvent_t>(s->ev_join), st(k.stream)));
    CUDA_TRY(cudaStreamWaitEvent(st(stream), static_cast<cudaEvent_t>(s->ev_join), 0));
    return 0;
}

int admm_gram(const nmfb_problem* p, const nmfb_admm_state* s, const nmfb_factor* f, int side, const KxK& k) {
    // fp64 Gram on the FP64 pipe of the SAME three planes the numerator GEMM contracts (admm.cu: the tensor-core Gram
    // is good to 4e-9, which the inverse of G + rho I turns into a systematic 7e-8; a Gram of the master would not
    // match the numerator's operand)
    (void)s;
    const size_t gsz = static_cast<size_t>(p->KP) * p->KP;
    if (gram_fp64_scratch_bytes(p->KP, f->Lp) > k.partial_bytes)
        return fail(NMFB_ERR_WORKSPACE, "admm: Gram scratch %zu < %zu bytes", k.partial_bytes,
                    gram_fp64_scratch_bytes(p->KP, f->Lp));
    CUDA_TRY(gram_fp64_launch(bf(f->cplanes), 256L * f->Lp, p->KP, f->Lp, reinterpret_cast<double*>(k.partial),
                              p->gram32 + side * gsz, p->gram64 + side * gsz, st(k.stream)));
    return 0;
}
int admm_gemm(const nmfb_problem* p, const nmfb_admm_state* s, const nmfb_data* data, const nmfb_factor* other,
              const nmfb_factor* self, bool leave_room, void* stream) {
    // sharded runs: this rank's rows of the product land at its window of the full-pitch partial
    float* out = p->partial + window_of(self).w0;
    // leave_room: the k x k kernels run next to this GEMM on the side stream; the single-CTA Cholesky needs an SM
    // of its own (it cannot share one with a GEMM CTA: 132 KB + ~200 KB of shared memory), so the persistent grid
    // leaves one CTA pair out
    const int sms = leave_room ? p->num_sms - 2 : p->num_sms;
    if (s->exact) {
        if (!data->fixed) return fail(NMFB_ERR_ARG, "admm: exact mode needs fixed-point data planes");
        return nmfb_panel_gemm_exact(data, other->cplanes, other->Lp, p->KP, out, p->partial_bytes, self->Lp,
                                     sms, stream);
    }
    return nmfb_panel_gemm(data, other->cplanes, other->Lp, p->KP, 3, out, p->partial_bytes, self->Lp,
                           p->num_sms, 0, 0, stream);
}

int check_prox(int prox) {
    if (prox != NMFB_PROX_NN && prox != NMFB_PROX_L1N && prox != NMFB_PROX_L2N)
        return fail(NMFB_ERR_ARG, "unknown prox code %d", prox);
    return 0;
}

}  // namespace

int nmfb_admm_prepare_l2n(const nmfb_problem* p, const nmfb_admm_state* s, int side, double lambda, void* stream) {
    if (!p || !s || (side != 0 && side != 1)) return fail(NMFB_ERR_ARG, "admm_prepare_l2n: bad argument");
    if (!s->binv || !s->tik_scratch) return fail(NMFB_ERR_ARG, "admm_prepare_l2n: binv / tik_scratch workspace missing");
    if (p->KP > 128) return fail(NMFB_ERR_ARG, "prox l2n with lambda > 0 supports k <= 128 in this build");
    const size_t gsz = static_cast<size_t>(p->KP) * p->KP;
    // a = 1/rho (lambda T^T T + rho I) = (lambda/rho) T^T T + I                            admm.py:148
    CUDA_TRY(tikhonov_gram_launch(p->k, p->KP, lambda, s->scal + side, s->tik_scratch, st(stream)));
    return nmfb_chol_inverse(s->tik_scratch, p->k, p->KP, 0, 1.0, s->chol_scratch, s->binv + side * gsz,
                             s->tik_scratch + gsz, s->flags, stream);
}

int nmfb_admm_half_step(const nmfb_problem* p, const nmfb_admm_state* s, int side, double rho, float lambda,
                        int prox, void* stream) {
    if (!p || !s || (side != 0 && side != 1)) return fail(NMFB_ERR_ARG, "admm_half_step: bad argument");
    int rc;
    if ((rc = check_prox(prox)) < 0) return rc;
    const int KP = p->KP;
    const size_t gsz = static_cast<size_t>(KP) * KP;
    // side 0: h_aux = (W_aux^T W_aux + rho I)^-1 (W_aux^T V + rho (H + U_h))               admm.py:295
    //         h = prox(h_aux, U_h), U_h += h - h_aux                                        admm.py:299,320
    // side 1: the same for W^T with the NEW h_aux                                           admm.py:296-297,300,321
    const nmfb_factor* other_aux = side == 0 ? &s->w_aux : &s->h_aux;
    const nmfb_factor* self = side == 0 ? &p->h : &p->w;
    const nmfb_factor* self_aux = side == 0 ? &s->h_aux : &s->w_aux;
    const nmfb_data* data = side == 0 ? &p->v_cols : &p->v_rows;
    float* dual = side == 0 ? s->dual_h : s->dual_w;
    if ((rc = admm_gram(p, s, other_aux, side, kxk_main(p, s, stream))) < 0) return rc;
    const KxK kx = kxk_begin(p, s, stream);
    if ((rc = nmfb_chol_inverse(p->gram64 + side * gsz, p->k, KP, 0, rho, s->chol_scratch, s->ainv + side * gsz,
                                s->scal + side, s->flags, kx.stream)) < 0) return rc;
    if (prox == NMFB_PROX_L2N && lambda != 0.f && (rc = nmfb_admm_prepare_l2n(p, s, side, lambda, kx.stream)) < 0) return rc;
    mark(p, 1 + 3 * side, stream);
    const int slots = admm_gemm(p, s, data, other_aux, self, kx.forked, stream);
    if (slots < 0) return slots;
    if ((rc = kxk_join(s, kx, stream)) < 0) return rc;
    mark(p, 2 + 3 * side, stream);
    return admm_apply(p, s, p->partial, slots, false, side, self, dual, self_aux, lambda, prox, false, false, stream);
}

int nmfb_admm_publish(const nmfb_admm_state* s, double* mailbox, void* stream) {
    if (!s || !mailbox) return fail(NMFB_ERR_ARG, "admm_publish: null argument");
    CUDA_TRY(admm_flags_publish_launch(s->flags, mailbox, st(stream)));
    return 0;
}

int nmfb_kl_vaux(const nmfb_data* v_rows, const float* prod, float* d_out, long ld, float* dual_v,
                 const int* done_flag, void* stream) {
    if (!v_rows || !v_rows->planes || !prod || !d_out || !dual_v) return fail(NMFB_ERR_ARG, "kl_vaux: null argument");
    if (!v_rows->blocked || v_rows->nv < 1 || v_rows->nv > 2 || ld < v_rows->cols)
        return fail(NMFB_ERR_ARG, "kl_vaux: needs K-blocked data planes (nv = 1 or 2) and ld >= cols");
    CUDA_TRY(kl_vaux_launch(prod, d_out, dual_v, ld, v_rows->rows, v_rows->cols, bf(v_rows->planes),
                            static_cast<long>(v_rows->rows_p) * v_rows->ld, v_rows->rows_p, v_rows->nv, done_flag,
                            st(stream)));
    return 0;
}

int nmfb_row_planes3(int KP, const nmfb_factor* f, void* planes3, void* stream) {
    if (!f || !planes3) return fail(NMFB_ERR_ARG, "row_planes3: null argument");
    CUDA_TRY(row_planes3_launch(f->master, KP, f->Lp, bf(planes3), st(stream)));
    return 0;
}

int nmfb_kl_data_step(const nmfb_data* v_rows, const void* a_planes3, int a_Lp, int a_row0, const void* b_planes3,
                      int b_Lp, int KP, float* dual_v, float* d_out, long ld, uint32_t* rowmax, uint32_t* colmax,
                      const int* done_flag, int num_sms, void* stream) {
    if (!v_rows || !v_rows->planes || !a_planes3 || !b_planes3 || !dual_v || !d_out || !rowmax || !colmax)
        return fail(NMFB_ERR_ARG, "kl_data_step: null argument");
    if (!v_rows->blocked || v_rows->nv < 1 || v_rows->nv > 2)
        return fail(NMFB_ERR_ARG, "kl_data_step: needs K-blocked data planes (nv = 1 or 2)");
    if (kl_data_smem_bytes(KP, v_rows->nv) < 0)
        return fail(NMFB_ERR_ARG, "kl_data_step: k <= 128 in this build (KP=%d)", KP);
    KlDataArgs a;
    memset(&a, 0, sizeof(a));
    a.mtiles = (v_rows->rows + 127) / 128;
    a.ntiles = (v_rows->cols + 127) / 128;
    if (a.mtiles < 1 || a.ntiles < 1) return fail(NMFB_ERR_ARG, "kl_data_step: empty shard");
    if (ld < static_cast<long>(a.ntiles) * 128 || (ld & 63) || a.mtiles * 128 > v_rows->rows_p ||
        a.ntiles * 128 > b_Lp || a_row0 + a.mtiles * 128 > a_Lp)
        return fail(NMFB_ERR_ARG, "kl_data_step: padded extents too small (ld=%ld rows_p=%d)", ld, v_rows->rows_p);
    int cs = 2;
    while (cs > 1 && a.mtiles < 2 * cs) cs >>= 1;
    a.cluster = cs;
    int rc;
    for (int q = 0; q < 3; ++q) {
        a.a_rpl[q] = bf(a_planes3) + (static_cast<long>(q) * a_Lp + a_row0) * KP;
        if ((rc = make_map(&a.br[q], bf(b_planes3) + static_cast<long>(q) * b_Lp * KP, b_Lp, KP, 128 / cs))) return rc;
    }
    for (int q = 0; q < v_rows->nv; ++q)
        if ((rc = make_map(&a.v[q], bf(v_rows->planes) + static_cast<long>(q) * v_rows->rows_p * v_rows->ld,
                           static_cast<long>(v_rows->ld / 64) * v_rows->rows_p, 64, 128))) return rc;
    a.dual_v = dual_v; a.d_out = d_out; a.rows_p = v_rows->rows_p;
    a.rowmax = rowmax; a.colmax = colmax; a.done_flag = done_flag; a.row_origin = v_rows->origin;
    a.KP = KP; a.NV = v_rows->nv; a.v_rows_p = v_rows->rows_p;
    const long ntot = static_cast<long>((a.mtiles + cs - 1) / cs) * a.ntiles;
    int ncl = residual_max_clusters(cs, num_sms);
    if (ntot < ncl) ncl = static_cast<int>(ntot);
    CUDA_TRY(zero_words_guarded_launch(rowmax + v_rows->origin, v_rows->rows, colmax, v_rows->cols, done_flag,
                                       st(stream)));
    CUDA_TRY(kl_data_launch(a, ncl * cs, st(stream)));
    return 0;
}

int nmfb_admm_step(const nmfb_problem* p, const nmfb_admm_state* s, double rho, float lambda_w, int prox_w,
                   float lambda_h, int prox_h, double* mailbox, void* stream) {
    if (!p || !s || !mailbox) return fail(NMFB_ERR_ARG, "admm_step: null argument");
    int rc;
    const int KP = p->KP;
    mark(p, 0, stream);
    if ((rc = nmfb_admm_half_step(p, s, 0, rho, lambda_h, prox_h, stream)) < 0) return rc;
    if (s->exact && (rc = nmfb_fixed_planes(KP, &s->h_aux, stream)) < 0) return rc;
    mark(p, 3, stream);
    if ((rc = nmfb_admm_half_step(p, s, 1, rho, lambda_w, prox_w, stream)) < 0) return rc;
    if (s->exact && (rc = nmfb_fixed_planes(KP, &s->w_aux, stream)) < 0) return rc;
    mark(p, 6, stream);
    if ((rc = nmfb_objective_eu(p, 0, mailbox, stream)) < 0) return rc;               // admm.py:324
    if ((rc = nmfb_admm_publish(s, mailbox, stream)) < 0) return rc;
    mark(p, 7, stream);
    return 0;
}

namespace {
void ao_roles(const nmfb_problem* p, const nmfb_admm_state* s, int side, const nmfb_factor** other,
              const nmfb_factor** self, const nmfb_data** data, float** dual) {
    // side 0: H update with a = W, y = V (ao_admm.py:261-264); side 1: W update with a = H^T, y = V^T (:265-270)
    *other = side == 0 ? &p->w : &p->h;
    *self = side == 0 ? &p->h : &p->w;
    *data = side == 0 ? &p->v_cols : &p->v_rows;
    *dual = side == 0 ? s->dual_h : s->dual_w;
}
}  // namespace

int nmfb_aoadmm_side_begin(const nmfb_problem* p, const nmfb_admm_state* s, int side, void* stream) {
    if (!p || !s || (side != 0 && side != 1)) return fail(NMFB_ERR_ARG, "aoadmm_side_begin: bad argument");
    const nmfb_factor *other, *self; const nmfb_data* data; float* dual;
    ao_roles(p, s, side, &other, &self, &data, &dual);
    const int KP = p->KP;
    const size_t gsz = static_cast<size_t>(KP) * KP;
    int rc;
    CUDA_TRY(admm_flags_reset_launch(s->flags, (1 << 1) | (1 << (2 + side)), st(stream)));
    if ((rc = admm_gram(p, s, other, side, kxk_main(p, s, stream))) < 0) return rc;   // ao_admm.py:53-55
    const KxK kx = kxk_begin(p, s, stream);
    if ((rc = nmfb_chol_inverse(p->gram64 + side * gsz, p->k, KP, 1, 0.0, s->chol_scratch, s->ainv + side * gsz,
                                s->scal + side, s->flags, kx.stream)) < 0) return rc;
    mark(p, 1 + 3 * side, stream);
    const int slots = admm_gemm(p, s, data, other, self, kx.forked, stream);          // ao_admm.py:56
    if (slots < 0) return slots;
    if ((rc = kxk_join(s, kx, stream)) < 0) return rc;
    mark(p, 2 + 3 * side, stream);
    return slots;
}

int nmfb_aoadmm_inner(const nmfb_problem* p, const nmfb_admm_state* s, int side, int first, int slots,
                      float lambda, int prox, void* stream) {
    if (!p || !s || (side != 0 && side != 1)) return fail(NMFB_ERR_ARG, "aoadmm_inner: bad argument");
    int rc;
    if ((rc = check_prox(prox)) < 0) return rc;
    const nmfb_factor *other, *self; const nmfb_data* data; float* dual;
    ao_roles(p, s, side, &other, &self, &data, &dual);
    // the auxiliary itself is kept only when the state carries buffers for it (KL branch: a.aux, ao_admm.py:88)
    const nmfb_factor* aux = side == 0 ? &s->h_aux : &s->w_aux;
    // The k x k apply stays in fp64 for every inner iteration.  Splitting it (base = Ainv . C in fp64 once per factor
    // update, the iteration's own term rho * Ainv . (x + u) in fp32) was tried in round 2: 1.6x faster inner loops, but
    // the transient of the k = 64 trajectory check moved from 9.3e-5 to 1.5e-4 of the objective -- Ainv has
    // cancelling entries (cond ~ k + 1) and fp32 accumulation of the product is amplified by that factor.
    return admm_apply(p, s, first ? p->partial : nullptr, slots, true, side, self, dual,
                      aux->master ? aux : nullptr, lambda, prox, true, true, stream);  // ao_admm.py:59-62
}

int nmfb_aoadmm_inner_flag(const nmfb_problem* p, const nmfb_admm_state* s, int side, void* stream) {
    if (!p || !s || (side != 0 && side != 1)) return fail(NMFB_ERR_ARG, "aoadmm_inner_flag: bad argument");
    const nmfb_factor* self = side == 0 ? &p->h : &p->w;
    CUDA_TRY(admm_inner_flag_launch(s->norm_partial, self->Lp / 64, s->flags, 2 + side, st(stream)));   // ao_admm.py:64
    return 0;
}

int nmfb_aoadmm_step(const nmfb_problem* p, const nmfb_admm_state* s, int admm_iter, float lambda_w,
                     int prox_w, float lambda_h, int prox_h, double* mailbox, void* stream) {
    if (!p || !s || !mailbox) return fail(NMFB_ERR_ARG, "aoadmm_step: null argument");
    if (admm_iter < 0) return fail(NMFB_ERR_ARG, "aoadmm_step: admm_iter < 0");
    int rc;
    if ((rc = check_prox(prox_w)) < 0 || (rc = check_prox(prox_h)) < 0) return rc;
    const int KP = p->KP;
    mark(p, 0, stream);
    for (int side = 0; side < 2; ++side) {
        const nmfb_factor* self = side == 0 ? &p->h : &p->w;
        const int slots = nmfb_aoadmm_side_begin(p, s, side, stream);
        if (slots < 0) return slots;
        {
            const float lam = side == 0 ? lambda_h : lambda_w;                      // rho = tr(G)/k is known now
            if ((side == 0 ? prox_h : prox_w) == NMFB_PROX_L2N && lam != 0.f &&
                (rc = nmfb_admm_prepare_l2n(p, s, side, lam, stream)) < 0) return rc;
        }
        for (int it = 0; it < admm_iter; ++it) {
            if ((rc = nmfb_aoadmm_inner(p, s, side, it == 0, slots, side == 0 ? lambda_h : lambda_w,
                                        side == 0 ? prox_h : prox_w, stream)) < 0) return rc;
            if ((rc = nmfb_aoadmm_inner_flag(p, s, side, stream)) < 0) return rc;
        }
        if (s->exact && (rc = nmfb_fixed_planes(KP, self, stream)) < 0) return rc;
        mark(p, 3 + 3 * side, stream);
    }
    if ((rc = nmfb_objective_eu(p, 0, mailbox, stream)) < 0) return rc;               // ao_admm.py:291
    if ((rc = nmfb_admm_publish(s, mailbox, stream)) < 0) return rc;
    mark(p, 7, stream);
    return 0;
}

}  // extern "C"
