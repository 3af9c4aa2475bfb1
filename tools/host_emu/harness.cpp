// C entry points (ctypes) that launch the extracted kernels on host threads with the grids / shared-memory
// sizes their real launchers use (admm.cu admm_apply_launch, elementwise.cu *_launch).
#include "gen_kernels.h"

using namespace nmfb;
using bf16 = __nv_bfloat16;

extern "C" {

// admm_apply_launch: grid = window / 64, 256 threads, (KP*64 + 32*KP) doubles of dynamic shared memory
void emu_admm_apply(int l2n, int KP, int k, int L, int Lp, int nrows_p, const float* part, int slots, double* c64,
                    const double* ainv, const double* binv, const double* rho, float lambda, int prox, float* x,
                    float* u, float* aux_master, std::uint16_t* aux_cpl, std::uint16_t* x_cpl,
                    std::uint16_t* x_rpl, double* norm_partial, const int* done_flag) {
    AdmmApplyArgs a;
    std::memset(&a, 0, sizeof(a));
    a.part = part; a.slots = slots; a.c64 = c64; a.ainv = ainv; a.rho = rho; a.lambda = lambda; a.prox = prox;
    a.x = x; a.u = u; a.aux_master = aux_master;
    a.aux_cpl = reinterpret_cast<bf16*>(aux_cpl); a.aux_cpl_plane = 256L * Lp;
    a.x_cpl = reinterpret_cast<bf16*>(x_cpl); a.x_cpl_plane = 256L * Lp;
    a.x_rpl = reinterpret_cast<bf16*>(x_rpl); a.x_rpl_plane = static_cast<long>(Lp) * KP;
    a.norm_partial = norm_partial; a.done_flag = done_flag;
    a.KP = KP; a.k = k; a.L = L; a.Lp = Lp; a.nrows_p = nrows_p; a.binv = l2n ? binv : nullptr;
    const unsigned grid = static_cast<unsigned>((nrows_p > 0 ? nrows_p : Lp) / 64);
    const size_t smem = (static_cast<size_t>(KP) * 64 + 32 * static_cast<size_t>(KP)) * sizeof(double);
    std::function<void()> body;
    if (l2n) {
        if (KP == 64) body = [&] { admm_apply_l2n_kernel<4>(a); };
        else body = [&] { admm_apply_l2n_kernel<8>(a); };
    } else {
        if (KP == 64) body = [&] { admm_apply_kernel<4>(a); };
        else if (KP == 128) body = [&] { admm_apply_kernel<8>(a); };
        else body = [&] { admm_apply_kernel<16>(a); };
    }
    emu::launch({grid, 1, 1}, {256, 1, 1}, smem, body);
}

// chol_inverse_launch: one block of 1024 threads; the working matrix in shared memory up to KP = 128
void emu_chol_inverse(int KP, const double* gram, int k, int rho_mode, double rho_in, double* gscratch, double* ainv,
                      double* rho_out, int* flags) {
    std::function<void()> body;
    size_t smem = 0;
    if (KP == 64) { smem = 64 * 65 * sizeof(double); body = [&] { chol_inverse_kernel<64>(gram, k, rho_mode, rho_in, gscratch, ainv, rho_out, flags); }; }
    else if (KP == 128) { smem = 128 * 129 * sizeof(double); body = [&] { chol_inverse_kernel<128>(gram, k, rho_mode, rho_in, gscratch, ainv, rho_out, flags); }; }
    else body = [&] { chol_inverse_kernel<256>(gram, k, rho_mode, rho_in, gscratch, ainv, rho_out, flags); };
    emu::launch({1, 1, 1}, {1024, 1, 1}, smem, body);
}

void emu_tikhonov_gram(int k, int KP, double lambda, const double* rho, double* out) {
    emu::launch({static_cast<unsigned>((KP * KP + 255) / 256), 1, 1}, {256, 1, 1}, 0,
                [&] { tikhonov_gram_kernel(k, KP, lambda, rho, out); });
}

void emu_inner_flag(const double* norm_partial, int nblocks, int* flags, int used_idx) {
    emu::launch({1, 1, 1}, {256, 1, 1}, 0, [&] { admm_inner_flag_kernel(norm_partial, nblocks, flags, used_idx); });
}

void emu_kl_vaux(const float* prod, float* d_out, float* dual_v, long ld, int rows, int cols,
                 const std::uint16_t* vpl, long v_plane, int v_rows_p, int nv, const int* done_flag) {
    emu::launch({static_cast<unsigned>(rows), static_cast<unsigned>((cols + 1023) / 1024), 1}, {256, 1, 1}, 0, [&] {
        kl_vaux_kernel(prod, d_out, dual_v, ld, rows, cols, reinterpret_cast<const bf16*>(vpl), v_plane, v_rows_p,
                       nv, done_flag);
    });
}

void emu_data_absmax(const float* src, long ld, int row0, int nrows, int ncols, unsigned* rowmax, unsigned* colmax) {
    emu::launch({static_cast<unsigned>((ncols + 63) / 64), static_cast<unsigned>((nrows + 63) / 64), 1}, {256, 1, 1}, 0,
                [&] { data_absmax_kernel<float>(src, ld, row0, nrows, ncols, rowmax, colmax); });
}

void emu_ingest(const float* src, long ld, int row0, int nrows, int ncols, std::uint16_t* rowp, long rowp_plane,
                int rowp_rows_p, std::uint16_t* colp, long colp_plane, int colp_rows_p, int nv, unsigned* stats,
                const unsigned* rowmax, const unsigned* colmax, int row_lo, int row_n, int col_lo, int col_n) {
    emu::launch({static_cast<unsigned>((ncols + 63) / 64), static_cast<unsigned>((nrows + 63) / 64), 1}, {256, 1, 1}, 0,
                [&] {
                    ingest_data_kernel<float>(src, ld, row0, nrows, ncols, reinterpret_cast<bf16*>(rowp), rowp_plane,
                                              rowp_rows_p, reinterpret_cast<bf16*>(colp), colp_plane, colp_rows_p,
                                              nv, stats, rowmax, colmax, row_lo, row_n, col_lo, col_n);
                });
}

void emu_fixed_planes(const float* master, int KP, int Lp, std::uint16_t* cpl, long cpl_plane) {
    emu::launch({static_cast<unsigned>(KP), 1, 1}, {256, 1, 1}, 0,
                [&] { fixed_planes_kernel(master, Lp, reinterpret_cast<bf16*>(cpl), cpl_plane); });
}

// mur_eu_update_launch / mur_kl_update_launch: grid = window / 64, 256 threads
void emu_mur_eu_update(int KP, const float* part, int slots, const float* gram, float* master, std::uint16_t* cpl,
                       std::uint16_t* rpl, int k, int L, int Lp, int nrows_p, float lambda) {
    const unsigned grid = static_cast<unsigned>(nrows_p / 64);
    const size_t smem = (static_cast<size_t>(KP) * 64 + 32 * static_cast<size_t>(KP)) * sizeof(float);
    bf16 *c = reinterpret_cast<bf16*>(cpl), *r = reinterpret_cast<bf16*>(rpl);
    const long cp = 256L * Lp, rp = static_cast<long>(Lp) * KP;
    std::function<void()> body;
    if (KP == 64) body = [&] { mur_eu_update_kernel<4>(part, slots, gram, master, c, cp, r, rp, k, L, Lp, lambda); };
    else if (KP == 128) body = [&] { mur_eu_update_kernel<8>(part, slots, gram, master, c, cp, r, rp, k, L, Lp, lambda); };
    else body = [&] { mur_eu_update_kernel<16>(part, slots, gram, master, c, cp, r, rp, k, L, Lp, lambda); };
    emu::launch({grid, 1, 1}, {256, 1, 1}, smem, body);
}

void emu_mur_kl_update(int KP, const float* part, int slots, const double* bsum, float* master, std::uint16_t* cpl,
                       std::uint16_t* rpl, int k, int L, int Lp, int nrows_p, float lambda) {
    const unsigned grid = static_cast<unsigned>(nrows_p / 64);
    bf16 *c = reinterpret_cast<bf16*>(cpl), *r = reinterpret_cast<bf16*>(rpl);
    const long cp = 256L * Lp, rp = static_cast<long>(Lp) * KP;
    std::function<void()> body;
    if (KP == 64) body = [&] { mur_kl_update_kernel<4>(part, slots, bsum, master, c, cp, r, rp, k, L, Lp, lambda); };
    else if (KP == 128) body = [&] { mur_kl_update_kernel<8>(part, slots, bsum, master, c, cp, r, rp, k, L, Lp, lambda); };
    else body = [&] { mur_kl_update_kernel<16>(part, slots, bsum, master, c, cp, r, rp, k, L, Lp, lambda); };
    emu::launch({grid, 1, 1}, {256, 1, 1}, 0, body);
}

void emu_factor_rowsum(const float* master, int KP, int L, int Lp, double* sums) {
    emu::launch({static_cast<unsigned>(KP), 1, 1}, {256, 1, 1}, 0, [&] { factor_rowsum_kernel(master, L, Lp, sums); });
}

void emu_gram_reduce(const float* part, int slots, int KP, int Lp, float* g32, double* g64) {
    emu::launch({static_cast<unsigned>((KP * KP + 255) / 256), 1, 1}, {256, 1, 1}, 0,
                [&] { gram_reduce_kernel(part, slots, KP, Lp, g32, g64); });
}

void emu_objective_simt(const std::uint16_t* vpl, long v_plane, int v_rows_p, int nv, const float* wt, int Mp,
                        const float* ht, int Np, int k, int m, int n, double* partial) {
    emu::launch({static_cast<unsigned>((n + 63) / 64), static_cast<unsigned>((m + 63) / 64), 1}, {256, 1, 1}, 0, [&] {
        objective_simt_kernel(reinterpret_cast<const bf16*>(vpl), v_plane, v_rows_p, nv, wt, Mp, ht, Np, k, m, n, partial);
    });
}

// gram_fp64_launch: upper-triangle 64 x 64 tiles x row splits on the FP64 tensor cores, then the fixed-order reduction
void emu_gram_fp64(const std::uint16_t* cpl, int KP, int Lp, int rows_per, int nsplit, double* scratch, float* g32,
                   double* g64) {
    const int nt = KP / 64;
    emu::launch({static_cast<unsigned>(nt * (nt + 1) / 2), static_cast<unsigned>(nsplit), 1}, {256, 1, 1}, 0, [&] {
        gram_fp64_kernel(reinterpret_cast<const bf16*>(cpl), 256L * Lp, Lp, KP, rows_per, scratch);
    });
    emu::launch({static_cast<unsigned>((KP * KP + 255) / 256), 1, 1}, {256, 1, 1}, 0,
                [&] { gram_fp64_reduce_kernel(scratch, nsplit, KP, g32, g64); });
}

// refresh_planes_launch: grid = (Lp / 256, KP / 16)
void emu_refresh_planes(const float* src, float* adopt, int KP, int Lp, std::uint16_t* cpl, std::uint16_t* rpl,
                        int skip_lo, int skip_hi) {
    emu::launch({static_cast<unsigned>((Lp + 255) / 256), static_cast<unsigned>(KP / 16), 1}, {256, 1, 1}, 0, [&] {
        refresh_planes_kernel(src, adopt, KP, Lp, reinterpret_cast<bf16*>(cpl), 256L * Lp, reinterpret_cast<bf16*>(rpl),
                              static_cast<long>(Lp) * KP, skip_lo, skip_hi);
    });
}

}  // extern "C"
