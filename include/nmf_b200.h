/* nmf_b200 -- C ABI of the B200 (sm_100a) NMF solver hot path.
 *
 * The reference (raleng/nmf) is pure Python and has no FFI of its own; its solver loops are the
 * boundary: `mur.mur` (nmf/mur.py:119-136), `admm.admm` (nmf/admm.py:292-334) and
 * `ao_admm.ao_admm` (nmf/ao_admm.py:259-301), reached from `NMF.factorize` (nmf/nmf.py:59-73).
 * Each entry point below names the reference lines whose arithmetic it replaces.  A maintainer
 * binds this library with ctypes; `nmf_b200/_lib.py` is that binding.
 *
 * Conventions
 *   - every function returns 0 on success and a negative nmfb_status otherwise; the message is in
 *     nmfb_last_error() (thread-local).  Nothing throws across the boundary.
 *   - the CALLER owns all memory.  Pointers are device pointers unless a name ends in `_host`.
 *     The library allocates nothing and keeps no global state besides one-time kernel attributes.
 *   - all work is enqueued on the `stream` argument (a cudaStream_t passed as void*); no call
 *     synchronises the device.
 *   - scalars that the host needs (objective, flags) are written to caller-provided device slots;
 *     the caller copies them out when it wants them.
 *
 * Device layouts (all padded regions must be zero-initialised by the caller; the library keeps
 * them zero):  k <= 256 components are padded to KP = 64, 128 or 256;  a length L (m or n) is
 * padded to Lp = round_up(L, 256).
 *   data planes  bf16 [nv][ld/64][rows_p][64]  K-blocked planes of V (rows = m) or V^T (rows = n):
 *                                        element (r, c) lives at ((c/64)*rows_p + r)*64 + c%64, so the
 *                                        [rows x 64] box of one contraction block is contiguous in HBM;
 *                                        rows_p = round_up(rows, 256), ld = round_up(cols, 256);
 *                                        plane 0 = hi, plane 1 = lo (nv = 1 when V is bf16-exact)
 *   master       fp32 [KP][Lp]           solver state of a factor, component-major (W^T or H)
 *   cplanes      bf16 [3][256][Lp]       hi / lo / lo2 planes of the master, component-major
 *   rplanes      bf16 [2][Lp][KP]        hi / lo planes, row-major (W or H^T)
 *   gram         fp32 [KP][KP] (+ fp64)  Gram matrix of a factor
 *   partial      fp32 [slots][KP][Lp]    per-(K-split) partial sums written by the panel GEMM
 */
#ifndef NMF_B200_H
#define NMF_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef enum nmfb_status {
    NMFB_OK = 0,
    NMFB_ERR_ARG = -1,       /* bad argument / unsupported shape */
    NMFB_ERR_CUDA = -2,      /* a CUDA runtime / driver call failed */
    NMFB_ERR_WORKSPACE = -3, /* caller workspace too small */
    NMFB_ERR_NOT_PD = -4     /* Gram + rho I not positive definite (numpy.linalg.LinAlgError) */
} nmfb_status;

typedef struct nmfb_data {          /* V (rows = m, cols = n) or V^T (rows = n, cols = m) */
    void* planes;                   /* bf16 [nv][ld/64][rows_p][64] (K-blocked) or [nv][rows_p][ld] */
    int rows, cols, rows_p, ld, nv;
    int blocked;                    /* 1: K-blocked layout (V, V^T); 0: plain row-major           */
    int fixed;                      /* 1: plane 0 is the fixed-point slice of each row (exact-mode   */
                                    /*    ADMM contractions), planes 1.. the floating remainder       */
    int origin;                     /* sharded runs: global index (row of V, or column of V for V^T) of   */
                                    /* local row 0 of these planes; 0 otherwise                           */
} nmfb_data;

typedef struct nmfb_factor {        /* W (L = m) or H (L = n) */
    float* master;                  /* fp32 [KP][Lp] */
    void* cplanes;                  /* bf16 [3][256][Lp] */
    void* rplanes;                  /* bf16 [2][Lp][KP] */
    int L, Lp;
    int win0, win_p;                /* sharded runs: this rank updates rows [win0, win0 + win_p) of the   */
                                    /* factor and holds exactly those rows of V (win_p = 0: all rows)     */
} nmfb_factor;

typedef struct nmfb_problem {
    int m, n, k, KP;
    nmfb_data v_rows;               /* V   : data for the W-side contraction V.H^T            */
    nmfb_data v_cols;               /* V^T : data for the H-side contraction (W^T.V)^T        */
    nmfb_factor w, h;
    float* gram32;                  /* fp32 [2][KP][KP]: [0] = W^T W, [1] = H H^T             */
    double* gram64;                 /* fp64 [2][KP][KP]                                       */
    float* partial;                 /* GEMM partial-sum workspace                             */
    size_t partial_bytes;
    double* obj_partial;            /* fp64 [obj_partial_len] per-CTA objective partials      */
    int obj_partial_len;
    int num_sms;                    /* SM count of the device (148 on B200)                   */
    void** phase_events;            /* optional cudaEvent_t[n_phase_events] recorded at the phase */
    int n_phase_events;             /* boundaries of a solver step (profiling; NULL = off)        */
    double* ksum;                   /* fp64 [2][KP]: row sums of H / column sums of W (MUR-KL)    */
} nmfb_problem;

const char* nmfb_last_error(void);
int nmfb_abi_version(void);
/* kernels launched by this process through the library so far (for benchmarks' launch counts) */
unsigned long long nmfb_launch_count(void);

/* Padding rules, so that callers size buffers the way the kernels expect. */
int nmfb_padded_components(int k);                 /* 64 / 128 / 256, or NMFB_ERR_ARG if k > 256 */
int nmfb_padded_length(int L);                     /* round_up(L, 256) */
size_t nmfb_partial_bytes(int m, int n, int k, int num_sms);   /* workspace for any GEMM of the problem */
size_t nmfb_gram_partial_bytes(int k, int L, int num_sms);     /* scratch of a Gram over a factor of length L */
int nmfb_obj_partial_len(int m, int n, int num_sms);

/* ---- ingest / export ------------------------------------------------------------------------ */
/* Split a chunk of `nrows` rows of the caller's V (fp32 or fp64, row-major, pitch src_ld elements -- or, with
 * src_ld < 0, K-blocked [cols / 64][-src_ld][64] like the data planes --, already on the device) starting at global row `row0` into the row-major planes of `v_rows` and the
 * transposed planes of `v_cols`.  stats[0] = order-preserving key of min(V) (init 0xFFFFFFFF),
 * stats[1] = 1 if a lo plane is needed, stats[2] = 1 if a non-finite value was seen.
 * Replaces the implicit float64 view of `x` in mur.py:52 / admm.py:233 / ao_admm.py:201. */
int nmfb_ingest_data(const void* src, int src_is_f64, long src_ld, int row0, int nrows,
                     const nmfb_data* v_rows, const nmfb_data* v_cols, uint32_t* stats, void* stream);
/* Fixed-point variant (v_rows->fixed / v_cols->fixed): plane 0 = x rounded to a multiple of 2^(E-6), E the
 * exponent of max|V| over the row of that copy (row of V for v_rows, column of V for v_cols), plane 1 =
 * the bf16 remainder; stats[1] = 1 when two planes do not hold the data exactly.  rowmax [m] / colmax [n]
 * are filled by nmfb_data_absmax over ALL chunks first (zero-initialised by the caller). */
int nmfb_ingest_data_fixed(const void* src, int src_is_f64, long src_ld, int row0, int nrows,
                           const nmfb_data* v_rows, const nmfb_data* v_cols, uint32_t* stats,
                           const uint32_t* rowmax, const uint32_t* colmax, void* stream);
int nmfb_data_absmax(const void* src, int src_is_f64, long src_ld, int row0, int nrows, int ncols,
                     uint32_t* rowmax, uint32_t* colmax, void* stream);
/* Re-derive the cplanes of a factor in fixed-point form (plane 0 = 7-bit slice relative to the row
 * maximum, planes 1-2 = floating remainders): operands of nmfb_panel_gemm_exact / nmfb_gram_exact. */
int nmfb_fixed_planes(int KP, const nmfb_factor* f, void* stream);

/* Import an initial factor (mur.py:108-109, admm.py:23-24, ao_admm.py:22-23 or utils.nndsvd).
 * Element (r, c) of the factor is src[r*stride_l + c*stride_c]; W [m][k]: (k, 1); H [k][n]: (1, n).
 * Fills master, cplanes and rplanes. */
int nmfb_import_factor(const void* src, int src_is_f64, long stride_l, long stride_c, int k, int KP,
                       const nmfb_factor* f, void* stream);
/* Re-derive cplanes / rplanes from the master. */
int nmfb_refresh_planes(int KP, const nmfb_factor* f, void* stream);
/* Export the master to a caller buffer (device) with the same stride convention. */
int nmfb_export_factor(const nmfb_factor* f, int k, int KP, void* dst, int dst_is_f64, long stride_l,
                       long stride_c, void* stream);

/* ---- building blocks ------------------------------------------------------------------------ */
/* partial[slot][c][r] = sum_t factor_plane[c][t] * data_plane[r][t] over the slot's K range
 * (`x @ h.T` mur.py:29, `w.T @ x` mur.py:45, `other_aux.T @ data_aux` admm.py:221, `w.T @ y`
 * ao_admm.py:56).  nb = number of factor planes used (2 for non-negative factors, 3 for signed
 * ADMM auxiliaries).  Returns the number of slots written (> 0) or a negative status.
 * force_bn / force_split = 0 lets the library choose. */
int nmfb_panel_gemm(const nmfb_data* data, const void* factor_cplanes, int factor_Lp, int KP, int nb,
                    float* partial, size_t partial_bytes, int out_Lp, int num_sms, int force_bn,
                    int force_split, void* stream);

/* Exact-hi contraction for the ADMM right-hand sides (admm.py:221, ao_admm.py:56): fixed-point planes on
 * both sides, the leading plane pair accumulated without rounding, (hi, lo) fp32 slot pairs out.
 * Returns the number of slots (sum them in fp64) or a negative status. */
int nmfb_panel_gemm_exact(const nmfb_data* data, const void* factor_cplanes, int factor_Lp, int KP,
                          float* partial, size_t partial_bytes, int out_Lp, int num_sms, void* stream);
int nmfb_gram_exact(const nmfb_factor* f, int KP, float* partial, size_t partial_bytes, float* gram32,
                    double* gram64, int num_sms, void* stream);

/* Gram matrix G = F F^T of a factor from its cplanes (`w.T @ w` ao_admm.py:53, `other_aux.T @
 * other_aux` admm.py:220, and the W (H H^T) / (W^T W) H re-association of mur.py:29,45).
 * nplanes = 2: hi/lo planes (16-bit operands, MUR), 3: hi/lo/lo2 (24-bit operands, ADMM / AO-ADMM).
 * Uses `partial` as scratch; writes gram32 [KP][KP] and gram64 [KP][KP]. */
int nmfb_gram(const nmfb_factor* f, int KP, int nplanes, float* partial, size_t partial_bytes, float* gram32,
              double* gram64, int num_sms, void* stream);

/* The Gram matrix of the ADMM / AO-ADMM systems (`other_aux.T @ other_aux` admm.py:220, `w.T @ w` ao_admm.py:53):
 * products and sums in fp64 on the FP64 tensor cores, operand = the sum of the factor's three bf16 cplanes (the very
 * operand the numerator contraction reads, so that the k x k solve sees a consistent system).
 * scratch >= nmfb_gram_partial_bytes(k, L, num_sms). */
int nmfb_gram_fp64(const nmfb_factor* f, int KP, double* scratch, size_t scratch_bytes, float* gram32, double* gram64,
                   void* stream);

/* nmfb_gram without a caller workspace, for callers that cannot allocate device memory (CUDA graph capture, serving
 * loops, harnesses that forbid cudaMalloc).  NO COUNTERPART IN THE REFERENCE.
 * Same kernels and results as nmfb_gram; the K-split slots live in a device array built into the library
 * (nmfb_gram_static_max_bytes() = 8 MiB, held by every GPU that loads the library until the process exits).  Calls no
 * cudaMalloc / cudaFree / allocating routine.  A shape whose slots exceed the array (k > 128 with L above ~32 K) is
 * rejected with NMFB_ERR_ARG and a message naming both sizes -- use nmfb_gram there.  All calls on one GPU share the
 * array: calls that overlap on the GPU (different streams, or graphs replayed concurrently) race; calls on one stream
 * are ordered.  The array keeps the last call's contents; nothing needs clearing, every slot word read is written by
 * the same call. */
size_t nmfb_gram_static_max_bytes(void);
int nmfb_gram_static(const nmfb_factor* f, int KP, int nplanes, float* gram32, double* gram64, int num_sms,
                     void* stream);

/* Fused multiplicative update  x <- x * num / (G x + lambda x + 1e-9)  (mur.py:29 / mur.py:45):
 * num = sum of `slots` partial slots, G = Gram of the other factor.  Writes master + all planes. */
int nmfb_mur_eu_update(const float* partial, int slots, const float* gram_other, const nmfb_factor* f,
                       int k, int KP, float lambda, void* stream);

/* 0.5 * ||V - W H||_F^2 (utils.py:29) in residual form, fp64 accumulation, written to *obj_out
 * (device).  impl = 0: the library's choice -- the wide-tile tcgen05 kernel (128 x 256 tiles) for KP = 64 / 128 with
 * one V plane, the tcgen05 kernel with the W planes as the TMEM-resident A operand for every other shape;
 * 1: CUDA-core kernel (cross-check); 2 / 3: force the wide-tile / TMEM-operand kernel (NMFB_ERR_ARG when the shape
 * does not fit it). */
int nmfb_objective_eu(const nmfb_problem* p, int impl, double* obj_out, void* stream);

/* ---- solver steps --------------------------------------------------------------------------- */
/* One outer iteration of MUR-Euclidean (mur.py:122-127): W update with the old H, H update with the
 * new W, then the objective of the new pair into *obj_out (skipped when obj_out == NULL). */
int nmfb_mur_eu_step(const nmfb_problem* p, float lambda_w, float lambda_h, double* obj_out,
                     void* stream);
/* One factor update of the iteration above (side 0 = W, 1 = H).  In a sharded run (factor windows set,
 * v_rows / v_cols holding this rank's rows / columns) it updates the window only; the caller then
 * exchanges the windows (nmfb_peer_push / nmfb_peer_barrier / nmfb_adopt_exchange) before the next half step. */
int nmfb_mur_eu_half_step(const nmfb_problem* p, int side, float lambda, void* stream);

/* Generalised Kullback-Leibler objective sum(v log(v/wh) - v + wh) with the reference's special values
 * (utils.py:23-26), W.H formed tile by tile on the tensor cores, fp64 accumulation. */
int nmfb_objective_kl(const nmfb_problem* p, double* obj_out, void* stream);

/* One outer iteration of MUR-Kullback-Leibler (mur.py:122-127 with distance_type='kl'): both updates
 * through the fused GEMM -> quotient -> GEMM kernel (W.H never reaches HBM), row / column sums by
 * shuffle reductions, then the objective of the new pair.  k <= 128 in this build. */
int nmfb_mur_kl_step(const nmfb_problem* p, float lambda_w, float lambda_h, double* obj_out, void* stream);
/* One factor update of the iteration above (side 0 = W, 1 = H); window semantics as nmfb_mur_eu_half_step.
 * obj_start_out != NULL: the pass also writes the KL objective (utils.py:23-26) of the pair it STARTS from
 * (this rank's rows / columns of V only in a sharded run) -- S = W.H and V are in registers anyway, so the
 * objective of outer iteration i costs no sweep of its own when taken from the W pass of iteration i + 1. */
int nmfb_mur_kl_half_step(const nmfb_problem* p, int side, float lambda, double* obj_start_out, void* stream);
/* W update (emitting the objective of the incoming pair, i.e. of the previous outer iteration, into
 * *obj_start_out) followed by the H update; no objective of the new pair.  The host loop reads objective
 * i - 1 after enqueueing iteration i and calls nmfb_objective_kl once for the last iteration. */
int nmfb_mur_kl_step_lagged(const nmfb_problem* p, float lambda_w, float lambda_h, double* obj_start_out,
                            void* stream);

/* ---- ADMM / AO-ADMM ------------------------------------------------------------------------- */
typedef enum nmfb_prox { NMFB_PROX_NN = 0, NMFB_PROX_L1N = 1, NMFB_PROX_L2N = 2 } nmfb_prox;   /* admm.py:126-156 */

typedef struct nmfb_admm_state {
    nmfb_factor w_aux, h_aux;       /* ADMM only: signed auxiliaries (admm.py:27-28); cplanes hold 3 planes */
    float* dual_w;                  /* fp32 [KP][Mp]  scaled dual of W (component-major, like the master) */
    float* dual_h;                  /* fp32 [KP][Np]                                                     */
    double* ainv;                   /* fp64 [2][KP][KP]: inverse of (Gram + rho I); [0] H-side, [1] W-side */
    double* chol_scratch;           /* fp64 [KP][KP+1]  working matrix when KP = 256 (else unused)        */
    double* c64;                    /* fp64 [KP][max(Mp,Np)]  reduced numerator (AO-ADMM inner loop)      */
    double* scal;                   /* fp64 [4]: [0] rho of the H-side system, [1] rho of the W-side      */
    int* flags;                     /* int  [8]: [0] not-PD (sticky) [1] inner done [2] inner iterations  */
                                    /*           used by the H update [3] ... by the W update             */
    double* norm_partial;           /* fp64 [4 * max(Mp,Np)/64]                                           */
    int exact;                      /* 1: V planes and factor cplanes are fixed-point; use the exact-hi   */
                                    /*    contractions (needs problem.v_rows.fixed / v_cols.fixed)        */
    double* binv;                   /* fp64 [2][KP][KP] or NULL: inverse Tikhonov systems of prox l2n,    */
                                    /* ((lambda/rho) T^T T + I)^-1, [0] H-side, [1] W-side (admm.py:141-156) */
    double* tik_scratch;            /* fp64 [KP][KP] + [4] scratch for building them (NULL iff binv NULL) */
    /* Optional overlap of the k x k work with the m x n x k numerator GEMM (all NULL / 0: everything on `stream`).
     * The Gram, its Cholesky inverse (a single-CTA latency chain, ~0.2 ms at k = 128) and the Tikhonov system do not
     * depend on the numerator; with a side stream they run next to it: the half step records ev_fork on `stream`,
     * the side stream waits for it, does the k x k work with gram_partial as its scratch, records ev_join, and
     * `stream` waits for ev_join before the solve / prox / dual kernel.  The numerator GEMM then leaves two SMs
     * free for the single-CTA kernels. */
    void* side_stream;              /* cudaStream_t or NULL                                               */
    void* ev_fork;                  /* cudaEvent_t (timing disabled is fine)                              */
    void* ev_join;                  /* cudaEvent_t                                                        */
    float* gram_partial;            /* scratch of the Gram contraction on the side stream                 */
    size_t gram_partial_bytes;      /* >= nmfb_gram_partial_bytes(k, max(m, n), num_sms)                  */
} nmfb_admm_state;

/* prox l2n with lambda > 0 (k <= 128): (re)build s->binv[side] from lambda and the rho of that side's system
 * (s->scal[side], valid once the side's Cholesky has been enqueued).  nmfb_admm_half_step, nmfb_admm_step and
 * nmfb_aoadmm_step call it themselves; a caller of nmfb_aoadmm_side_begin calls it right after that. */
int nmfb_admm_prepare_l2n(const nmfb_problem* p, const nmfb_admm_state* s, int side, double lambda, void* stream);

/* ainv = (gram + rho I)^-1 by a shared-memory Cholesky (fp64), rho = rho_in (rho_mode 0, admm.py:220)
 * or trace(gram)/k (rho_mode 1, ao_admm.py:54); *rho_out = rho.  flags[0] is raised when the matrix is
 * not positive definite (scipy.linalg.cholesky / numpy.linalg.solve -> LinAlgError). */
int nmfb_chol_inverse(const double* gram64, int k, int KP, int rho_mode, double rho_in, double* scratch,
                      double* ainv, double* rho_out, int* flags, void* stream);

/* One outer iteration of ADMM, Euclidean (admm.py:294-301, 320-324): h_aux, w_aux (with the new h_aux),
 * both prox steps with the old duals, both dual updates, objective of the prox'd pair.
 * mailbox[0] = objective, mailbox[3] = not-PD flag. */
int nmfb_admm_step(const nmfb_problem* p, const nmfb_admm_state* s, double rho, float lambda_w, int prox_w,
                   float lambda_h, int prox_h, double* mailbox, void* stream);

/* Sharded runs drive the iteration above in two halves.  side 0: h_aux / h / U_h (admm.py:295,299,320),
 * side 1: w_aux / w / U_w with the new h_aux (admm.py:296-297,300,321); only the rank's window of the
 * factor, its dual and its auxiliary is touched.  The caller then all-gathers the auxiliary (and the
 * factor) masters and re-derives their planes (nmfb_refresh_planes, nmfb_fixed_planes in exact mode). */
int nmfb_admm_half_step(const nmfb_problem* p, const nmfb_admm_state* s, int side, double rho, float lambda,
                        int prox, void* stream);
/* mailbox[1], [2] = inner iterations used by the H / W update (AO-ADMM), mailbox[3] = not-PD flag. */
int nmfb_admm_publish(const nmfb_admm_state* s, double* mailbox, void* stream);

/* Kullback-Leibler branch of ADMM / AO-ADMM, the element-wise data step (admm.py:310-313,
 * ao_admm.py:89-94).  prod [rows][ld] fp32 holds P = W_aux.H_aux (admm) or a.aux (ao_admm), rows =
 * v_rows->rows of this rank; with V read from the planes of v_rows:  v_bar = P - dual_v,
 * v_aux = ((v_bar - 1) + sqrt((v_bar - 1)^2 + 4 V)) / 2,  dual_v += v_aux - P,  d_out <- v_aux + dual_v
 * (d_out may be prod): the data term a^T (v_aux + dual_v) of the next right-hand sides (admm.py:224,
 * ao_admm.py:83).  Ingest d_out with nmfb_data_absmax + nmfb_ingest_data_fixed and run the half-steps on a
 * problem whose v_rows / v_cols are those planes.  done_flag (device, may be NULL): nothing is touched
 * when *done_flag != 0 (AO-ADMM: the inner loop has broken, ao_admm.py:96-98). */
int nmfb_kl_vaux(const nmfb_data* v_rows, const float* prod, float* d_out, long ld, float* dual_v,
                 const int* done_flag, void* stream);

/* One outer iteration of AO-ADMM, Euclidean (ao_admm.py:261-270, 291): per factor Gram, rho = tr/k,
 * Cholesky, numerator GEMM, then up to admm_iter x (solve, prox, dual, residual test ao_admm.py:33-43)
 * with the early break decided on the device.  mailbox[0] = objective, [1] / [2] = inner iterations
 * completed without a break for H / W (== admm_iter when no break), [3] = not-PD flag. */
int nmfb_aoadmm_step(const nmfb_problem* p, const nmfb_admm_state* s, int admm_iter, float lambda_w,
                     int prox_w, float lambda_h, int prox_h, double* mailbox, void* stream);

/* The pieces of nmfb_aoadmm_step, for sharded runs (side 0 = H update, 1 = W update):
 *   side_begin  flags reset, Gram of the fixed factor, rho = tr/k, Cholesky inverse, numerator GEMM for the
 *               rank's window (ao_admm.py:53-56); returns the number of partial slots
 *   inner       one solve / prox / dual pass over the window (ao_admm.py:59-62) leaving the window's blocks of
 *               the four norm partials in s->norm_partial; `first` != 0 reduces the GEMM slots into s->c64
 *   inner_flag  the residual test of ao_admm.py:33-43 over ALL blocks of norm_partial (the caller all-gathers
 *               the windows first), deciding the early break on the device
 * After the inner loop the caller all-gathers the factor and re-derives its planes. */
int nmfb_aoadmm_side_begin(const nmfb_problem* p, const nmfb_admm_state* s, int side, void* stream);
int nmfb_aoadmm_inner(const nmfb_problem* p, const nmfb_admm_state* s, int side, int first, int slots,
                      float lambda, int prox, void* stream);
int nmfb_aoadmm_inner_flag(const nmfb_problem* p, const nmfb_admm_state* s, int side, void* stream);

/* ---- Kullback-Leibler branch of ADMM / AO-ADMM: the data step ----------------------------------- */
/* planes3 [3][Lp][KP] <- three row-major bf16 planes (24-bit significands) of the factor's master. */
int nmfb_row_planes3(int KP, const nmfb_factor* f, void* planes3, void* stream);
/* admm.py:309-313 (`w_aux @ h_aux`, v_aux, dual_v) and ao_admm.py:88-94 (`a @ aux`, ...) fused on this rank's rows
 * of V: P = A.B is formed tile by tile in tensor memory from the planes of nmfb_row_planes3 (A: rows a_row0.. of the
 * left factor, B: the right factor) and never written; dual_v is updated in place, d_out = v_aux + dual_v (both fp32,
 * K-blocked like the data planes: [ld / 64][v_rows->rows_p][64], ld >= round_up(cols, 128)); rowmax [origin + rows] /
 * colmax [cols] receive the |max| bit patterns of D for the fixed-point re-ingest (nmfb_ingest_data_fixed with
 * src_ld = -rows_p).  *done_flag != 0: nothing is touched.
 * k <= 128.  Replaces the cuBLAS SGEMM + element-wise kernel + |max| pass of round 1. */
int nmfb_kl_data_step(const nmfb_data* v_rows, const void* a_planes3, int a_Lp, int a_row0, const void* b_planes3,
                      int b_Lp, int KP, float* dual_v, float* d_out, long ld, uint32_t* rowmax, uint32_t* colmax,
                      const int* done_flag, int num_sms, void* stream);

/* ---- sharded runs: exchange over NVLink peer memory ------------------------------------------- */
/* SURVEY.md 8e (1-D MPI-FAUN layout): after a half step every rank owns fresh rows [win0, win0 + win_p) of the
 * factor; the reference has no counterpart (it is single-process).  The exchange is done by the ranks' own
 * kernels on peer pointers -- `peers_host[p]` is the device address, valid on THIS device, of rank p's exchange
 * arena (symmetric memory or CUDA IPC, opened by the caller; index `rank` is the local arena) --, never by a
 * collective library:
 *   peer_push       copies `rows` rows of `row_bytes` bytes from src (pitch src_pitch) to byte offset dst_off of
 *                   EVERY rank's arena (pitch dst_pitch): the window of a master, a Gram term, a scalar
 *   peer_barrier    cross-GPU barrier on the stream: publishes `seq` (a counter the callers advance together) to
 *                   every rank's flag array and waits until all ranks have published >= seq; the stores of
 *                   earlier kernels on the stream are visible to the peers that pass it
 *   sum_rank_slots  out[i] = sum_p slots[p * stride + i] in rank order (bit-identical on all ranks): the k x k
 *                   Gram all-reduce (`w.T @ w` ao_admm.py:53, the re-associated Grams of mur.py:29,45) and the
 *                   objective sum (utils.py:29 over the ranks' rows)
 *   adopt_exchange  the factor's master <- the exchange buffer ([KP][Lp] like the master, all windows current);
 *                   refresh != 0 also re-derives cplanes / rplanes in the same pass; refresh == 2 leaves the rows
 *                   of the rank's own window alone (its update kernel wrote master and planes itself: MUR-eu)
 * Buffers must be double-buffered by the caller (see nmf_b200/csrc/peer.cu for the ordering argument). */
int nmfb_peer_push(const void* src, long src_pitch, void* const* peers_host, int world, long dst_off,
                   long dst_pitch, int rows, long row_bytes, void* stream);
/* All-to-all of row blocks over peer memory (the sharded upload): rank q's arena receives, at dst_off / dst_pitch,
 * `rows` rows of row_bytes_host[q] bytes taken from src + q * src_peer_stride (pitch src_pitch) -- the columns of
 * this rank's rows of V that belong to rank q.  After nmfb_peer_barrier every rank holds ALL rows of its columns,
 * although each row shard crossed PCIe exactly once. */
int nmfb_peer_scatter(const void* src, long src_pitch, long src_peer_stride, void* const* peers_host, int world,
                      long dst_off, long dst_pitch, int rows, const long* row_bytes_host, void* stream);
/* peer_push through the NVSwitch multicast mapping of the arenas (mc_base: the multicast device address that aliases
 * every rank's arena, from the caller's symmetric-memory handle): one multimem.st per 16 bytes, replicated by the
 * switch into all ranks -- 1/world of the unicast version's NVLink egress. */
int nmfb_peer_push_mc(const void* src, long src_pitch, void* mc_base, long dst_off, long dst_pitch, int rows,
                      long row_bytes, void* stream);
int nmfb_peer_barrier(void* const* flags_host, int rank, int world, unsigned int seq, void* stream);
int nmfb_sum_rank_slots(const double* slots, int world, long stride, int n, double* out64, float* out32,
                        void* stream);
int nmfb_adopt_exchange(int KP, const nmfb_factor* f, const float* xchg, int refresh, void* stream);

/* The whole exchange of a factor window in ONE call (what the building blocks above do, sequenced here so that the
 * host issues one call instead of eight: at 8 GPUs an iteration is ~1 ms and the per-launch cost of the caller's
 * language decides whether the GPUs wait for their host):
 *   1. store this rank's window of the master into every rank's exchange buffer at byte offset xchg_off of the arenas
 *      (multicast when ctx->mc_base != NULL, unicast otherwise);
 *   2. gram_nplanes != 0: on ctx->side_stream (forked / joined with ctx->ev_fork / ev_join), next to 1., contract this
 *      rank's term of F F^T over its window (nmfb_gram_window) and store it at gram_off + rank * KP*KP*8 of every arena;
 *   3. cross-GPU barrier `seq` (the caller's per-arena counter);
 *   4. adopt all windows (refresh as in nmfb_adopt_exchange), fixed != 0: nmfb_fixed_planes on top;
 *   5. gram_nplanes != 0: sum the ranks' terms in rank order into gram64 / gram32 (k x k all-reduce).
 * probe_events: NULL or five cudaEvent_t recorded on `stream` before 1. and after 1.+2., 3., 4., 5. */
typedef struct {
    void* const* peers;             /* [world] base pointers of the ranks' arenas (host array)                 */
    void* mc_base;                  /* multicast alias of the arenas, or NULL                                   */
    int rank, world;
    void* side_stream;              /* stream / events of step 2 (all NULL: step 2 runs on `stream`)            */
    void* ev_fork;
    void* ev_join;
    float* gram_scratch;            /* >= nmfb_gram_partial_bytes(k, window length, num_sms)                    */
    size_t gram_scratch_bytes;
    double* gram_term64;            /* [KP][KP] this rank's term (source of the store)                          */
    float* gram_term32;             /* [KP][KP]                                                                 */
    int num_sms;
} nmfb_peer_ctx;
int nmfb_exchange_factor(const nmfb_peer_ctx* ctx, int KP, const nmfb_factor* f, long xchg_off, int refresh, int fixed,
                         int gram_nplanes, long gram_off, double* gram64, float* gram32, unsigned int seq,
                         void* const* probe_events, void* stream);
/* Sum of the double at src (16-byte aligned; the granule src[0..1] travels) over the ranks: stored into slot `rank`
 * of the table at table_off (64-byte slots) of every arena, barrier `seq`, summed in rank order into out[0] (the
 * objective: utils.py:29 over the ranks' rows). */
int nmfb_exchange_scalar(const nmfb_peer_ctx* ctx, const double* src, long table_off, unsigned int seq, double* out,
                         void* stream);
/* This rank's term of the Gram matrix: nmfb_gram over the rows of the factor's window only. */
int nmfb_gram_window(const nmfb_factor* f, int KP, int nplanes, float* partial, size_t partial_bytes,
                     float* gram32, double* gram64, int num_sms, void* stream);
/* nmfb_mur_eu_half_step without its Gram launch: gram32 already holds the Gram of the other factor (reduced from
 * the ranks' window terms during that factor's exchange). */
int nmfb_mur_eu_window_step(const nmfb_problem* p, int side, float lambda, void* stream);
/* Strided host -> device copy (cudaMemcpy2DAsync) of a column panel of the caller's host V: a sharded upload reads
 * the rank's rows AND its columns straight from host memory (2/P of V per PCIe link, no inter-GPU step). */
int nmfb_upload_2d(void* dst, long dst_pitch, const void* src_host, long src_pitch, long width_bytes, long height,
                   void* stream);

#ifdef __cplusplus
}
#endif
#endif /* NMF_B200_H */
