"""ctypes binding of libnmf_b200.so (the C ABI declared in include/nmf_b200.h).

There is deliberately no fallback: if the shared library is missing or a call fails, an
exception is raised -- the product path never computes on the CPU.
"""
from __future__ import annotations

import ctypes as C
import os

from . import _build

c_void_p, c_int, c_long, c_float, c_size_t = C.c_void_p, C.c_int, C.c_long, C.c_float, C.c_size_t


class NmfbError(RuntimeError):
    def __init__(self, code: int, msg: str):
        super().__init__(f"nmf_b200 error {code}: {msg}")
        self.code = code


P = C.POINTER


class Data(C.Structure):            # struct nmfb_data
    _fields_ = [("planes", c_void_p), ("rows", c_int), ("cols", c_int), ("rows_p", c_int),
                ("ld", c_int), ("nv", c_int), ("blocked", c_int), ("fixed", c_int), ("origin", c_int)]


class PeerCtx(C.Structure):         # struct nmfb_peer_ctx
    _fields_ = [("peers", c_void_p), ("mc_base", c_void_p), ("rank", c_int), ("world", c_int),
                ("side_stream", c_void_p), ("ev_fork", c_void_p), ("ev_join", c_void_p),
                ("gram_scratch", c_void_p), ("gram_scratch_bytes", c_size_t),
                ("gram_term64", c_void_p), ("gram_term32", c_void_p), ("num_sms", c_int)]


class Factor(C.Structure):          # struct nmfb_factor
    _fields_ = [("master", c_void_p), ("cplanes", c_void_p), ("rplanes", c_void_p),
                ("L", c_int), ("Lp", c_int), ("win0", c_int), ("win_p", c_int)]


class Problem(C.Structure):         # struct nmfb_problem
    _fields_ = [("m", c_int), ("n", c_int), ("k", c_int), ("KP", c_int),
                ("v_rows", Data), ("v_cols", Data), ("w", Factor), ("h", Factor),
                ("gram32", c_void_p), ("gram64", c_void_p),
                ("partial", c_void_p), ("partial_bytes", c_size_t),
                ("obj_partial", c_void_p), ("obj_partial_len", c_int), ("num_sms", c_int),
                ("phase_events", P(c_void_p)), ("n_phase_events", c_int), ("ksum", c_void_p)]


class AdmmState(C.Structure):       # struct nmfb_admm_state
    _fields_ = [("w_aux", Factor), ("h_aux", Factor), ("dual_w", c_void_p), ("dual_h", c_void_p),
                ("ainv", c_void_p), ("chol_scratch", c_void_p), ("c64", c_void_p), ("scal", c_void_p),
                ("flags", c_void_p), ("norm_partial", c_void_p), ("exact", c_int),
                ("binv", c_void_p), ("tik_scratch", c_void_p),
                ("side_stream", c_void_p), ("ev_fork", c_void_p), ("ev_join", c_void_p),
                ("gram_partial", c_void_p), ("gram_partial_bytes", c_size_t)]


_SIGNATURES = {
    "nmfb_last_error": (C.c_char_p, []),
    "nmfb_abi_version": (c_int, []),
    "nmfb_launch_count": (C.c_ulonglong, []),
    "nmfb_padded_components": (c_int, [c_int]),
    "nmfb_padded_length": (c_int, [c_int]),
    "nmfb_partial_bytes": (c_size_t, [c_int, c_int, c_int, c_int]),
    "nmfb_obj_partial_len": (c_int, [c_int, c_int, c_int]),
    "nmfb_gram_partial_bytes": (c_size_t, [c_int, c_int, c_int]),
    "nmfb_ingest_data": (c_int, [c_void_p, c_int, c_long, c_int, c_int, P(Data), P(Data), c_void_p, c_void_p]),
    "nmfb_ingest_data_fixed": (c_int, [c_void_p, c_int, c_long, c_int, c_int, P(Data), P(Data), c_void_p, c_void_p,
                                       c_void_p, c_void_p]),
    "nmfb_data_absmax": (c_int, [c_void_p, c_int, c_long, c_int, c_int, c_int, c_void_p, c_void_p, c_void_p]),
    "nmfb_fixed_planes": (c_int, [c_int, P(Factor), c_void_p]),
    "nmfb_panel_gemm_exact": (c_int, [P(Data), c_void_p, c_int, c_int, c_void_p, c_size_t, c_int, c_int, c_void_p]),
    "nmfb_gram_exact": (c_int, [P(Factor), c_int, c_void_p, c_size_t, c_void_p, c_void_p, c_int, c_void_p]),
    "nmfb_import_factor": (c_int, [c_void_p, c_int, c_long, c_long, c_int, c_int, P(Factor), c_void_p]),
    "nmfb_refresh_planes": (c_int, [c_int, P(Factor), c_void_p]),
    "nmfb_export_factor": (c_int, [P(Factor), c_int, c_int, c_void_p, c_int, c_long, c_long, c_void_p]),
    "nmfb_panel_gemm": (c_int, [P(Data), c_void_p, c_int, c_int, c_int, c_void_p, c_size_t, c_int, c_int,
                                c_int, c_int, c_void_p]),
    "nmfb_gram": (c_int, [P(Factor), c_int, c_int, c_void_p, c_size_t, c_void_p, c_void_p, c_int, c_void_p]),
    "nmfb_gram_fp64": (c_int, [P(Factor), c_int, c_void_p, c_size_t, c_void_p, c_void_p, c_void_p]),
    "nmfb_gram_static_max_bytes": (c_size_t, []),
    "nmfb_gram_static": (c_int, [P(Factor), c_int, c_int, c_void_p, c_void_p, c_int, c_void_p]),
    "nmfb_mur_eu_update": (c_int, [c_void_p, c_int, c_void_p, P(Factor), c_int, c_int, c_float, c_void_p]),
    "nmfb_objective_eu": (c_int, [P(Problem), c_int, c_void_p, c_void_p]),
    "nmfb_mur_eu_step": (c_int, [P(Problem), c_float, c_float, c_void_p, c_void_p]),
    "nmfb_mur_eu_half_step": (c_int, [P(Problem), c_int, c_float, c_void_p]),
    "nmfb_objective_kl": (c_int, [P(Problem), c_void_p, c_void_p]),
    "nmfb_mur_kl_step": (c_int, [P(Problem), c_float, c_float, c_void_p, c_void_p]),
    "nmfb_mur_kl_half_step": (c_int, [P(Problem), c_int, c_float, c_void_p, c_void_p]),
    "nmfb_mur_kl_step_lagged": (c_int, [P(Problem), c_float, c_float, c_void_p, c_void_p]),
    "nmfb_admm_half_step": (c_int, [P(Problem), P(AdmmState), c_int, C.c_double, c_float, c_int, c_void_p]),
    "nmfb_admm_prepare_l2n": (c_int, [P(Problem), P(AdmmState), c_int, C.c_double, c_void_p]),
    "nmfb_admm_publish": (c_int, [P(AdmmState), c_void_p, c_void_p]),
    "nmfb_kl_vaux": (c_int, [P(Data), c_void_p, c_void_p, c_long, c_void_p, c_void_p, c_void_p]),
    "nmfb_aoadmm_side_begin": (c_int, [P(Problem), P(AdmmState), c_int, c_void_p]),
    "nmfb_aoadmm_inner": (c_int, [P(Problem), P(AdmmState), c_int, c_int, c_int, c_float, c_int, c_void_p]),
    "nmfb_aoadmm_inner_flag": (c_int, [P(Problem), P(AdmmState), c_int, c_void_p]),
    "nmfb_chol_inverse": (c_int, [c_void_p, c_int, c_int, c_int, C.c_double, c_void_p, c_void_p, c_void_p,
                                  c_void_p, c_void_p]),
    "nmfb_admm_step": (c_int, [P(Problem), P(AdmmState), C.c_double, c_float, c_int, c_float, c_int, c_void_p,
                               c_void_p]),
    "nmfb_aoadmm_step": (c_int, [P(Problem), P(AdmmState), c_int, c_float, c_int, c_float, c_int, c_void_p,
                                 c_void_p]),
}

_SIGNATURES.update({
    "nmfb_peer_push": (c_int, [c_void_p, c_long, c_void_p, c_int, c_long, c_long, c_int, c_long, c_void_p]),
    "nmfb_peer_scatter": (c_int, [c_void_p, c_long, c_long, c_void_p, c_int, c_long, c_long, c_int, c_void_p, c_void_p]),
    "nmfb_peer_push_mc": (c_int, [c_void_p, c_long, c_void_p, c_long, c_long, c_int, c_long, c_void_p]),
    "nmfb_peer_barrier": (c_int, [c_void_p, c_int, c_int, C.c_uint, c_void_p]),
    "nmfb_sum_rank_slots": (c_int, [c_void_p, c_int, c_long, c_int, c_void_p, c_void_p, c_void_p]),
    "nmfb_adopt_exchange": (c_int, [c_int, P(Factor), c_void_p, c_int, c_void_p]),
    "nmfb_gram_window": (c_int, [P(Factor), c_int, c_int, c_void_p, c_size_t, c_void_p, c_void_p, c_int, c_void_p]),
    "nmfb_exchange_factor": (c_int, [P(PeerCtx), c_int, P(Factor), c_long, c_int, c_int, c_int, c_long, c_void_p,
                                     c_void_p, C.c_uint, c_void_p, c_void_p]),
    "nmfb_exchange_scalar": (c_int, [P(PeerCtx), c_void_p, c_long, C.c_uint, c_void_p, c_void_p]),
    "nmfb_mur_eu_window_step": (c_int, [P(Problem), c_int, c_float, c_void_p]),
    "nmfb_row_planes3": (c_int, [c_int, P(Factor), c_void_p, c_void_p]),
    "nmfb_kl_data_step": (c_int, [P(Data), c_void_p, c_int, c_int, c_void_p, c_int, c_int, c_void_p, c_void_p, c_long,
                                  c_void_p, c_void_p, c_void_p, c_int, c_void_p]),
    "nmfb_upload_2d": (c_int, [c_void_p, c_long, c_void_p, c_long, c_long, c_long, c_void_p]),
})

_OPTIONAL = set()

_lib = None


def exported_symbols():
    """Names this binding expects libnmf_b200.so to export (checked by the CPU tests)."""
    return sorted(_SIGNATURES)


def load(build_if_needed: bool = True):
    """Load (building first if the in-tree .so is stale and nvcc is present) the C-ABI library."""
    global _lib
    if _lib is not None:
        return _lib
    path = _build.LIB_PATH
    if build_if_needed and _build.is_stale():
        try:
            _build.build()
        except _build.NvccMissing as exc:       # no compiler on the box: the prebuilt .so stands in, loudly
            if not os.path.exists(path):
                raise RuntimeError(f"libnmf_b200.so is missing and could not be built: {exc}") from exc
            import warnings
            warnings.warn("nmf_b200: sources are newer than libnmf_b200.so and there is no nvcc to rebuild it; "
                          "using the prebuilt library", RuntimeWarning)
        # compile and link errors propagate: a stale binary must never stand in for an edit that does not build
    if not os.path.exists(path):
        raise RuntimeError("libnmf_b200.so not found; run `python -c 'import __graft_entry__ as g; g.build()'`")
    lib = C.CDLL(path)
    for name, (res, args) in _SIGNATURES.items():
        try:
            fn = getattr(lib, name)
        except AttributeError:
            if name in _OPTIONAL:
                continue
            raise
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib


def check(rc: int) -> int:
    """Raise on a negative status, pass non-negative results through."""
    if rc < 0:
        msg = load().nmfb_last_error().decode("utf-8", "replace")
        raise NmfbError(rc, msg)
    return rc
