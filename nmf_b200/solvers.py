"""Steppers: one object per solver that enqueues ONE outer iteration on an Engine.

They are what the drivers (mur.py / admm.py / ao_admm.py) and bench.py share, so that the
benchmark times exactly the call sequence the public API runs.
"""
from __future__ import annotations


class MurEuStepper:
    """mur.py:122-127 with distance_type='eu' -> nmfb_mur_eu_step."""
    phases = ["gram_h", "gemm_w", "update_w", "gram_w", "gemm_h", "update_h", "objective"]

    def __init__(self, eng, lambda_w=0.0, lambda_h=0.0):
        self.eng, self.lw, self.lh = eng, float(lambda_w), float(lambda_h)

    def prepare(self):
        e = self.eng
        if e.world > 1 and e.peer is not None:
            e.full_grams()                               # from here on the exchanges keep both Grams current
        e.objective_eu(slot=0)
        e.reduce_scalar(0)

    def step(self, ring_slot: int):
        e = self.eng
        if e.world == 1:
            e.mur_eu_step(self.lw, self.lh, obj_slot=ring_slot * 8)
            return
        if e.peer is None:
            # host stand-in of the CPU tests: update my window of W, all-gather W, my window of H, all-gather H
            e.mur_eu_half_step(0, self.lw)
            e.gather_factor(e.w)
            e.mur_eu_half_step(1, self.lh)
            e.gather_factor(e.h)
        else:
            # GPUs: my window of W (GEMM on my rows of V + update), then ONE exchange that stores the window and
            # my term of W^T W into the peers' memory (k x k Gram all-reduce, north_star); the same for H
            e.mur_eu_window_step(0, self.lw)
            e.gather_factor(e.w, gram_index=0)
            e.mur_eu_window_step(1, self.lh)
            e.gather_factor(e.h, gram_index=1)
        e.objective_eu(slot=ring_slot * 8)
        e.reduce_scalar(ring_slot * 8)
        if e.peer is not None:
            e._mark(7)

    def phase_names(self):
        return self.phases

    def phase_pairs(self):
        """(name, first event, last event) triples that are recorded in this mode."""
        if self.eng.world == 1:
            return [(nm, j, j + 1) for j, nm in enumerate(self.phases)]
        # update_*: the wait for the Gram terms of the side streams (cross-GPU barrier + sum) and the fused update;
        # exchange_w: window push, barrier, adoption of all windows; the last interval also holds the objective
        if self.eng.peer is None:
            return [("gemm_w", 1, 2), ("update_w", 2, 3), ("gemm_h", 4, 5), ("update_h", 5, 6)]
        return [("gemm_w", 1, 2), ("update_w", 2, 3), ("exchange_w", 3, 4), ("gemm_h", 4, 5), ("update_h", 5, 6),
                ("exchange_h_objective", 6, 7)]

    def roofline(self, phase_ms, peaks):
        """Dominant kernel = the split-bf16 panel GEMM (two launches per iteration: V.H^T and
        W^T.V).  Algorithmic work of one launch: 2*m*n*k FLOP over m*n*2 bytes of V."""
        e = self.eng
        ms = 0.5 * (phase_ms["gemm_w"] + phase_ms["gemm_h"])
        mpp = getattr(self, "mma_per_product", 2)
        if mpp == 3 and getattr(e, "fixed", False):
            mpp = 5                                          # exact-hi contraction (panel_gemm_exact.cu)
        flops = 2.0 * e.m * e.n * e.k / e.world              # per rank (its shard of V)
        vbytes = 2.0 * e.m * e.n * e.nv / e.world
        tf = flops / (ms * 1e-3) / 1e12
        gbs = vbytes / (ms * 1e-3) / 1e9
        peak = peaks["bf16_tflops_sustained"]
        tensor_frac, hbm_frac = tf / peak, gbs / peaks["hbm_gbs"]
        # the binding roof is the one whose lower bound on the ALGORITHMIC work is larger (with
        # 2-byte V read once these shapes sit on the ridge or on the HBM side); the executed-MMA fraction
        # (mpp MMAs per algorithmic product) is reported next to it
        bound = "tensor" if tensor_frac >= hbm_frac else "hbm"
        out = {"kernel": "panel_gemm_exact_kernel" if mpp == 5 else "panel_gemm_kernel", "bound": bound, "ms_per_launch": ms,
               "executed_mma_per_product": mpp, "traffic": None,
               "tensor": {"achieved": tf, "peak": peak, "unit": "TFLOP/s", "frac": tensor_frac,
                          "executed_frac": mpp * tensor_frac},
               "hbm": {"achieved": gbs, "peak": peaks["hbm_gbs"], "unit": "GB/s", "frac": hbm_frac}}
        pick = out["tensor"] if bound == "tensor" else out["hbm"]
        out.update(achieved=pick["achieved"], peak=pick["peak"], unit=pick["unit"], frac=pick["frac"],
                   peak_source=peaks["source"] + (" (sustained bf16)" if bound == "tensor" else ""))
        return out


class MurKlStepper:
    """mur.py:122-127 with distance_type='kl' -> nmfb_mur_kl_step."""
    phases = ["kl_w", "kl_h"]

    def __init__(self, eng, lambda_w=0.0, lambda_h=0.0):
        self.eng, self.lw, self.lh = eng, float(lambda_w), float(lambda_h)

    def prepare(self):
        self.eng.objective_kl(slot=0)
        self.eng.reduce_scalar(0)

    lagged = True       # _loop.run: the objective of iteration i-1 comes out of the W pass of iteration i

    def step(self, ring_slot: int, prev_slot: int, mid=None):
        """One outer iteration (mur.py:122-123).  The W pass also produces the objective (mur.py:127) of
        the pair it starts from -- the previous iteration's -- into scalars[prev_slot*8]; `mid` is called
        right after it so that the host can read that value while the H pass runs."""
        e = self.eng
        e.mur_kl_half_step(0, self.lw, obj_start_slot=prev_slot * 8)   # my rows of W (all of them unsharded)
        e.reduce_scalar(prev_slot * 8)
        if mid is not None:
            mid()
        e.gather_factor(e.w)
        e.mur_kl_half_step(1, self.lh)                                 # my columns of H with the new W
        e.gather_factor(e.h)

    def finish(self, ring_slot: int):
        """Stand-alone objective of the current pair (after the last iteration)."""
        self.eng.objective_kl(slot=ring_slot * 8)
        self.eng.reduce_scalar(ring_slot * 8)

    def phase_names(self):
        return self.phases

    def phase_pairs(self):
        return [("kl_w", 0, 1), ("kl_h", 2, 3)]

    def roofline(self, phase_ms, peaks):
        """Dominant kernel = the fused GEMM -> quotient -> GEMM kernel (one launch per side):
        algorithmic 4*m*n*k FLOP per launch (S tile + Q.Y^T), executed 3 MMAs per product."""
        e = self.eng
        if "kl_w" not in phase_ms:
            return None
        ms = 0.5 * (phase_ms["kl_w"] + phase_ms["kl_h"])     # W pass carries the objective terms as well
        tf = 4.0 * e.m * e.n * e.k / e.world / (ms * 1e-3) / 1e12      # per rank (its shard of V)
        gbs = 2.0 * e.m * e.n * e.nv / e.world / (ms * 1e-3) / 1e9
        peak = peaks["bf16_tflops_sustained"]
        return {"kernel": "kl_panel_ts2_kernel" if e.KP == 128 else "kl_panel_ts_kernel", "bound": "tensor", "ms_per_launch": ms, "executed_mma_per_product": 3,
                "traffic": None, "achieved": tf, "peak": peak, "unit": "TFLOP/s", "frac": tf / peak,
                "executed_frac": 3 * tf / peak,
                "hbm": {"achieved": gbs, "peak": peaks["hbm_gbs"], "unit": "GB/s", "frac": gbs / peaks["hbm_gbs"]},
                "peak_source": peaks["source"] + " (sustained bf16)"}


PROX_CODES = {"nn": 0, "l1n": 1, "l2n": 2}


def prox_code(reg, k=None):
    """(lambda, name) -> (lambda, C-ABI prox code).  'l2n' with lambda == 0 is the identity system of
    admm.py:141-156, i.e. plain non-negativity; with lambda > 0 it is the Tikhonov solve along the
    component axis (k <= 128 in this build).  'l1inf' / 'l1inf_transpose' are not part of this build."""
    lam, name = reg
    if name == "l2n" and lam == 0:
        name = "nn"
    if name == "l2n" and k is not None and k > 128:
        raise NotImplementedError("prox operator 'l2n' with lambda > 0 supports k <= 128 in the B200 build")
    if name in PROX_CODES:
        return float(lam), PROX_CODES[name]
    if name in ("l1inf", "l1inf_transpose"):
        raise NotImplementedError(f"prox operator {name!r} is not provided by the B200 build "
                                  "(available: 'nn', 'l1n', 'l2n')")
    raise TypeError('Unknown prox_type.')


class _GemmRoofline:
    """Shared roofline report: the dominant kernel is the panel GEMM (two launches per iteration)."""
    mma_per_product = 2
    gemm_phases = ("gemm_w", "gemm_h")

    def roofline(self, phase_ms, peaks):
        return MurEuStepper.roofline(self, phase_ms, peaks)


class AdmmStepper(_GemmRoofline):
    """admm.py:292-334 (Euclidean) -> nmfb_admm_step."""
    phases = ["gram_chol_h", "gemm_h", "solve_prox_dual_h", "gram_chol_w", "gemm_w", "solve_prox_dual_w", "objective"]
    mma_per_product = 3

    def __init__(self, eng, rho=1, reg_w=(0, "nn"), reg_h=(0, "l2n")):
        self.eng, self.rho = eng, float(rho)
        self.lw, self.pw = prox_code(reg_w, eng.k)
        self.lh, self.ph = prox_code(reg_h, eng.k)

    def prepare(self):
        self.eng.alloc_admm(with_aux=True)
        self.eng.objective_eu(slot=0)
        self.eng.reduce_scalar(0)

    def step(self, ring_slot):
        e = self.eng
        if e.world == 1:
            e.admm_step(self.rho, self.lw, self.pw, self.lh, self.ph, ring_slot)
            return
        # sharded: my columns of h_aux / H / U_h, exchange, my rows of w_aux / W / U_w, exchange
        e.admm_half_step(0, self.rho, self.lh, self.ph)
        e.gather_factor(e.h_aux, refresh=not e.fixed, fixed=e.fixed)     # operand of the W-side Gram and GEMM
        e.gather_factor(e.h)                                            # operand of the objective
        e.admm_half_step(1, self.rho, self.lw, self.pw)
        e.gather_factor(e.w_aux, refresh=not e.fixed, fixed=e.fixed)
        e.gather_factor(e.w, refresh=False)                             # keeps the replicated W current
        e.objective_eu(slot=ring_slot * 8)
        e.reduce_scalar(ring_slot * 8)
        e.admm_publish(ring_slot)

    def phase_pairs(self):
        if self.eng.world == 1:
            return [(nm, j, j + 1) for j, nm in enumerate(self.phases)]
        return [("gemm_h", 1, 2), ("gemm_w", 4, 5)]

    def phase_names(self):
        return self.phases


class AdmmKlStepper:
    """admm.py:292-334 with distance_type='kl' (:303-315): the half-steps of the Euclidean branch on the data
    term v_aux + dual_v, then the element-wise v_aux / dual_v step on the rank's rows of V."""
    phases = []

    def __init__(self, eng, rho=1, reg_w=(0, "nn"), reg_h=(0, "l2n")):
        self.eng, self.rho = eng, float(rho)
        self.lw, self.pw = prox_code(reg_w, eng.k)
        self.lh, self.ph = prox_code(reg_h, eng.k)

    def prepare(self):
        self.eng.alloc_admm_kl()
        self.eng.objective_kl(slot=0)
        self.eng.reduce_scalar(0)

    def step(self, ring_slot):
        self.eng.admm_kl_step(self.rho, self.lw, self.pw, self.lh, self.ph, ring_slot)

    def phase_names(self):
        return self.phases

    def phase_pairs(self):
        return []

    def roofline(self, phase_ms, peaks):
        return None


class AoAdmmStepper(_GemmRoofline):
    """ao_admm.py:259-301 (Euclidean) -> nmfb_aoadmm_step."""
    phases = ["gram_chol_h", "gemm_h", "inner_h", "gram_chol_w", "gemm_w", "inner_w", "objective"]
    mma_per_product = 3

    def __init__(self, eng, reg_w=(0, "nn"), reg_h=(0, "l2n"), admm_iter=10):
        self.eng, self.admm_iter = eng, int(admm_iter)
        self.lw, self.pw = prox_code(reg_w, eng.k)
        self.lh, self.ph = prox_code(reg_h, eng.k)

    def prepare(self):
        self.eng.alloc_admm(with_aux=False)
        self.eng.objective_eu(slot=0)
        self.eng.reduce_scalar(0)

    def step(self, ring_slot):
        e = self.eng
        if e.world == 1:
            e.aoadmm_step(self.admm_iter, self.lw, self.pw, self.lh, self.ph, ring_slot)
            return
        e.aoadmm_side(0, self.admm_iter, self.lh, self.ph)              # H update, ao_admm.py:261-264
        e.aoadmm_side(1, self.admm_iter, self.lw, self.pw)              # W update, ao_admm.py:265-270
        e.objective_eu(slot=ring_slot * 8)
        e.reduce_scalar(ring_slot * 8)
        e.admm_publish(ring_slot)

    def phase_pairs(self):
        if self.eng.world == 1:
            return [(nm, j, j + 1) for j, nm in enumerate(self.phases)]
        return [("gemm_h", 1, 2), ("gemm_w", 4, 5)]

    def phase_names(self):
        return self.phases


class AoAdmmKlStepper(AdmmKlStepper):
    """ao_admm.py:259-301 with distance_type='kl' (:272-285, admm_kl_update :71-101).  The inner loop is
    enqueued from the host, its early break is taken on the device (flag-guarded kernels)."""

    def __init__(self, eng, reg_w=(0, "nn"), reg_h=(0, "l2n"), admm_iter=10):
        self.eng, self.admm_iter = eng, int(admm_iter)
        self.lw, self.pw = prox_code(reg_w, eng.k)
        self.lh, self.ph = prox_code(reg_h, eng.k)

    def prepare(self):
        self.eng.alloc_admm_kl(ao=True)
        self.eng.objective_kl(slot=0)
        self.eng.reduce_scalar(0)

    def step(self, ring_slot):
        self.eng.aoadmm_kl_step(self.admm_iter, self.lw, self.pw, self.lh, self.ph, ring_slot)


def make_stepper(eng, method, distance_type="eu", **kw):
    if method == "mur" and distance_type == "eu":
        return MurEuStepper(eng, kw.get("lambda_w", 0.0), kw.get("lambda_h", 0.0))
    if method == "mur" and distance_type == "kl":
        return MurKlStepper(eng, kw.get("lambda_w", 0.0), kw.get("lambda_h", 0.0))
    if method == "admm" and distance_type == "eu":
        return AdmmStepper(eng, kw.get("rho", 1), kw.get("reg_w", (0, "nn")), kw.get("reg_h", (0, "l2n")))
    if method == "admm" and distance_type == "kl":
        return AdmmKlStepper(eng, kw.get("rho", 1), kw.get("reg_w", (0, "nn")), kw.get("reg_h", (0, "l2n")))
    if method == "ao_admm" and distance_type == "eu":
        return AoAdmmStepper(eng, kw.get("reg_w", (0, "nn")), kw.get("reg_h", (0, "l2n")), kw.get("admm_iter", 10))
    if method == "ao_admm" and distance_type == "kl":
        return AoAdmmKlStepper(eng, kw.get("reg_w", (0, "nn")), kw.get("reg_h", (0, "l2n")), kw.get("admm_iter", 10))
    raise NotImplementedError(f"no stepper for method={method!r} distance_type={distance_type!r}")
