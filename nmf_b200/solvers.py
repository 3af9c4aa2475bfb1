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
        self.eng.objective_eu(slot=0)

    def step(self, ring_slot: int):
        self.eng.mur_eu_step(self.lw, self.lh, obj_slot=ring_slot * 8)

    def phase_names(self):
        return self.phases

    def roofline(self, phase_ms, peaks):
        """Dominant kernel = the split-bf16 panel GEMM (two launches per iteration: V.H^T and
        W^T.V).  Algorithmic work of one launch: 2*m*n*k FLOP over m*n*2 bytes of V."""
        e = self.eng
        ms = 0.5 * (phase_ms["gemm_w"] + phase_ms["gemm_h"])
        flops = 2.0 * e.m * e.n * e.k
        vbytes = 2.0 * e.m * e.n * e.nv
        tf = flops / (ms * 1e-3) / 1e12
        gbs = vbytes / (ms * 1e-3) / 1e9
        peak = peaks["bf16_tflops_sustained"]
        tensor_frac, hbm_frac = tf / peak, gbs / peaks["hbm_gbs"]
        bound = "tensor" if tensor_frac * 2 >= hbm_frac else "hbm"     # 2 MMAs execute per product
        out = {"kernel": "panel_gemm_kernel", "bound": bound, "ms_per_launch": ms,
               "executed_mma_per_product": 2, "traffic": None,
               "tensor": {"achieved": tf, "peak": peak, "unit": "TFLOP/s", "frac": tensor_frac,
                          "executed_frac": 2 * tensor_frac},
               "hbm": {"achieved": gbs, "peak": peaks["hbm_gbs"], "unit": "GB/s", "frac": hbm_frac}}
        pick = out["tensor"] if bound == "tensor" else out["hbm"]
        out.update(achieved=pick["achieved"], peak=pick["peak"], unit=pick["unit"], frac=pick["frac"],
                   peak_source=peaks["source"] + (" (sustained bf16)" if bound == "tensor" else ""))
        return out


def make_stepper(eng, method, distance_type="eu", **kw):
    if method == "mur" and distance_type == "eu":
        return MurEuStepper(eng, kw.get("lambda_w", 0.0), kw.get("lambda_h", 0.0))
    raise NotImplementedError(f"no stepper for method={method!r} distance_type={distance_type!r}")
