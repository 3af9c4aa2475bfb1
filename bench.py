#!/usr/bin/env python3
"""Headline benchmark: NMF outer iterations / second on B200 (BASELINE.json `metric`).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--workload NAME] [--impl reference]

A "step" is one outer iteration of the solver (both factor updates + the objective, exactly the
body of the reference loop mur.py:119-136 / admm.py:292-334 / ao_admm.py:259-301) on synthetic
planted non-negative data.  Prints ONE JSON line (rank 0).

  value        iterations/s with V resident in HBM (CUDA events, max over ranks)
  e2e          iterations/s through NMF(V_host, k).factorize(...): upload of the host V, K
               iterations, download of W and H -- all inside the timed region (N > 1: the same call
               with comm=True on every rank, barrier to slowest rank)
  roofline     dominant kernel (the split-bf16 tcgen05 panel GEMM) timed live with CUDA events
  time_to_tol  seconds until the reference's stop test fires (tol2 = 1e-5 x initial objective), end to end
  cpu_baseline the reference's NMF.factorize (oracle/_ref; else the oracle port) timed on this box's host cores

`--impl reference` times the reference's own NMF.factorize (vendored into oracle/_ref by
__graft_entry__.build(); the oracle port when that is absent) on the host cores, on a bounded block of the
same workload, scaled to the full size.
"""
from __future__ import annotations

import argparse
import json
import os
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

# name -> (method, distance, m, n, k, extra kwargs)
WORKLOADS = {
    "mur_eu_65536x65536_k128": ("mur", "eu", 65536, 65536, 128, {}),
    "mur_eu_32768x32768_k64": ("mur", "eu", 32768, 32768, 64, {}),
    "mur_kl_20000x50000_k128": ("mur", "kl", 20000, 50000, 128, {}),
    "admm_65536x16384_k128": ("admm", "eu", 65536, 16384, 128,
                              dict(rho=100, reg_w=(0, "nn"), reg_h=(0.1, "l1n"))),
    "ao_admm_65536x16384_k128": ("ao_admm", "eu", 65536, 16384, 128,
                                 dict(reg_w=(0, "nn"), reg_h=(0.1, "l1n"), admm_iter=5)),
    "mur_eu_262144x131072_k256": ("mur", "eu", 262144, 131072, 256, {}),
    # Kullback-Leibler branch of admm / ao_admm (single GPU, first version; no roofline claim yet)
    "admm_kl_65536x16384_k128": ("admm", "kl", 65536, 16384, 128,
                                 dict(rho=100, reg_w=(0, "nn"), reg_h=(0.1, "l1n"))),
    "ao_admm_kl_65536x16384_k128": ("ao_admm", "kl", 65536, 16384, 128,
                                    dict(reg_w=(0, "nn"), reg_h=(0.1, "l1n"), admm_iter=5)),
    "mur_eu_4096x4096_k64": ("mur", "eu", 4096, 4096, 64, {}),          # quick functional check
    "mur_eu_65536x32768_k256": ("mur", "eu", 65536, 32768, 256, {}),    # 1/16 of config 5: the KP = 256 kernels
}
DEFAULT_WORKLOAD = "mur_eu_65536x65536_k128"

# dram__bytes_read.sum + dram__bytes_write.sum of ONE launch of the dominant kernel, from an `ncu --set full` capture
# of the same kernel at the same shape (per launch, like `achieved`), with the capture it comes from.
NCU_TRAFFIC = {
    ("panel_gemm_kernel", "mur_eu_65536x65536_k128"):
        (8.624068e9 + 85.743104e6, "profiles/r01b_all_paths.md (kernel unchanged since)"),
    ("kl_panel_ts2_kernel", "mur_kl_20000x50000_k128"):
        (2.098872e9 + 65.014016e6, "profiles/r02z_ncu_full_summary.md"),
    ("panel_gemm_exact_kernel", "admm_65536x16384_k128"):
        (4.393657e9 + 188.351488e6, "profiles/r01b_all_paths.md (kernel unchanged since)"),
    ("panel_gemm_exact_kernel", "ao_admm_65536x16384_k128"):
        (4.393657e9 + 188.351488e6, "profiles/r01b_all_paths.md (kernel unchanged since)"),
}


def load_peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        with open(path) as f:
            p = json.load(f)
        return dict(hbm_gbs=p["hbm_gbs"], bf16_tflops=p["bf16_tflops"],
                    bf16_tflops_sustained=p.get("bf16_tflops_sustained", p["bf16_tflops"]),
                    source="measured")
    return dict(hbm_gbs=6650.0, bf16_tflops=1590.0, bf16_tflops_sustained=1400.0, source="fallback")


class ClockSampler:
    """Samples SM clock + throttle reasons through NVML while the timed region runs."""

    BAD = {"hw_slowdown": 0x8, "hw_thermal_slowdown": 0x40, "sw_thermal_slowdown": 0x20}
    NOTE = {"sw_power_cap": 0x4, "hw_power_brake": 0x80}

    def __init__(self, index: int):
        self.samples, self.reasons, self.max_mhz = [], set(), None
        self._stop = threading.Event()
        self._thread = None
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = int(pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM))
        except Exception:                                   # NVML missing: report nulls
            self.nv = None

    def _run(self):
        nv = self.nv
        while not self._stop.is_set():
            try:
                self.samples.append(int(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM)))
                mask = int(nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h))
                for name, bit in {**self.BAD, **self.NOTE}.items():
                    if mask & bit:
                        self.reasons.add(name)
            except Exception:
                pass
            time.sleep(0.01)

    def __enter__(self):
        if self.nv is not None:
            self._thread = threading.Thread(target=self._run, daemon=True)
            self._thread.start()
        return self

    def __exit__(self, *exc):
        self._stop.set()
        if self._thread is not None:
            self._thread.join()

    def summary(self):
        med = int(np.median(self.samples)) if self.samples else None
        return {"sm_mhz": med, "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons),
                "samples": len(self.samples)}


# ----------------------------------------------------------------------------------------------
# synthetic workload on the device: planted low-rank + 1% noise on the bf16 grid
# ----------------------------------------------------------------------------------------------
def synth_blocks(dev, m: int, n: int, k: int, seed: int, chunk: int = 2048):
    """Returns (wt, ht, blocks): blocks() yields (r0, rows r0.. of V as fp32 on the bf16 grid).  Every chunk
    is drawn from its own counter (seed, r0), so any rank -- and the host rebuild of the end-to-end leg --
    regenerates the same V chunk by chunk."""
    import torch
    g = torch.Generator(device=dev)
    g.manual_seed(seed)
    wt = torch.rand((m, k), device=dev, generator=g)
    ht = torch.rand((k, n), device=dev, generator=g)

    def blocks():
        for r0 in range(0, m, chunk):
            r1 = min(m, r0 + chunk)
            g.manual_seed(seed * 1000003 + 17 + r0)
            blk = wt[r0:r1] @ ht
            blk.add_(torch.rand((r1 - r0, n), device=dev, generator=g), alpha=0.01)
            yield r0, blk.bfloat16().float()               # values on the bf16 grid, fp32 container

    return wt, ht, blocks


def synth_fill_engine(eng, seed: int, fixed: bool = False):
    """Fill eng.v_rows / eng.v_cols (1 lossless bf16 plane) for the engine's row / column shard.
    Every rank walks the same row chunks with the same per-chunk generator, so shards of
    different ranks are slices of one and the same V."""
    wt, ht, blocks = synth_blocks(eng.device, eng.m, eng.n, eng.k, seed)
    # through the library's ingest kernels; a sharded engine keeps its rows / columns of every chunk
    eng.begin_ingest(1, fixed=fixed)
    if fixed:
        for r0, blk in blocks():
            eng.absmax_chunk(blk, r0)
    for r0, blk in blocks():
        eng.ingest_chunk(blk, r0)
    vmin, inexact, bad = eng.finish_ingest()
    assert not bad and not inexact
    return wt, ht


def synth_init(eng, seed: int):
    """|randn| initial factors generated on the device (reference order W then H)."""
    import ctypes as C
    import torch
    from nmf_b200._lib import check
    g = torch.Generator(device=eng.device)
    g.manual_seed(seed)
    eng.w.master.zero_()
    eng.h.master.zero_()
    eng.w.master[:eng.k, :eng.m] = torch.randn((eng.k, eng.m), device=eng.device, generator=g).abs_()
    eng.h.master[:eng.k, :eng.n] = torch.randn((eng.k, eng.n), device=eng.device, generator=g).abs_()
    for f in (eng.w, eng.h):
        fs = f.struct()
        check(eng.lib.nmfb_refresh_planes(eng.KP, C.byref(fs), eng.stream))


# ----------------------------------------------------------------------------------------------
# CPU arm: the oracle port of the reference loop on a bounded sample of the workload
# ----------------------------------------------------------------------------------------------
def cpu_sample_shape(m, n, k, budget_elems):
    """Largest (m/2^a, n/2^b) block with at most budget_elems elements (halve the longer side)."""
    ms, ns = m, n
    while ms * ns > budget_elems:
        if ms >= ns:
            ms //= 2
        else:
            ns //= 2
    return ms, ns


def cpu_step_fn(method, dist, ms, ns, k, extra, seed=0):
    """Returns (step, cores): `step()` runs ONE outer iteration of the oracle on the sample."""
    from oracle import nmf_oracle as orc
    v = orc.planted(ms, ns, k, seed).astype(np.float64)
    w0, h0 = orc.seeded_init(ms, ns, k, seed + 1)
    st = {"w": w0, "h": h0}
    if method == "mur":                          # loop body of mur.py:122-127
        st["wh"] = w0 @ h0

        def step():
            w = orc.mur_w_step(dist, v, st["w"], st["h"], st["wh"], 0.0)
            h = orc.mur_h_step(dist, v, w, st["h"], w @ st["h"], 0.0)
            st["wh"] = w @ h
            st["w"], st["h"] = w, h
            return orc.objective(v, st["wh"], dist)
    elif method == "admm" and dist == "kl":      # loop body of admm.py:303-324
        rho, reg_w, reg_h = extra["rho"], extra["reg_w"], extra["reg_h"]
        st.update(w_aux=w0.copy(), h_aux=h0.copy(), dw=np.zeros_like(w0), dh=np.zeros_like(h0),
                  va=np.zeros_like(v), dv=np.zeros_like(v))

        def step():
            h_aux = orc.admm_aux_kl(st["h"], st["dh"], st["w_aux"], st["va"], st["dv"], rho)
            w_aux = orc.admm_aux_kl(st["w"].T, st["dw"].T, h_aux.T, st["va"].T, st["dv"].T, rho).T
            h = orc.prox(reg_h[1], h_aux, st["dh"], rho, reg_h[0])
            w = orc.prox(reg_w[1], w_aux.T, st["dw"].T, rho, reg_w[0]).T
            wh_aux = w_aux @ h_aux
            va = orc.kl_v_aux(v, wh_aux - st["dv"])
            st["dv"] = st["dv"] + va - wh_aux
            st["dh"] = st["dh"] + h - h_aux
            st["dw"] = st["dw"] + w - w_aux
            st.update(w=w, h=h, w_aux=w_aux, h_aux=h_aux, va=va)
            return orc.objective(v, w @ h, "kl")
    elif method == "ao_admm" and dist == "kl":   # loop body of ao_admm.py:272-291
        reg_w, reg_h, inner = extra["reg_w"], extra["reg_h"], extra["admm_iter"]
        st.update(dw=np.zeros_like(w0), dh=np.zeros_like(h0), va=np.zeros_like(v), dv=np.zeros_like(v))

        def step():
            h, dh, va, dv, _ = orc.ao_factor_update_kl(v, st["va"], st["dv"], st["w"], st["h"], st["dh"], k,
                                                       reg_h[1], reg_h[0], inner)
            wt, dwt, vat, dvt, _ = orc.ao_factor_update_kl(v.T, va.T, dv.T, h.T, st["w"].T, st["dw"].T, k,
                                                           reg_w[1], reg_w[0], inner)
            st.update(w=wt.T, h=h, dw=dwt.T, dh=dh, va=vat.T, dv=dvt.T)
            return orc.objective(v, st["w"] @ h, "kl")
    elif method == "admm":                       # loop body of admm.py:294-301,320-324
        rho, reg_w, reg_h = extra["rho"], extra["reg_w"], extra["reg_h"]
        st.update(w_aux=w0.copy(), h_aux=h0.copy(), dw=np.zeros_like(w0), dh=np.zeros_like(h0))

        def step():
            h_aux = orc.admm_aux(st["h"], st["dh"], st["w_aux"], v, rho)
            w_aux = orc.admm_aux(st["w"].T, st["dw"].T, h_aux.T, v.T, rho).T
            h = orc.prox(reg_h[1], h_aux, st["dh"], rho, reg_h[0])
            w = orc.prox(reg_w[1], w_aux, st["dw"], rho, reg_w[0])
            st["dh"] = st["dh"] + h - h_aux
            st["dw"] = st["dw"] + w - w_aux
            st.update(w=w, h=h, w_aux=w_aux, h_aux=h_aux)
            return orc.objective(v, w @ h)
    else:                                        # loop body of ao_admm.py:261-270,291
        reg_w, reg_h, inner = extra["reg_w"], extra["reg_h"], extra["admm_iter"]
        st.update(dw=np.zeros_like(w0), dh=np.zeros_like(h0))

        def step():
            h, dh, _ = orc.ao_factor_update(v, st["w"], st["h"], st["dh"], k, reg_h[1], reg_h[0], inner)
            wt, dwt, _ = orc.ao_factor_update(v.T, h.T, st["w"].T, st["dw"].T, k, reg_w[1], reg_w[0], inner)
            st.update(w=wt.T, h=h, dw=dwt.T, dh=dh)
            return orc.objective(v, st["w"] @ h)
    try:
        from threadpoolctl import threadpool_info
        cores = max([p.get("num_threads", 1) for p in threadpool_info()] or [1])
    except Exception:
        cores = os.cpu_count() or 1
    return step, cores


def use_all_host_threads():
    """torch.distributed.run exports OMP_NUM_THREADS=1; the CPU arm is supposed to use the box's cores."""
    n = os.cpu_count() or 1
    try:
        from threadpoolctl import threadpool_limits
        threadpool_limits(limits=n)
    except Exception:                                        # noqa: BLE001
        pass
    return n


def blas_threads():
    try:
        from threadpoolctl import threadpool_info
        return max([p.get("num_threads", 1) for p in threadpool_info()] or [1])
    except Exception:                                        # noqa: BLE001
        return os.cpu_count() or 1


def load_reference():
    """The UNMODIFIED reference package, vendored by __graft_entry__.build() from /root/reference/nmf into
    oracle/_ref/nmf (git-ignored, travels to the GPU box), imported under a private name so that it cannot be
    mistaken for the drop-in shim `nmf/`.  None when it is not there."""
    import importlib.util
    pkg = os.path.join(ROOT, "oracle", "_ref", "nmf")
    init = os.path.join(pkg, "__init__.py")
    if not os.path.exists(init):
        return None
    name = "_nmf_reference"
    if name in sys.modules:
        return sys.modules[name]
    spec = importlib.util.spec_from_file_location(name, init, submodule_search_locations=[pkg])
    mod = importlib.util.module_from_spec(spec)
    sys.modules[name] = mod
    spec.loader.exec_module(mod)
    return mod


def reference_run_seconds(ref, v, k, method, dist, extra, max_iter):
    """Wall time of NMF(V, k).factorize(method, max_iter=...) of the reference: stop test off, stdout dropped."""
    import contextlib
    kw = dict(extra, distance_type=dist, min_iter=10**9, max_iter=max_iter)
    if method != "mur":
        kw["nndsvd_init"] = (False, "zero")
    np.random.seed(1)
    model = ref.NMF(v, k)
    t0 = time.perf_counter()
    with open(os.devnull, "w") as sink, contextlib.redirect_stdout(sink):
        model.factorize(method=method, **kw)
    return time.perf_counter() - t0


def time_reference(ref, method, dist, m, n, k, extra, steps, warmup, budget_s):
    """The reference's own NMF.factorize on a block of the workload, by the two-run difference of BASELINE.md
    section 3: t(max_iter = warmup + steps) - t(max_iter = warmup) excludes the initialisation and the first
    `warmup` iterations.  The block is the largest power-of-two fraction of V whose two runs fit budget_s (the
    reference needs ~8 fp64 copies of V; 65536^2 does not fit this host at all); work is proportional to m*n, so
    the full-size rate is the block's rate divided by the area ratio -- the line says so."""
    from oracle import nmf_oracle as orc                     # data generator only
    warmup = max(1, warmup)
    pm, pn = cpu_sample_shape(m, n, k, 1 << 20)
    pv = orc.planted(pm, pn, k, 0).astype(np.float64)
    reference_run_seconds(ref, pv, k, method, dist, extra, 1)
    per_elem = (reference_run_seconds(ref, pv, k, method, dist, extra, 3)
                - reference_run_seconds(ref, pv, k, method, dist, extra, 1)) / 2 / (pm * pn)
    total_iters = 2 * warmup + steps + 2                     # both runs + roughly one iteration per init
    budget_elems = max(1 << 22, int(budget_s / max(per_elem, 1e-12) / total_iters))
    budget_elems = min(budget_elems, 1 << 28)                # <= 2 GiB of fp64 V (x ~8 temporaries)
    ms, ns = cpu_sample_shape(m, n, k, budget_elems)
    v = orc.planted(ms, ns, k, 0).astype(np.float64)
    t_a = reference_run_seconds(ref, v, k, method, dist, extra, warmup)
    t_b = reference_run_seconds(ref, v, k, method, dist, extra, warmup + steps)
    dt = max(t_b - t_a, 1e-9) / steps
    scale = (m * n) / (ms * ns)
    desc = (f"reference NMF.factorize (oracle/_ref) on a {ms}x{ns} block of the {m}x{n} workload (1/{scale:g} of V), "
            f"k={k}: t(max_iter={warmup + steps}) - t(max_iter={warmup}) = {t_b - t_a:.2f} s over {steps} iterations "
            f"= {dt * 1e3:.1f} ms each; SCALED by {scale:g} to the full size (work is proportional to m*n)")
    return 1.0 / (dt * scale), dt * 1e3, blas_threads(), desc


def time_cpu(method, dist, m, n, k, extra, steps, warmup, budget_s):
    """-> (full-size iterations/s, ms per sample step, cores, description, kind).  kind = "reference": the vendored
    reference itself (time_reference); "port": the oracle port, when oracle/_ref is absent."""
    use_all_host_threads()
    ref = load_reference()
    if ref is not None:
        return (*time_reference(ref, method, dist, m, n, k, extra, steps, warmup, budget_s), "reference")
    # calibrate on a small block, then pick the sample size
    probe_m, probe_n = cpu_sample_shape(m, n, k, 1 << 20)
    step, cores = cpu_step_fn(method, dist, probe_m, probe_n, k, extra)
    step()
    t0 = time.perf_counter()
    step()
    per_elem = (time.perf_counter() - t0) / (probe_m * probe_n)
    total_steps = steps + warmup
    budget_elems = max(1 << 23, int(budget_s / max(per_elem, 1e-12) / total_steps))
    budget_elems = min(budget_elems, 1 << 28)                  # <= 2 GiB of fp64 V
    ms, ns = cpu_sample_shape(m, n, k, budget_elems)
    step, cores = cpu_step_fn(method, dist, ms, ns, k, extra)
    for _ in range(warmup):
        step()
    t0 = time.perf_counter()
    for _ in range(steps):
        step()
    dt = (time.perf_counter() - t0) / steps
    scale = (m * n) / (ms * ns)
    desc = (f"oracle port on a {ms}x{ns} block of the {m}x{n} workload (1/{scale:g} of V), k={k}, {steps} iterations "
            f"at {dt * 1e3:.1f} ms each; SCALED by {scale:g} to the full size (work is proportional to m*n)")
    return 1.0 / (dt * scale), dt * 1e3, blas_threads(), desc, "port"


# ----------------------------------------------------------------------------------------------
# algorithmic work
# ----------------------------------------------------------------------------------------------
def alg_flops_per_iter(method, dist, m, n, k, extra):
    if method == "mur" and dist == "eu":
        return 4.0 * m * n * k + 4.0 * (m + n) * k * k
    if method == "mur":
        return 8.0 * m * n * k
    if method == "admm":
        return 6.0 * m * n * k + 4.0 * (m + n) * k * k + 4.0 / 3.0 * k ** 3
    inner = extra.get("admm_iter", 10)
    return 4.0 * m * n * k + (2.0 * inner + 2.0) * (m + n) * k * k + 4.0 / 3.0 * k ** 3


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=100)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--workload", default=DEFAULT_WORKLOAD, choices=sorted(WORKLOADS))
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--no-e2e", action="store_true", help="skip the host-buffer end-to-end leg")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--cpu-budget", type=float, default=20.0, help="seconds of CPU work for cpu_baseline")
    return ap.parse_args()


def run_reference(args):
    """`--impl reference`: the reference algorithm (oracle port, NumPy fp64, all host threads)."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    method, dist, m, n, k, extra = WORKLOADS[args.workload]
    steps, warmup = max(1, args.steps), max(0, args.warmup)
    budget = min(150.0, 4.0 * (steps + warmup))
    its, ms_step, cores, desc, kind = time_cpu(method, dist, m, n, k, extra, steps, warmup, budget)
    line = {
        "impl": "reference", "metric": "nmf_outer_iterations_per_second", "value": its, "unit": "it/s",
        "n_gpus": args.gpus, "steps": steps, "warmup": warmup, "ms_per_step": 1e3 / its,
        "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f64",
        "data": "synthetic",
        "config": {"workload": args.workload, "method": method, "distance": dist, "m": m, "n": n, "k": k,
                   **{kk: (list(vv) if isinstance(vv, tuple) else vv) for kk, vv in extra.items()}},
        "cpu_baseline": {"value": its, "unit": "it/s", "cores": cores, "kind": kind, "sample": desc,
                         "scaled_from_sample": True},
        "e2e": {"value": its, "unit": "it/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    emit(line)


# ----------------------------------------------------------------------------------------------
# GPU arm
# ----------------------------------------------------------------------------------------------
PHASES = {  # names of the intervals between consecutive phase events of a solver step
    ("mur", "eu"): ["gram_h", "gemm_w", "update_w", "gram_w", "gemm_h", "update_h", "objective"],
}


def run_b200(args):
    import torch
    import torch.distributed as dist
    from nmf_b200 import solvers
    from nmf_b200.engine import Engine

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    method, dtype_, m, n, k, extra = WORKLOADS[args.workload]
    steps, warmup = max(1, args.steps), max(3, args.warmup)
    peaks = load_peaks()

    eng = Engine(m, n, k, device=dev, rank=rank, world=world)
    synth_fill_engine(eng, seed=0, fixed=(method in ('admm', 'ao_admm') and dtype_ == 'eu'))
    synth_init(eng, seed=1)
    stepper = solvers.make_stepper(eng, method, dtype_, **extra)
    stepper.prepare()
    obj0 = eng.read_scalar(0)                                  # objective of the initial pair
    launches0 = eng.launch_count()
    lagged = getattr(stepper, "lagged", False)

    def do_step(i):
        if lagged:          # MUR-KL: the W pass of step i also emits the objective of step i-1 (solvers.py)
            stepper.step(i % 8, (i - 1) % 8)
        else:
            stepper.step(i % 8)

    for i in range(warmup):
        do_step(i)
    torch.cuda.synchronize(dev)
    launches_per_step = (eng.launch_count() - launches0) // warmup

    names = PHASES.get((method, dtype_), stepper.phase_names())
    pairs = stepper.phase_pairs() if hasattr(stepper, "phase_pairs") else [(nm, j, j + 1) for j, nm in enumerate(names)]
    nev = 8
    evs = [[torch.cuda.Event(enable_timing=True) for _ in range(nev)] for _ in range(steps)]
    for row in evs:
        for e in row:
            e.record()                                   # materialise the cudaEvent_t handles
    t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize(dev)
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize(dev)
    with ClockSampler(local_rank) as clk:
        if world > 1 and getattr(eng, "peer", None) is not None:
            # the ranks' hosts leave the collective above milliseconds apart; without this every rank's first exchange
            # barrier waits for the last host to start, and at 8 GPUs (1 ms per step) that start-up skew was 15 % of a
            # 20-step measurement.  The GPUs leave this device-side barrier together; t0 is taken behind it.
            eng.peer.barrier(eng.stream)
        t0.record()
        h0 = time.perf_counter()
        for i in range(steps):
            eng.set_phase_events(evs[i])
            do_step(i)
        host_ms_step = (time.perf_counter() - h0) * 1e3 / steps      # time the HOST needs to enqueue one step
        t1.record()
        torch.cuda.synchronize(dev)
    eng.set_phase_events(None)
    ms_total = torch.tensor([t0.elapsed_time(t1)], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(ms_total, op=dist.ReduceOp.MAX)
        dist.barrier()
    ms_step = float(ms_total.item()) / steps
    value = 1e3 / ms_step
    obj_last = eng.read_scalar(((steps - (2 if lagged else 1)) % 8) * 8)

    phase_ms = {nm: float(np.mean([evs[i][a].elapsed_time(evs[i][b]) for i in range(steps)]))
                for nm, a, b in pairs}
    roof = stepper.roofline(phase_ms, peaks)
    if roof is not None and world == 1:
        t = NCU_TRAFFIC.get((roof.get("kernel"), args.workload))
        if t is not None:
            roof["traffic"] = t[0]
            roof["traffic_source"] = "ncu --set full, one launch: " + t[1]

    line = {
        "metric": "nmf_outer_iterations_per_second", "value": value, "unit": "it/s", "n_gpus": world,
        "steps": steps, "warmup": warmup, "ms_per_step": ms_step, "higher_is_better": True,
        "scaling": "strong", "vs_baseline": None, "dtype": "bf16x2-split/f32-accumulate",
        "data": "synthetic",
        "config": {"workload": args.workload, "method": method, "distance": dtype_, "m": m, "n": n, "k": k,
                   "v_storage": "bf16 (lossless, V and V^T resident)", "l2": "inputs_exceed_l2",
                   "objective_every_iteration": True, "parallelism": f"rows/cols x{world}",
                   "exchange": ("none" if world == 1 else
                                ("NVSwitch multicast stores (multimem.st)" if getattr(eng.peer, "mc_ptr", 0) else
                                 "peer-memory stores") + " + device barrier (csrc/peer.cu), k x k Gram terms reduced in "
                                "the same exchange" if getattr(eng, "peer", None) is not None else "torch.distributed"),
                   **{kk: (list(vv) if isinstance(vv, tuple) else vv) for kk, vv in extra.items()}},
        "clocks": clk.summary(),
        "gpu_launches": int(launches_per_step * steps),
        # enqueue time of one step on this rank's host thread: when it approaches ms_per_step the GPU waits for its host
        "host_enqueue_ms_per_step": round(host_ms_step, 4),
        "phase_ms": {kk: round(vv, 4) for kk, vv in phase_ms.items()},
        "roofline": roof,
        "alg_tflops": alg_flops_per_iter(method, dtype_, m, n, k, extra) / (ms_step * 1e-3) / 1e12 / world,
        "objective_last": obj_last,
        "peaks": peaks["source"],
    }

    # where an exchange spends its time (sharded runs over peer memory): a few extra, untimed steps with an event
    # after every leg of nmfb_exchange_factor -- window store (+ Gram term) | cross-GPU barrier | adoption | Gram sum
    if world > 1 and getattr(eng, "peer", None) is not None:
        eng.exchange_probe = []
        for i in range(5):
            do_step(i)
        torch.cuda.synchronize(dev)
        probe, eng.exchange_probe = eng.exchange_probe, None
        legs = {lab: [evs[j].elapsed_time(evs[j + 1]) for evs, _ in probe] for j, lab in enumerate(eng.PROBE_LABELS)}
        line["exchange_ms"] = {lab: round(float(np.mean(v)), 4) for lab, v in legs.items()}
        line["exchange_ms"]["exchanges_per_step"] = round(len(probe) / 5, 2)

    # ---- end to end through the public API with HOST buffers ----
    line["e2e"] = None
    _install_term_handler(line, rank)                        # a killed run still prints its device-timed line
    if not args.no_e2e and world == 1:
        try:
            line["e2e"] = run_e2e(eng, stepper, method, dtype_, m, n, k, extra, steps, obj0)
            line["time_to_tol"] = line["e2e"].pop("time_to_tol", None)
        except Exception as exc:                             # noqa: BLE001 -- reported in the line
            line["e2e"] = {"value": None, "error": f"{type(exc).__name__}: {exc}"[:300]}
    elif not args.no_e2e:
        # a failure or a stall of this leg must not cost the device-timed line: after E2E_LIMIT_S seconds
        # every rank prints / exits with what it has
        guard = threading.Timer(E2E_LIMIT_S, _e2e_timed_out, args=(line, rank))
        guard.daemon = True
        guard.start()
        try:
            line["e2e"] = run_e2e_sharded(eng, method, dtype_, m, n, k, extra, steps, world, rank, obj0)
            line["time_to_tol"] = line["e2e"].pop("time_to_tol", None)
        except Exception as exc:                             # noqa: BLE001 -- reported in the line
            line["e2e"] = {"value": None, "error": f"{type(exc).__name__}: {exc}"[:300]}
        guard.cancel()
    if not args.no_cpu and rank == 0 and world == 1:
        try:
            its, ms_cpu, cores, desc, kind = time_cpu(method, dtype_, m, n, k, extra, 3, 1, args.cpu_budget)
            line["cpu_baseline"] = {"value": its, "unit": "it/s", "cores": cores, "kind": kind, "sample": desc,
                                    "scaled_from_sample": True}
        except Exception as exc:                             # noqa: BLE001 -- reported in the line
            line["cpu_baseline"] = {"value": None, "error": f"{type(exc).__name__}: {exc}"[:300]}
    _EMITTED.set()
    if rank == 0:
        emit(line)
    if world > 1:
        dist.destroy_process_group()


TTT_MAX_ITER = 300


def run_time_to_tol(v_np, method, dtype_, k, extra, obj0):
    """End-to-end time to tolerance (BASELINE.json north_star) through the public API, host V in, host W / H
    out, the reference's own stop test switched on (utils.py:4-15: stop when the objective falls by less than
    tol2 in one iteration, or rises): tol2 = 1e-5 x the initial objective, tol1 = 0, min_iter = 10, the
    reference's default random init from the seeded global RNG (mur.py:108-109), at most TTT_MAX_ITER."""
    import contextlib
    import io

    import torch
    from nmf_b200 import NMF

    tol2 = 1e-5 * float(obj0)
    kw = dict(extra, distance_type=dtype_)
    if method != "mur":
        kw["nndsvd_init"] = (False, "zero")
    np.random.seed(1)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    model = NMF(v_np, k)
    sink = io.StringIO()
    with contextlib.redirect_stdout(sink):
        model.factorize(method=method, min_iter=10, max_iter=TTT_MAX_ITER, tol1=0.0, tol2=tol2, verbose=False, **kw)
    torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    res = model.results
    return {"seconds": dt, "iterations": res.i + 1, "converged": "Algorithm converged" in sink.getvalue(),
            "tol1": 0.0, "tol2": tol2, "min_iter": 10, "max_iter": TTT_MAX_ITER,
            "objective_first": res.obj_history[0], "objective_last": res.obj_history[-1],
            "rule": "reference stop test (utils.py:4-15) with tol2 = 1e-5 x initial objective; host V in, host W/H out"}


def run_e2e(eng, stepper, method, dtype_, m, n, k, extra, steps, obj0=None):
    """NMF(V_host, k).factorize(...) with V, W0, H0 in (pinned) host memory; the timed region
    covers upload, `steps` iterations (objective every iteration) and the download of W, H."""
    import contextlib
    import io

    import torch
    from nmf_b200 import NMF

    dev = eng.device
    v_host = torch.empty((m, n), dtype=torch.float32).pin_memory()
    chunk = 4096
    for r0 in range(0, m, chunk):
        r1 = min(m, r0 + chunk)
        blk = eng.v_rows[0, :, r0:r1, :].permute(1, 0, 2).reshape(r1 - r0, -1)[:, :n].float()
        if eng.fixed:
            blk = blk + eng.v_rows[1, :, r0:r1, :].permute(1, 0, 2).reshape(r1 - r0, -1)[:, :n].float()
        v_host[r0:r1].copy_(blk)
    w0 = eng.w.master[:k, :m].t().contiguous().cpu().numpy()     # current factors as the start
    h0 = eng.h.master[:k, :n].contiguous().cpu().numpy()
    torch.cuda.synchronize(dev)
    eng.release()
    torch.cuda.empty_cache()
    v_np = v_host.numpy()
    kw = dict(extra, distance_type=dtype_)
    if method != "mur":
        kw["nndsvd_init"] = (False, "zero")
    torch.cuda.synchronize(dev)
    t0 = time.perf_counter()
    model = NMF(v_np, k)
    with contextlib.redirect_stdout(io.StringIO()):
        model.factorize(method=method, min_iter=10**9, max_iter=steps, w0=w0, h0=h0, verbose=False, **kw)
    torch.cuda.synchronize(dev)
    dt = time.perf_counter() - t0
    res = model.results
    assert res.i == steps - 1 and len(res.obj_history) == steps + 1
    h2d = v_np.nbytes + w0.nbytes + h0.nbytes
    d2h = res.w.nbytes // 2 + res.h.nbytes // 2 + 8 * (steps + 1)    # factors travel as fp32
    out = {"value": steps / dt, "unit": "it/s", "seconds_total": dt,
           "h2d_bytes_per_step": h2d // steps, "d2h_bytes_per_step": d2h // steps,
           "what": "NMF(V_host fp32 pinned, k).factorize(max_iter=steps): ingest + iterations + factor download"}
    if obj0 is not None:
        try:
            out["time_to_tol"] = run_time_to_tol(v_np, method, dtype_, k, extra, obj0)
        except Exception as exc:                             # noqa: BLE001 -- reported, never costs the line
            out["time_to_tol"] = {"seconds": None, "error": f"{type(exc).__name__}: {exc}"[:300]}
    return out


E2E_LIMIT_S = 300.0
_EMITTED = threading.Event()


def _install_term_handler(line, rank):
    """From here on the device-timed fields are complete: if the launcher tears the job down during one of the
    later legs (another rank died, an outer time limit), rank 0 still prints them before it exits."""
    import signal

    def on_term(signum, _frame):
        if not _EMITTED.is_set():
            _EMITTED.set()
            if line.get("e2e") is None:
                line["e2e"] = {"value": None, "error": f"terminated by signal {signum} during a host-side leg; "
                                                       "device-timed fields are complete"}
            if rank == 0:
                emit(line)
        os._exit(0 if rank == 0 else 1)

    try:
        signal.signal(signal.SIGTERM, on_term)
    except (ValueError, OSError):                            # not the main thread (tests drive run_b200 from one)
        pass


def _e2e_timed_out(line, rank):
    if _EMITTED.is_set():
        return
    _EMITTED.set()
    line["e2e"] = {"value": None, "error": f"end-to-end leg exceeded {E2E_LIMIT_S:.0f} s; device-timed fields are complete"}
    if rank == 0:
        emit(line)
    os._exit(0)


def run_e2e_sharded(eng, method, dtype_, m, n, k, extra, steps, world, rank, obj0=None):
    """N > 1: the same call on every rank, NMF(V_host, k).factorize(..., comm=True) -- each rank is handed the
    whole host V (the public API of a sharded run) and reads ONLY its rows of it
    (Engine._shard_and_panel: 1/P of V per PCIe link; the column copy is assembled over NVLink by peer-memory
    stores).  Timed from a
    barrier to the slowest rank's return (engine set-up incl. the peer arena, upload, `steps` iterations with the
    objective read back every iteration, download of W and H on every rank); `breakdown` = the slowest rank's
    wall seconds per leg.  The host V is ONE shared-memory file that rank 0 fills and every rank maps, so the node
    holds one copy instead of N pinned ones."""
    import contextlib
    import io

    import psutil
    import torch
    import torch.distributed as dist
    from nmf_b200 import NMF
    dev = eng.device
    engine_cls = type(eng)                                   # the product Engine (a host stand-in in the CPU tests)

    def agreed(flag: bool) -> bool:                          # every rank takes the same branch
        t = torch.tensor([1 if flag else 0], dtype=torch.int32, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MIN)
        return bool(t.item())

    need = 4 * m * n
    dist.barrier()
    if not agreed(psutil.virtual_memory().available > 1.3 * need):
        return {"value": None, "error": f"host memory: {need / 2**30:.1f} GiB of V do not fit"}
    w0 = eng.w.master[:k, :m].t().contiguous().cpu().numpy()     # current (replicated) factors as the start
    h0 = eng.h.master[:k, :n].contiguous().cpu().numpy()
    torch.cuda.synchronize(dev)
    eng.release()
    torch.cuda.empty_cache()
    import shutil
    path = f"/dev/shm/nmfb_bench_v_{os.environ.get('MASTER_PORT', '0')}.npy"
    v_np, pinned = None, False
    shared = agreed(os.path.isdir("/dev/shm") and shutil.disk_usage("/dev/shm").free > 1.1 * need)
    try:
        if shared and rank == 0:
            v_np = np.lib.format.open_memmap(path, mode="w+", dtype=np.float32, shape=(m, n))
        elif not shared:                                     # no room in /dev/shm: a private copy per rank
            v_np = np.empty((m, n), dtype=np.float32)
        if v_np is not None:
            _wt, _ht, blocks = synth_blocks(dev, m, n, k, seed=0)
            for r0, blk in blocks():
                v_np[r0:r0 + blk.shape[0]] = blk.cpu().numpy()
            del _wt, _ht, blk
            if shared:
                v_np.flush()
        dist.barrier()
        if shared and rank != 0:
            v_np = np.load(path, mmap_mode="r+")
    except Exception:                                        # noqa: BLE001 -- agreed on below
        v_np = None
    if v_np is not None:
        try:   # page-lock the mapping in this process (every rank maps the same physical pages)
            pinned = int(torch.cuda.cudart().cudaHostRegister(v_np.ctypes.data, v_np.nbytes, 0)) == 0
        except Exception:                                    # noqa: BLE001 -- pageable memory is slower, not wrong
            pinned = False
    if not agreed(v_np is not None):
        return {"value": None, "error": "host V (shared memory) could not be set up on a rank"}
    torch.cuda.empty_cache()
    kw = dict(extra, distance_type=dtype_)
    if method != "mur":
        kw["nndsvd_init"] = (False, "zero")

    def timed(**run_kw):
        torch.cuda.synchronize(dev)
        dist.barrier()
        t0 = time.perf_counter()
        engine = engine_cls.for_comm(m, n, k, True)
        t_engine = time.perf_counter() - t0
        model = NMF(v_np, k)
        sink = io.StringIO()
        with contextlib.redirect_stdout(sink):
            model.factorize(method=method, verbose=False, comm=True, engine=engine, **run_kw, **kw)
        torch.cuda.synchronize(dev)
        total = time.perf_counter() - t0
        legs = [total, t_engine, engine.timings.get("upload_s", 0.0), engine.timings.get("download_s", 0.0)]
        t = torch.tensor(legs, dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        total, t_engine, t_up, t_down = (float(x) for x in t.tolist())
        engine.release()
        bd = {"engine_and_peer_arena_s": round(t_engine, 4), "upload_s": round(t_up, 4),
              "iterations_s": round(max(total - t_engine - t_up - t_down, 0.0), 4), "download_s": round(t_down, 4)}
        return model.results, total, bd, sink.getvalue()

    res, dt, breakdown, _ = timed(min_iter=10**9, max_iter=steps, w0=w0, h0=h0)
    assert res.i == steps - 1 and len(res.obj_history) == steps + 1
    # each rank uploads its rows of V (the columns travel between the GPUs) and both initial factors
    h2d = v_np.nbytes + world * (w0.nbytes + h0.nbytes)
    d2h = world * (res.w.nbytes // 2 + res.h.nbytes // 2 + 8 * (steps + 1))
    out = {"value": steps / dt, "unit": "it/s", "seconds_total": dt, "breakdown": breakdown,
           "h2d_bytes_per_step": h2d // steps, "d2h_bytes_per_step": d2h // steps, "host_v_pinned": pinned,
           "what": f"on each of {world} ranks NMF(V_host fp32 in shared memory, k).factorize(max_iter=steps, comm=True): "
                   "engine + peer arena, upload of the rank's rows of V (columns over NVLink), iterations, factor download; "
                   "slowest rank"}
    if obj0 is not None:
        try:
            tol2 = 1e-5 * float(obj0)
            np.random.seed(1)
            r2, dt2, bd2, text = timed(min_iter=10, max_iter=TTT_MAX_ITER, tol1=0.0, tol2=tol2)
            out["time_to_tol"] = {"seconds": dt2, "iterations": r2.i + 1, "converged": "Algorithm converged" in text
                                  or r2.i + 1 < TTT_MAX_ITER, "tol1": 0.0, "tol2": tol2, "min_iter": 10,
                                  "max_iter": TTT_MAX_ITER, "breakdown": bd2,
                                  "objective_first": r2.obj_history[0], "objective_last": r2.obj_history[-1],
                                  "rule": "reference stop test (utils.py:4-15) with tol2 = 1e-5 x initial objective; "
                                          "host V in, host W/H out, slowest rank"}
        except Exception as exc:                             # noqa: BLE001 -- reported, never costs the line
            out["time_to_tol"] = {"seconds": None, "error": f"{type(exc).__name__}: {exc}"[:300]}
    if pinned:
        with contextlib.suppress(Exception):
            torch.cuda.cudart().cudaHostUnregister(v_np.ctypes.data)
    del v_np
    dist.barrier()
    if rank == 0 and shared:
        with contextlib.suppress(OSError):
            os.remove(path)
    out["host_v_shared"] = shared
    return out


_JSON_FD = None


def emit(line: dict):
    """The contract is ONE JSON line on stdout: libraries (NCCL prints its version banner there when
    NCCL_DEBUG is set on the box) write to fd 2 for the duration of the run, the line goes to the real fd 1."""
    data = (json.dumps(line) + "\n").encode()
    if _JSON_FD is None:
        sys.stdout.write(data.decode())
        sys.stdout.flush()
    else:
        os.write(_JSON_FD, data)


def main():
    global _JSON_FD
    args = parse_args()
    sys.stdout.flush()
    _JSON_FD = os.dup(1)
    os.dup2(2, 1)
    if args.impl == "reference":
        run_reference(args)
    else:
        run_b200(args)


if __name__ == "__main__":
    main()
