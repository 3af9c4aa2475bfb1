#!/usr/bin/env python3
"""Kernel-level self tests + micro-benchmarks, one sub-command per process so that a faulting
kernel cannot poison the CUDA context of the others.  Run on the GPU box:

    python tools/selftest.py gemm | gram | update | objective | mur | bench_gemm | bench_obj
"""
import ctypes as C
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from nmf_b200 import _lib  # noqa: E402
from nmf_b200._lib import Data, check  # noqa: E402
from nmf_b200.engine import Engine  # noqa: E402


def bf16_round(a):
    u = np.ascontiguousarray(a, dtype=np.float32).view(np.uint32)
    r = ((u >> 16) & 1) + np.uint32(0x7FFF)
    return ((u + r) & np.uint32(0xFFFF0000)).view(np.float32)


def rel(a, b):
    return float(np.linalg.norm(np.asarray(a, np.float64) - b) / (np.linalg.norm(b) + 1e-300))


def sum_slots(eng, slots, Lp):
    KP = eng.KP
    p = eng.partial[: slots * KP * Lp].view(slots, KP, Lp).double().sum(0)
    return p.cpu().numpy()


def make_engine(m, n, k, exact=True, seed=0, signed_w=False):
    rng = np.random.default_rng(seed)
    v = rng.random((m, n), dtype=np.float32) * 4
    if exact:
        v = bf16_round(v)
    eng = Engine(m, n, k)
    big = 16 * eng.KP * max(eng.Mp, eng.Np)                 # room for forced K-splits
    if eng.partial.numel() < big:
        eng.partial = torch.empty(big, dtype=torch.float32, device=eng.device)
        eng._problem = None
    eng.upload_data(v)
    w0 = rng.random((m, k)).astype(np.float64) + 0.05
    h0 = rng.random((k, n)).astype(np.float64) + 0.05
    if signed_w:
        w0 -= 0.5
    eng.set_factors(w0, h0)
    return eng, v.astype(np.float64), w0, h0


def test_gemm():
    ok = True
    lib = _lib.load()
    cases = [(300, 200, 10, True, 2), (1000, 500, 100, True, 2), (520, 900, 200, True, 2),
             (300, 333, 24, False, 2), (640, 384, 128, False, 2), (1000, 500, 100, True, 3),
             (300, 200, 48, True, 3), (4096, 8192, 64, True, 2), (8192, 4096, 128, True, 2)]
    for (m, n, k, exact, nb) in cases:
        eng, v, w0, h0 = make_engine(m, n, k, exact, signed_w=(nb == 3))
        w32, h32 = w0.astype(np.float32).astype(np.float64), h0.astype(np.float32).astype(np.float64)
        vr, vc = eng._data_structs()
        for side in ("W", "H"):
            for force_split in (0, 1, 3):
                if side == "W":      # P^T[c][i] = sum_j H[c][j] V[i][j]
                    slots = check(lib.nmfb_panel_gemm(C.byref(vr), eng.h.cplanes.data_ptr(), eng.h.Lp, eng.KP, nb,
                                                      eng.partial.data_ptr(), eng.partial.numel() * 4, eng.w.Lp,
                                                      eng.num_sms, 0, force_split, eng.stream))
                    got = sum_slots(eng, slots, eng.w.Lp)[:k, :m]
                    want = h32 @ v.T
                else:
                    slots = check(lib.nmfb_panel_gemm(C.byref(vc), eng.w.cplanes.data_ptr(), eng.w.Lp, eng.KP, nb,
                                                      eng.partial.data_ptr(), eng.partial.numel() * 4, eng.h.Lp,
                                                      eng.num_sms, 0, force_split, eng.stream))
                    got = sum_slots(eng, slots, eng.h.Lp)[:k, :n]
                    want = w32.T @ v
                torch.cuda.synchronize()
                e = rel(got, want)
                tol = 2e-5 if not exact else 3e-6
                flag = "OK " if e < tol else "BAD"
                ok &= e < tol
                print(f"gemm {flag} m={m} n={n} k={k} KP={eng.KP} nv={eng.nv} nb={nb} side={side} "
                      f"split={force_split} slots={slots} rel={e:.3e}", flush=True)
                if e >= tol:
                    d = np.abs(got - want)
                    idx = np.unravel_index(np.argmax(d), d.shape)
                    print("   worst", idx, got[idx], want[idx], " first row got/want", got[0, :4], want[0, :4])
        del eng
    return ok


def test_gram():
    ok = True
    lib = _lib.load()
    for (m, n, k) in [(300, 200, 10), (1000, 500, 100), (520, 900, 200), (5000, 3000, 128)]:
        eng, v, w0, h0 = make_engine(m, n, k)
        p = eng.problem()
        for which, fac, ref in ((0, eng.w, w0.astype(np.float32).astype(np.float64)),
                                (1, eng.h, h0.astype(np.float32).astype(np.float64).T)):
            fs = fac.struct()
            check(lib.nmfb_gram(C.byref(fs), eng.KP, 2, eng.partial.data_ptr(), eng.partial.numel() * 4,
                                eng.gram32[which].data_ptr(), eng.gram64[which].data_ptr(), eng.num_sms,
                                eng.stream))
            torch.cuda.synchronize()
            got = eng.gram64[which].cpu().numpy()[:k, :k]
            want = ref.T @ ref
            e = rel(got, want)
            pad = float(eng.gram64[which].abs().sum().item() - np.abs(got).sum())
            flag = "OK " if e < 3e-6 and abs(pad) < 1e-6 else "BAD"
            ok &= flag == "OK "
            print(f"gram {flag} m={m} n={n} k={k} which={which} rel={e:.3e} pad_leak={pad:.3e}", flush=True)
    return ok


def test_update():
    ok = True
    lib = _lib.load()
    for (m, n, k, lam) in [(300, 200, 10, 0.0), (1000, 500, 100, 0.1), (520, 900, 200, 0.05)]:
        eng, v, w0, h0 = make_engine(m, n, k)
        w32, h32 = w0.astype(np.float32).astype(np.float64), h0.astype(np.float32).astype(np.float64)
        p = eng.problem()
        KP = eng.KP
        hs = eng.h.struct()
        check(lib.nmfb_gram(C.byref(hs), KP, 2, eng.partial.data_ptr(), eng.partial.numel() * 4,
                            eng.gram32[1].data_ptr(), eng.gram64[1].data_ptr(), eng.num_sms, eng.stream))
        vr, vc = eng._data_structs()
        slots = check(lib.nmfb_panel_gemm(C.byref(vr), eng.h.cplanes.data_ptr(), eng.h.Lp, KP, 2,
                                          eng.partial.data_ptr(), eng.partial.numel() * 4, eng.w.Lp, eng.num_sms,
                                          0, 0, eng.stream))
        ws = eng.w.struct()
        check(lib.nmfb_mur_eu_update(eng.partial.data_ptr(), slots, eng.gram32[1].data_ptr(), C.byref(ws), k, KP,
                                     lam, eng.stream))
        torch.cuda.synchronize()
        want = w32 * (v @ h32.T) / (w32 @ (h32 @ h32.T) + lam * w32 + 1e-9)
        got_w, _ = eng.get_factors()
        e = rel(got_w, want)
        # planes must reproduce the master
        mast = eng.w.master.double()
        cp = eng.w.cplanes[:2, :KP].double().sum(0)
        rp = eng.w.rplanes.double().sum(0).t()
        e_c = float((cp - mast).norm() / mast.norm())
        e_r = float((rp - mast).norm() / mast.norm())
        padsum = float(mast[k:].abs().sum() + mast[:, m:].abs().sum())
        flag = "OK " if (e < 3e-6 and e_c < 2e-5 and e_r < 2e-5 and padsum == 0) else "BAD"
        ok &= flag == "OK "
        print(f"update {flag} m={m} n={n} k={k} rel={e:.3e} cplanes={e_c:.2e} rplanes={e_r:.2e} pad={padsum}",
              flush=True)
    return ok


def test_objective():
    ok = True
    for (m, n, k, exact) in [(300, 200, 10, True), (1000, 500, 100, True), (520, 900, 200, True),
                             (300, 333, 24, False), (2048, 4096, 128, True), (4096, 2048, 64, True)]:
        eng, v, w0, h0 = make_engine(m, n, k, exact)
        w32, h32 = w0.astype(np.float32).astype(np.float64), h0.astype(np.float32).astype(np.float64)
        want = 0.5 * np.sum((v - w32 @ h32) ** 2)
        eng.objective_eu(slot=0, impl=1)
        eng.objective_eu(slot=1, impl=0)
        torch.cuda.synchronize()
        simt, tc = eng.read_scalar(0), eng.read_scalar(1)
        e1, e2 = abs(simt - want) / want, abs(tc - want) / want
        tol = 1e-5 if exact else 1e-4
        flag = "OK " if (e1 < tol and e2 < tol) else "BAD"
        ok &= flag == "OK "
        print(f"objective {flag} m={m} n={n} k={k} nv={eng.nv} want={want:.6e} simt_rel={e1:.2e} tc_rel={e2:.2e}",
              flush=True)
    # planted: tiny residual relative to |V| -- the case the residual form exists for
    from oracle.nmf_oracle import planted
    m, n, k = 1024, 768, 32
    v = planted(m, n, k, 1)
    rng = np.random.default_rng(1)
    eng = Engine(m, n, k)
    eng.upload_data(v)
    wt = rng.random((m, k), dtype=np.float32).astype(np.float64)
    ht = np.linalg.lstsq(wt, v.astype(np.float64), rcond=None)[0]
    ht = np.abs(ht)
    eng.set_factors(wt, ht)
    w32, h32 = wt.astype(np.float32).astype(np.float64), ht.astype(np.float32).astype(np.float64)
    want = 0.5 * np.sum((v.astype(np.float64) - w32 @ h32) ** 2)
    eng.objective_eu(slot=0, impl=1)
    eng.objective_eu(slot=1, impl=0)
    simt, tc = eng.read_scalar(0), eng.read_scalar(1)
    print(f"objective planted want={want:.6e} simt_rel={abs(simt-want)/want:.2e} tc_rel={abs(tc-want)/want:.2e} "
          f"|V|^2/2={0.5*np.sum(v.astype(np.float64)**2):.3e}", flush=True)
    return ok


def test_mur():
    from oracle import nmf_oracle as orc
    from nmf_b200.mur import mur
    ok = True
    for (m, n, k, exact, iters) in [(1000, 500, 10, False, 100), (512, 384, 16, True, 100),
                                    (2048, 1536, 64, True, 60), (1500, 2500, 128, True, 40)]:
        if exact:
            v = orc.planted(m, n, k, 0).astype(np.float64)
        else:
            v = np.random.default_rng(0).random((m, n))
        w0, h0 = orc.seeded_init(m, n, k, 1)
        t0 = time.time()
        res = mur(v.copy(), k, distance_type='eu', min_iter=10**9, max_iter=iters, w0=w0, h0=h0, verbose=False)
        t1 = time.time()
        tr = orc.mur_run(v, k, w0, h0, distance_type='eu', min_iter=10**9, max_iter=iters)
        ew, eh = rel(res.w, tr.w), rel(res.h, tr.h)
        eo = max(abs(a - b) / b for a, b in zip(res.obj_history, tr.obj_history))
        flag = "OK " if max(ew, eh, eo) < 1e-4 else "BAD"
        ok &= flag == "OK "
        print(f"mur-eu {flag} m={m} n={n} k={k} exact={exact} iters={iters} relW={ew:.2e} relH={eh:.2e} "
              f"relObj={eo:.2e} gpu_s={t1-t0:.2f} obj_last={res.obj_history[-1]:.6f}/{tr.obj_history[-1]:.6f}",
              flush=True)
    return ok


def _time(fn, iters=5, warm=2):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters


def bench(m, n, k, what):
    lib = _lib.load()
    eng = Engine(m, n, k)
    big = 16 * eng.KP * max(eng.Mp, eng.Np)
    if eng.partial.numel() < big:
        eng.partial = torch.empty(big, dtype=torch.float32, device=eng.device)
        eng._problem = None
    g = torch.Generator(device='cuda').manual_seed(0)
    # synthesize bf16-exact planes directly on the device (values in [0, 4))
    eng._alloc_data(1)
    for r0 in range(0, m, 4096):
        r1 = min(m, r0 + 4096)
        eng.store_block((torch.rand((r1 - r0, n), device='cuda', generator=g) * 4).bfloat16(), r0)
    w0 = torch.rand((k, m), device='cuda', generator=g) + 0.05
    h0 = torch.rand((k, n), device='cuda', generator=g) + 0.05
    eng.w.master[:k, :m] = w0
    eng.h.master[:k, :n] = h0
    for f in (eng.w, eng.h):
        fs = f.struct()
        check(lib.nmfb_refresh_planes(eng.KP, C.byref(fs), eng.stream))
    vr, vc = eng._data_structs()
    KP = eng.KP
    if what in ("gemm", "all"):
        for fs in (0, 1, 2, 4, 8):
            f = lambda: check(lib.nmfb_panel_gemm(C.byref(vr), eng.h.cplanes.data_ptr(), eng.h.Lp, KP, 2,
                                                  eng.partial.data_ptr(), eng.partial.numel() * 4, eng.w.Lp,
                                                  eng.num_sms, 0, fs, eng.stream))
            slots = f()
            ms = _time(f)
            fl = 2.0 * m * n * k
            print(f"bench gemm W-side m={m} n={n} k={k} split={fs} slots={slots} {ms:.3f} ms  "
                  f"alg {fl/ms/1e9:.1f} TF/s  exec {2*fl/ms/1e9:.1f} TF/s  V {m*n*2/ms/1e6:.0f} GB/s", flush=True)
        f = lambda: check(lib.nmfb_panel_gemm(C.byref(vc), eng.w.cplanes.data_ptr(), eng.w.Lp, KP, 2,
                                              eng.partial.data_ptr(), eng.partial.numel() * 4, eng.h.Lp,
                                              eng.num_sms, 0, 0, eng.stream))
        ms = _time(f)
        print(f"bench gemm H-side {ms:.3f} ms alg {2.0*m*n*k/ms/1e9:.1f} TF/s", flush=True)
    if what in ("parts", "all"):
        hs = eng.h.struct()
        ws = eng.w.struct()
        f = lambda: check(lib.nmfb_gram(C.byref(hs), KP, 2, eng.partial.data_ptr(), eng.partial.numel() * 4,
                                        eng.gram32[1].data_ptr(), eng.gram64[1].data_ptr(), eng.num_sms, eng.stream))
        print(f"bench gram(H) {_time(f):.3f} ms", flush=True)
        slots = check(lib.nmfb_panel_gemm(C.byref(vr), eng.h.cplanes.data_ptr(), eng.h.Lp, KP, 2,
                                          eng.partial.data_ptr(), eng.partial.numel() * 4, eng.w.Lp, eng.num_sms,
                                          0, 0, eng.stream))
        f = lambda: check(lib.nmfb_mur_eu_update(eng.partial.data_ptr(), slots, eng.gram32[1].data_ptr(),
                                                 C.byref(ws), k, KP, 0.0, eng.stream))
        print(f"bench update(W) slots={slots} {_time(f):.3f} ms", flush=True)
    if what in ("obj", "all"):
        f = lambda: eng.objective_eu(slot=1, impl=0)
        ms = _time(f)
        print(f"bench objective tc m={m} n={n} k={k} {ms:.3f} ms alg {2.0*m*n*k/ms/1e9:.1f} TF/s "
              f"exec {6.0*m*n*k/ms/1e9:.1f} TF/s V {m*n*2/ms/1e6:.0f} GB/s", flush=True)
        # experimental variants (opt-in): 2 = wide tiles (residual_n256.cu), 3 = W planes in TMEM (residual_ts.cu);
        # same value first, then the time.  SELFTEST_VARIANTS="2,3" (default) / "" to skip.
        for impl in [int(x) for x in os.environ.get("SELFTEST_VARIANTS", "2,3").split(",") if x]:
            if impl == 2 and not (KP <= 128 and eng.nv == 1):
                continue
            name = {2: "tc-wide", 3: "tc-tmemA"}[impl]
            eng.objective_eu(slot=1, impl=0)
            eng.objective_eu(slot=2, impl=impl)
            a, b = eng.read_scalar(1), eng.read_scalar(2)
            print(f"objective {name} vs default: {a:.9e} {b:.9e} rel {abs(a - b) / a:.2e}", flush=True)
            f = lambda: eng.objective_eu(slot=2, impl=impl)
            ms = _time(f)
            print(f"bench objective {name} m={m} n={n} k={k} {ms:.3f} ms alg {2.0*m*n*k/ms/1e9:.1f} TF/s "
                  f"exec {6.0*m*n*k/ms/1e9:.1f} TF/s V {m*n*2/ms/1e6:.0f} GB/s", flush=True)
    if what in ("step", "all"):
        f = lambda: eng.mur_eu_step(0.0, 0.0, obj_slot=0)
        ms = _time(f, iters=5, warm=2)
        print(f"bench mur_eu_step m={m} n={n} k={k} {ms:.3f} ms/iter = {1000/ms:.1f} it/s", flush=True)
        f = lambda: eng.mur_eu_step(0.0, 0.0, obj_slot=None)
        ms = _time(f, iters=5, warm=1)
        print(f"bench mur_eu_step(no objective) {ms:.3f} ms/iter = {1000/ms:.1f} it/s", flush=True)


if __name__ == "__main__":
    cmd = sys.argv[1]
    if cmd == "bench":
        m, n, k = int(sys.argv[2]), int(sys.argv[3]), int(sys.argv[4])
        bench(m, n, k, sys.argv[5] if len(sys.argv) > 5 else "all")
        sys.exit(0)
    fn = {"gemm": test_gemm, "gram": test_gram, "update": test_update, "objective": test_objective,
          "mur": test_mur}[cmd]
    good = fn()
    print(f"SELFTEST {cmd}: {'PASS' if good else 'FAIL'}")
    sys.exit(0 if good else 1)
