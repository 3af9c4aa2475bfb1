"""CPU oracle for the NMF solver hot path -- TEST INFRASTRUCTURE ONLY.

This module is a float64 NumPy restatement of the solver loops of the reference
(`raleng/nmf`).  It exists to check the CUDA path (tests/, __graft_entry__.smoke)
and to be timed as the CPU baseline (bench.py cpu_baseline / --impl reference).
The product package `nmf_b200` never imports it.

Parity pinning: the reference ships no tests, fixtures or golden vectors
(SURVEY.md section 4), so this restatement is pinned by *executing the reference*
in the build container: `tests/golden/make_golden.py` imports `/root/reference`,
runs it from fixed seeds and freezes its outputs under `tests/golden/*.npz`;
`tests/test_oracle.py` replays the same seeds through this file and demands
agreement to ~1e-12.  Third-party arithmetic below the reference (not vendored in
/root/reference): numpy ==1.25.2 and scipy ==1.11.1 by its Pipfile.lock; this
image has numpy 2.3.5 / scipy 1.18.1 (dgemm / potrf / gesv results agree to ~1e-15).

Every function cites the reference lines it follows.  State is passed explicitly
(w0/h0 in, per-iteration trace out) because the parity tests need the trajectory,
which the reference API cannot provide.
"""
from __future__ import annotations

from dataclasses import dataclass, field
from typing import Callable, List, Optional

import numpy as np
import scipy.linalg as sla

EPS_MUR = 1e-9          # mur.py:25,29,41,45
INNER_TOL = 1e-2        # ao_admm.py:33 (hard-coded tol of `terminate`)


# --------------------------------------------------------------------------- #
# shared numerics (utils.py)
# --------------------------------------------------------------------------- #
def objective(v: np.ndarray, wh: np.ndarray, kind: str = "eu") -> float:
    """utils.py:18-33 -- 0.5*||V-WH||_F^2, or generalised KL with +inf/NaN
    terms of v*log(v/wh) mapped to 0 (but not -inf)."""
    if kind == "eu":
        r = v - wh
        return float(0.5 * np.sum(r * r))
    if kind == "kl":
        with np.errstate(divide="ignore", invalid="ignore"):
            t = v * np.log(v / wh)
        t[t == np.inf] = 0.0
        t[np.isnan(t)] = 0.0
        return float(np.sum(t - v + wh))
    raise KeyError('Distance type unknown: use "kl" or "eu"')


def stop_test(new: float, old: float, tol1: float, tol2: float) -> int:
    """utils.py:4-15 -- returns 0 (continue), 1 (new < tol1) or 2 (new >= old - tol2).
    Absolute tolerances on the raw objective."""
    if new < tol1:
        return 1
    if new >= old - tol2:
        return 2
    return 0


def reference_init(m: int, n: int, k: int):
    """mur.py:108-109, admm.py:23-24, ao_admm.py:22-23 -- |randn(m,k)| then |randn(k,n)|
    from the *global* NumPy RNG, in that order."""
    w = np.abs(np.random.randn(m, k))
    h = np.abs(np.random.randn(k, n))
    return w, h


def nndsvd(x: np.ndarray, rank: int, variant: str = "zero"):
    """utils.py:36-93 -- Boutsidis/Gallopoulos NNDSVD (zero / mean / random fill)."""
    u, s, vt = np.linalg.svd(x, full_matrices=False)
    m, n = x.shape
    w = np.zeros((m, rank))
    h = np.zeros((rank, n))
    w[:, 0] = np.sqrt(s[0]) * np.abs(u[:, 0])
    h[0, :] = np.sqrt(s[0]) * np.abs(vt[0, :])
    for j in range(1, rank):
        uj, vj = u[:, j], vt[j, :]
        up, un = np.maximum(uj, 0.0), np.maximum(-uj, 0.0)
        vp, vn = np.maximum(vj, 0.0), np.maximum(-vj, 0.0)
        nup, nun = np.linalg.norm(up), np.linalg.norm(un)
        nvp, nvn = np.linalg.norm(vp), np.linalg.norm(vn)
        pos, neg = nup * nvp, nun * nvn
        if pos >= neg:
            w[:, j] = np.sqrt(s[j] * pos) / nup * up
            h[j, :] = np.sqrt(s[j] * pos) / nvp * vp
        else:
            w[:, j] = np.sqrt(s[j] * neg) / nun * un
            h[j, :] = np.sqrt(s[j] * neg) / nvn * vn
    if variant == "mean":
        mu = np.mean(x)
        w[w == 0] = mu
        h[h == 0] = mu
    elif variant == "random":
        mu = np.mean(x)
        fill = mu * np.random.random_sample(w.shape) / 100
        w = np.where(w == 0, fill, w)
        fill = mu * np.random.random_sample(h.shape) / 100
        h = np.where(h == 0, fill, h)
    return w, h


@dataclass
class Trace:
    """What a run returns: final factors, last loop index, objective history
    (len = i + 2, like the reference) and, if requested, per-iteration snapshots."""
    w: np.ndarray
    h: np.ndarray
    i: int
    obj_history: List[float]
    stop_code: int = 0
    snaps: list = field(default_factory=list)
    inner_breaks: list = field(default_factory=list)


# --------------------------------------------------------------------------- #
# MUR  (mur.py)
# --------------------------------------------------------------------------- #
def mur_w_step(kind, v, w, h, wh, lam):
    """mur.py:20-33."""
    if kind == "eu":
        return w * (v @ h.T) / (wh @ h.T + lam * w + EPS_MUR)
    if kind == "kl":
        a = w * ((v / (wh + EPS_MUR)) @ h.T)
        b = np.broadcast_to(h.sum(axis=1), w.shape)      # == ones_like(v) @ h.T
        return 2 * a / (b + np.sqrt(b * b + 4 * lam * a))
    raise KeyError("Unknown distance type.")


def mur_h_step(kind, v, w, h, wh, lam):
    """mur.py:36-49."""
    if kind == "eu":
        return h * (w.T @ v) / (w.T @ wh + lam * h + EPS_MUR)
    if kind == "kl":
        c = h * (w.T @ (v / (wh + EPS_MUR)))
        d = np.broadcast_to(w.sum(axis=0)[:, None], h.shape)   # == w.T @ ones_like(v)
        return 2 * c / (d + np.sqrt(d * d + 4 * lam * c))
    raise KeyError("Unknown distance type.")


def mur_run(v, k, w0, h0, *, distance_type="kl", min_iter=100, max_iter=100000,
            tol1=1e-5, tol2=1e-5, lambda_w=0.0, lambda_h=0.0,
            snap: Optional[Callable[[int], bool]] = None) -> Trace:
    """mur.py:52-146 driver loop (W step with the old WH, H step with W_new @ H_old,
    objective on the new pair, stop test only when i > min_iter)."""
    v = np.asarray(v, dtype=np.float64)
    if v.min() < 0:                      # mur.py:99-102 (the reference shifts in place)
        v = v + abs(v.min())
    w, h = np.array(w0, dtype=np.float64), np.array(h0, dtype=np.float64)
    wh = w @ h
    hist = [objective(v, wh, distance_type)]
    out = Trace(w, h, -1, hist)
    for i in range(max_iter):
        w = mur_w_step(distance_type, v, w, h, wh, lambda_w)
        h = mur_h_step(distance_type, v, w, h, w @ h, lambda_h)
        wh = w @ h
        hist.append(objective(v, wh, distance_type))
        out.i = i
        if snap is not None and snap(i):
            out.snaps.append((i, w.copy(), h.copy()))
        if i > min_iter:
            code = stop_test(hist[-1], hist[-2], tol1, tol2)
            if code:
                out.stop_code = code
                break
    out.w, out.h = w, h
    return out


# --------------------------------------------------------------------------- #
# proximal operators shared by ADMM and AO-ADMM
# --------------------------------------------------------------------------- #
def tikhonov_system(kk: int, lam: float, rho: float) -> np.ndarray:
    """admm.py:142-150: a = 1/rho * (lam * T^T T + rho * I) with T = tridiag(-1, 2, -1) of order kk
    (the component axis: `n = mat_aux.shape[0]`, and prox is always handed k x cols arrays)."""
    t = 2.0 * np.eye(kk) - np.eye(kk, k=1) - np.eye(kk, k=-1)
    return (lam / rho) * (t.T @ t) + np.eye(kk)


def prox(kind: str, aux, dual, rho, lam):
    """admm.py:126-156 / ao_admm.py:113-135 on component-major (k x cols) arrays.
    'l2n': solve the Tikhonov system along the component axis, then clip (admm.py:141-156; with
    lam == 0 the system is the identity, i.e. 'nn').  The reference uses a sparse solve of the same
    k x k pentadiagonal SPD matrix; ao_admm.py:128 builds it without dtype=object and therefore raises
    on NumPy >= 1.24 -- for AO-ADMM this restates the intended arithmetic (parity unpinned there)."""
    if kind == "nn" or (kind == "l2n" and lam == 0):
        return np.maximum(aux - dual, 0.0)
    if kind == "l1n":
        return np.maximum(aux - dual - lam / rho, 0.0)
    if kind == "l2n":
        a = tikhonov_system(aux.shape[0], lam, rho)
        return np.maximum(np.linalg.solve(a, aux - dual), 0.0)
    raise TypeError("Unknown prox_type.")


# --------------------------------------------------------------------------- #
# ADMM  (admm.py)
# --------------------------------------------------------------------------- #
def admm_aux(x, dual, other_aux, data, rho):
    """admm.py:216-221,230 (eu): solve (A^T A + rho I) X = A^T D + rho (x + dual), LU."""
    kk = other_aux.shape[1]
    a = other_aux.T @ other_aux + rho * np.eye(kk)
    b = other_aux.T @ data + rho * (x + dual)
    return np.linalg.solve(a, b)


def admm_aux_kl(x, dual, other_aux, v_aux, dual_v, rho):
    """admm.py:223-225,230 (kl): the data term is the auxiliary copy of V plus its dual."""
    kk = other_aux.shape[1]
    a = other_aux.T @ other_aux + rho * np.eye(kk)
    b = other_aux.T @ (v_aux + dual_v) + rho * (x + dual)
    return np.linalg.solve(a, b)


def kl_v_aux(v, v_bar):
    """admm.py:311 / ao_admm.py:90: closed-form minimiser of KL(v | v_aux) + 1/2 |v_aux - v_bar|^2."""
    return 0.5 * ((v_bar - 1.0) + np.sqrt((v_bar - 1.0) ** 2 + 4.0 * v))


def admm_run(v, k, w0, h0, *, rho=1, distance_type="eu", reg_w=(0, "nn"), reg_h=(0, "l2n"), min_iter=10,
             max_iter=100000, tol1=1e-3, tol2=1e-3,
             snap: Optional[Callable[[int], bool]] = None) -> Trace:
    """admm.py:233-345: Euclidean branch (:294-301) or Kullback-Leibler branch (:303-315, with the m x n
    auxiliary v_aux and its dual, both starting at zero :33-35); duals :320-321, objective :324."""
    if distance_type not in ("eu", "kl"):
        raise TypeError("Unknown loss type.")
    v = np.asarray(v, dtype=np.float64)
    w, h = np.array(w0, dtype=np.float64), np.array(h0, dtype=np.float64)
    w_aux, h_aux = w.copy(), h.copy()                     # admm.py:27-28
    dual_w, dual_h = np.zeros_like(w), np.zeros_like(h)   # admm.py:31-32
    v_aux, dual_v = np.zeros_like(v), np.zeros_like(v)    # admm.py:33
    hist = [objective(v, w @ h, distance_type)]
    out = Trace(w, h, -1, hist)
    for i in range(max_iter):
        if distance_type == "eu":
            h_aux = admm_aux(h, dual_h, w_aux, v, rho)
            w_aux = admm_aux(w.T, dual_w.T, h_aux.T, v.T, rho).T
        else:
            h_aux = admm_aux_kl(h, dual_h, w_aux, v_aux, dual_v, rho)
            w_aux = admm_aux_kl(w.T, dual_w.T, h_aux.T, v_aux.T, dual_v.T, rho).T
        h = prox(reg_h[1], h_aux, dual_h, rho, reg_h[0])
        w = prox(reg_w[1], w_aux.T, dual_w.T, rho, reg_w[0]).T      # admm.py:300-301
        if distance_type == "kl":
            wh_aux = w_aux @ h_aux
            v_aux = kl_v_aux(v, wh_aux - dual_v)                    # admm.py:310-311
            dual_v = dual_v + v_aux - wh_aux                        # admm.py:313
        dual_h = dual_h + h - h_aux
        dual_w = dual_w + w - w_aux
        hist.append(objective(v, w @ h, distance_type))
        out.i = i
        if snap is not None and snap(i):
            out.snaps.append((i, w.copy(), h.copy()))
        if i > min_iter:
            code = stop_test(hist[-1], hist[-2], tol1, tol2)
            if code:
                out.stop_code = code
                break
    out.w, out.h = w, h
    return out


# --------------------------------------------------------------------------- #
# AO-ADMM  (ao_admm.py)
# --------------------------------------------------------------------------- #
def inner_done(x, x_prev, aux, dual, tol=INNER_TOL) -> bool:
    """ao_admm.py:33-43 -- relative primal and dual residual; inf/NaN compare False."""
    with np.errstate(divide="ignore", invalid="ignore"):
        r = np.linalg.norm(x - aux) / np.linalg.norm(x)
        s = np.linalg.norm(x - x_prev) / np.linalg.norm(dual)
    return bool(r < tol and s < tol)


def ao_factor_update(y, a, x, dual, k, prox_kind, lam, admm_iter):
    """ao_admm.py:46-68 -- one least-squares sub-problem: Gram, rho = tr(G)/k,
    Cholesky once, then up to `admm_iter` (solve, prox, dual, residual test)."""
    g = a.T @ a
    rho = np.trace(g) / k
    cho = sla.cholesky(g + rho * np.eye(g.shape[0]), lower=True)
    aty = a.T @ y
    used = admm_iter
    for it in range(admm_iter):
        aux = sla.cho_solve((cho, True), aty + rho * (x + dual))
        x_prev = x
        x = prox(prox_kind, aux, dual, rho, lam)
        dual = dual + x - aux
        if inner_done(x, x_prev, aux, dual):
            used = it
            break
    return x, dual, used


def ao_factor_update_kl(v, v_aux, dual_v, a, x, dual, k, prox_kind, lam, admm_iter):
    """ao_admm.py:71-101 -- the Kullback-Leibler sub-problem: the right-hand side is rebuilt from
    v_aux + dual_v in every inner iteration, v_aux / dual_v follow the closed form of :89-94."""
    g = a.T @ a
    rho = np.trace(g) / k
    cho = sla.cholesky(g + rho * np.eye(g.shape[0]), lower=True)
    used = admm_iter
    for it in range(admm_iter):
        aux = sla.cho_solve((cho, True), a.T @ (v_aux + dual_v) + rho * (x + dual))
        x_prev = x
        x = prox(prox_kind, aux, dual, rho, lam)
        a_aux = a @ aux
        v_aux = kl_v_aux(v, a_aux - dual_v)
        dual = dual + x - aux
        dual_v = dual_v + v_aux - a_aux
        if inner_done(x, x_prev, aux, dual):
            used = it
            break
    return x, dual, v_aux, dual_v, used


def ao_admm_run(v, k, w0, h0, *, distance_type="eu", reg_w=(0, "nn"), reg_h=(0, "l2n"), min_iter=10,
                max_iter=100000, admm_iter=10, tol1=1e-3, tol2=1e-3,
                snap: Optional[Callable[[int], bool]] = None) -> Trace:
    """ao_admm.py:201-311: Euclidean branch (:261-270) or Kullback-Leibler branch (:272-285); duals
    (and, for kl, v_aux / dual_v, starting at zero :28-30) are warm-started across outer iterations.
    (`reg_h=(0,'l2n')`, the reference's own default, crashes the reference on NumPy >= 1.24 at
    ao_admm.py:128; here it is the lam==0 identity.)"""
    if distance_type not in ("eu", "kl"):
        raise TypeError("Unknown loss function type.")
    v = np.asarray(v, dtype=np.float64)
    w, h = np.array(w0, dtype=np.float64), np.array(h0, dtype=np.float64)
    dual_w, dual_h = np.zeros_like(w), np.zeros_like(h)
    v_aux, dual_v = np.zeros_like(v), np.zeros_like(v)
    hist = [objective(v, w @ h, distance_type)]
    out = Trace(w, h, -1, hist)
    for i in range(max_iter):
        if distance_type == "eu":
            h, dual_h, used_h = ao_factor_update(v, w, h, dual_h, k, reg_h[1], reg_h[0], admm_iter)
            wt, dwt, used_w = ao_factor_update(v.T, h.T, w.T, dual_w.T, k, reg_w[1], reg_w[0],
                                               admm_iter)
        else:
            h, dual_h, v_aux, dual_v, used_h = ao_factor_update_kl(v, v_aux, dual_v, w, h, dual_h, k,
                                                                   reg_h[1], reg_h[0], admm_iter)
            wt, dwt, vat, dvt, used_w = ao_factor_update_kl(v.T, v_aux.T, dual_v.T, h.T, w.T, dual_w.T, k,
                                                            reg_w[1], reg_w[0], admm_iter)
            v_aux, dual_v = vat.T, dvt.T
        w, dual_w = wt.T, dwt.T
        out.inner_breaks.append((used_h, used_w))
        hist.append(objective(v, w @ h, distance_type))
        out.i = i
        if snap is not None and snap(i):
            out.snaps.append((i, w.copy(), h.copy()))
        if i > min_iter:
            code = stop_test(hist[-1], hist[-2], tol1, tol2)
            if code:
                out.stop_code = code
                break
    out.w, out.h = w, h
    return out


# --------------------------------------------------------------------------- #
# synthetic workloads -- shared by tests and bench
# --------------------------------------------------------------------------- #
def round_to_bf16_grid(a: np.ndarray) -> np.ndarray:
    """Round float32 values to the nearest bfloat16 (ties to even), returned as float32."""
    u = np.ascontiguousarray(a, dtype=np.float32).view(np.uint32)
    r = ((u >> 16) & 1) + np.uint32(0x7FFF)
    return ((u + r) & np.uint32(0xFFFF0000)).view(np.float32)


def planted(m: int, n: int, k: int, seed: int = 0) -> np.ndarray:
    """planted(m,n,k,seed): low-rank + 1% noise, rounded to the bf16 grid so that a
    2-byte device copy of V is lossless.  float32 result."""
    rng = np.random.default_rng(seed)
    wt = rng.random((m, k), dtype=np.float32)
    ht = rng.random((k, n), dtype=np.float32)
    v = wt @ ht
    v += np.float32(0.01) * rng.random((m, n), dtype=np.float32)
    return round_to_bf16_grid(v)


def planted_with_zeros(m: int, n: int, k: int, seed: int = 0, frac: float = 0.3) -> np.ndarray:
    """planted() with a fraction `frac` of the entries, row 3 and column 5 set to exactly zero: the data
    for the 0 * log(0 / x) and x / 0 rules of utils.py:23-26 and for all-zero rows / columns."""
    v = planted(m, n, k, seed)
    mask = np.random.default_rng(seed + 7919).random((m, n)) < frac
    v[mask] = 0.0
    v[min(3, m - 1), :] = 0.0
    v[:, min(5, n - 1)] = 0.0
    return v


def seeded_init(m: int, n: int, k: int, seed: int):
    """Reference-order init from a private seed (does not disturb the global RNG)."""
    st = np.random.get_state()
    np.random.seed(seed)
    w, h = reference_init(m, n, k)
    np.random.set_state(st)
    return w, h
