"""CPU: the CUDA-core kernels that were written without a GPU at hand -- the l2n (Tikhonov) variant of the
solve / prox / dual kernel, the Tikhonov Gram, the element-wise v_aux / dual_v kernel of the Kullback-Leibler
branch -- run on host threads (tools/host_emu: kernel text verbatim from nmf_b200/csrc, one std::thread per
CUDA thread, barriers for __syncthreads / shuffles) against NumPy restatements of the reference lines.  The
GPU-validated nn / l1n kernel, the fixed-point ingest and the |max| kernel go through the same emulator as
its control (and, for the ingest, to pin the behaviour on SIGNED data the KL branch feeds it).
This checks indexing, synchronisation structure and arithmetic; tcgen05 / TMA / mbarrier kernels are out of its
reach and stay with the GPU tests."""
import ctypes as C
import os
import subprocess
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

from oracle import nmf_oracle as orc  # noqa: E402

EMU = os.path.join(ROOT, "tools", "host_emu")
CUDA_INC = "/usr/local/cuda/include"


@pytest.fixture(scope="module")
def emu(tmp_path_factory):
    if not os.path.exists(os.path.join(CUDA_INC, "cuda_bf16.h")):
        pytest.skip("CUDA headers not found")
    d = tmp_path_factory.mktemp("host_emu")
    subprocess.check_call([sys.executable, os.path.join(EMU, "extract.py"), str(d / "gen_kernels.h")])
    so = str(d / "libemu.so")
    subprocess.check_call(["g++", "-std=c++20", "-O1", "-fPIC", "-shared", "-pthread", "-I", CUDA_INC, "-I", EMU,
                           "-I", str(d), os.path.join(EMU, "harness.cpp"), "-o", so])
    return C.CDLL(so)


def ptr(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


def bf16_to_f32(u16):
    return (u16.astype(np.uint32) << 16).view(np.float32)


def f32_to_bf16(x):
    return (orc.round_to_bf16_grid(x).view(np.uint32) >> 16).astype(np.uint16)


def spd_inverse(kk, KP, rng, shift):
    a = rng.standard_normal((kk, kk))
    g = a @ a.T / kk + shift * np.eye(kk)
    out = np.eye(KP)
    out[:kk, :kk] = np.linalg.inv(g)
    return out


# ----------------------------------------------------------------------------------------------
def run_apply(emu, *, l2n, KP, k, L, Lp, prox, lam, rho, seed, use_part=True, want_aux=True):
    rng = np.random.default_rng(seed)
    slots = 3
    part = (rng.standard_normal((slots, KP, Lp)) * 5).astype(np.float32)
    c64_in = part.astype(np.float64).sum(0)
    c64 = np.zeros((KP, Lp)) if use_part else c64_in.copy()
    ainv = spd_inverse(k, KP, rng, rho)
    binv = spd_inverse(k, KP, rng, 1.0)
    x = np.abs(rng.standard_normal((KP, Lp))).astype(np.float32)
    u = rng.standard_normal((KP, Lp)).astype(np.float32) * 0.3
    x[k:], u[k:] = 0, 0
    x[:, L:], u[:, L:] = 0, 0
    x0, u0 = x.copy(), u.copy()
    aux_m = np.full((KP, Lp), np.nan, np.float32) if want_aux else None
    x_rpl = np.zeros((2, Lp, KP), np.uint16)
    x_cpl = np.zeros((3, 256, Lp), np.uint16)
    nblk = Lp // 64
    norms = np.zeros((nblk, 4))
    rho_a = np.array([rho])
    emu.emu_admm_apply(C.c_int(int(l2n)), KP, k, L, Lp, 0, ptr(part) if use_part else None, slots, ptr(c64),
                       ptr(ainv), ptr(binv), ptr(rho_a), C.c_float(lam), prox, ptr(x), ptr(u), ptr(aux_m), None,
                       ptr(x_cpl), ptr(x_rpl), ptr(norms), None)
    # NumPy restatement (admm.py:216-230 solve, :126-156 prox, :320-321 dual; ao_admm.py:36-43 norms)
    b = c64_in + rho * (x0.astype(np.float64) + u0.astype(np.float64))
    aux = ainv @ b
    live = np.zeros((KP, Lp), bool)
    live[:k, :L] = True
    aux = np.where(live, aux, 0.0)
    if l2n:
        z = binv @ np.where(live, aux - u0, 0.0)
    else:
        z = aux - u0 - (lam / rho if prox == 1 else 0.0)
    xn = np.where(live & (z > 0), z, 0.0)
    un = np.where(live, u0 + xn - aux, 0.0)
    return dict(x=x, u=u, aux=aux_m, c64=c64, x_rpl=x_rpl, x_cpl=x_cpl, norms=norms,
                want_x=xn, want_u=un, want_aux=aux, want_c64=c64_in, x0=x0)


def check_apply(r, KP, Lp):
    tol = dict(rtol=2e-6, atol=2e-6)
    np.testing.assert_allclose(r["x"], r["want_x"], **tol)
    np.testing.assert_allclose(r["u"], r["want_u"], **tol)
    if r["aux"] is not None:
        np.testing.assert_allclose(r["aux"], r["want_aux"], **tol)
    np.testing.assert_allclose(r["c64"], r["want_c64"], rtol=1e-14, atol=1e-12)
    # planes are the bf16 split of the fp32 x that was stored: rplanes [2][Lp][KP], cplanes [3][256][Lp]
    hi = orc.round_to_bf16_grid(r["x"])
    lo = orc.round_to_bf16_grid(r["x"] - hi)
    lo2 = orc.round_to_bf16_grid(r["x"] - hi - lo)
    assert np.array_equal(bf16_to_f32(r["x_rpl"][0]), hi.T) and np.array_equal(bf16_to_f32(r["x_rpl"][1]), lo.T)
    for q, pl in enumerate((hi, lo, lo2)):
        assert np.array_equal(bf16_to_f32(r["x_cpl"][q, :KP]), pl)
    xs, au, us = r["x"].astype(np.float64), r["want_aux"], r["u"].astype(np.float64)
    want = np.array([[((xs - au)[:, s:s + 64] ** 2).sum(), (xs[:, s:s + 64] ** 2).sum(),
                      ((xs - r["x0"])[:, s:s + 64] ** 2).sum(), (us[:, s:s + 64] ** 2).sum()]
                     for s in range(0, Lp, 64)])
    np.testing.assert_allclose(r["norms"], want, rtol=1e-5, atol=1e-9)


@pytest.mark.parametrize("KP,k", [(64, 40), (128, 128), (256, 200)])
@pytest.mark.parametrize("prox,lam", [(0, 0.0), (1, 0.7)])
def test_control_nn_l1n_kernel_matches_numpy(emu, KP, k, prox, lam):
    """The kernel the GPU tests validated, through the emulator: if this fails the emulator is wrong."""
    r = run_apply(emu, l2n=False, KP=KP, k=k, L=150, Lp=256, prox=prox, lam=lam, rho=3.5, seed=KP + prox)
    check_apply(r, KP, 256)
    assert r["x"][:, 150:].max() == 0 and (k == KP or r["x"][k:].max() == 0)      # padding stays zero


@pytest.mark.parametrize("KP,k", [(64, 12), (64, 64), (128, 100)])
def test_l2n_kernel_matches_numpy(emu, KP, k):
    """admm.py:141-156: x = max(Binv (aux - u), 0) along the component axis, dual with that x."""
    r = run_apply(emu, l2n=True, KP=KP, k=k, L=200, Lp=256, prox=2, lam=0.4, rho=2.0, seed=7 * KP + k)
    check_apply(r, KP, 256)
    r = run_apply(emu, l2n=True, KP=KP, k=k, L=64, Lp=128, prox=2, lam=0.4, rho=50.0, seed=k, use_part=False,
                  want_aux=False)
    check_apply(r, KP, 128)


def test_l2n_pipeline_matches_the_oracle_prox(emu):
    """tikhonov_gram_kernel -> (host inverse standing in for the GPU-validated Cholesky kernel) -> l2n kernel
    against oracle.prox('l2n', ...), the restatement pinned by the `admm_l2n` golden."""
    KP, k, L, Lp, lam, rho = 64, 12, 100, 128, 0.5, 50.0
    rng = np.random.default_rng(5)
    tik = np.zeros((KP, KP))
    emu.emu_tikhonov_gram(k, KP, C.c_double(lam), ptr(np.array([rho])), ptr(tik))
    want_sys = orc.tikhonov_system(k, lam, rho)
    np.testing.assert_allclose(tik[:k, :k] + np.eye(k), want_sys, rtol=1e-14)
    assert not tik[k:].any() and not tik[:, k:].any()
    for kk in (1, 2, 3):                                     # corner cases of the pentadiagonal pattern
        t = np.zeros((KP, KP))
        emu.emu_tikhonov_gram(kk, KP, C.c_double(lam), ptr(np.array([rho])), ptr(t))
        np.testing.assert_allclose(t[:kk, :kk] + np.eye(kk), orc.tikhonov_system(kk, lam, rho), rtol=1e-14)
    binv = np.eye(KP)
    binv[:k, :k] = np.linalg.inv(tik[:k, :k] + np.eye(k))
    ainv = np.eye(KP)                                        # identity solve: aux = b / 1 with rho folded below
    c = rng.standard_normal((KP, Lp))
    c[k:], c[:, L:] = 0, 0
    x = np.zeros((KP, Lp), np.float32)
    u = (rng.standard_normal((KP, Lp)) * 0.2).astype(np.float32)
    u[k:], u[:, L:] = 0, 0
    u0 = u.copy()
    c64 = c - rho * u0.astype(np.float64)                    # so that b = c64 + rho (x + u) = c exactly
    aux_m = np.zeros((KP, Lp), np.float32)
    x_rpl = np.zeros((2, Lp, KP), np.uint16)
    emu.emu_admm_apply(1, KP, k, L, Lp, 0, None, 0, ptr(c64), ptr(ainv), ptr(binv), ptr(np.array([rho])),
                       C.c_float(lam), 2, ptr(x), ptr(u), ptr(aux_m), None, None, ptr(x_rpl), None, None)
    want = orc.prox("l2n", c[:k, :L], u0[:k, :L].astype(np.float64), rho, lam)
    np.testing.assert_allclose(x[:k, :L], want, rtol=2e-6, atol=2e-6)
    np.testing.assert_allclose(u[:k, :L], u0[:k, :L] + want - c[:k, :L], rtol=2e-6, atol=2e-6)


# ----------------------------------------------------------------------------------------------
def blocked(v, rows_p):
    """[col / 64][rows_p][64] layout of kernels.cuh:blocked_index for a (rows, cols) array."""
    rows, cols = v.shape
    nb = (cols + 63) // 64
    out = np.zeros((nb, rows_p, 64), v.dtype)
    for b in range(nb):
        w = min(64, cols - b * 64)
        out[b, :rows, :w] = v[:, b * 64:b * 64 + w]
    return out


@pytest.mark.parametrize("nv", [1, 2])
def test_kl_vaux_kernel_matches_the_oracle(emu, nv):
    """admm.py:310-313 / ao_admm.py:89-94 incl. a column tail, two data planes, the cancellation-prone
    v_bar << 1 regime, v == 0, and the done flag."""
    rows, cols, rows_p, ld = 9, 1100, 64, 1104
    rng = np.random.default_rng(11 + nv)
    v = rng.random((rows, cols)).astype(np.float32) * 3
    v[rng.random((rows, cols)) < 0.2] = 0.0
    if nv == 1:
        v = orc.round_to_bf16_grid(v)
    hi = orc.round_to_bf16_grid(v)
    lo = orc.round_to_bf16_grid(v - hi)
    planes = np.stack([blocked(f32_to_bf16(hi), rows_p), blocked(f32_to_bf16(lo), rows_p)])
    v_dev = (hi + lo).astype(np.float64) if nv == 2 else hi.astype(np.float64)
    prod = np.zeros((rows, ld), np.float32)
    prod[:, :cols] = rng.standard_normal((rows, cols)) * 4
    prod[0, :200] = -1e4 * rng.random(200)                  # t = v_bar - 1 very negative: root cancels
    prod[1, :200] = 1e4 * rng.random(200)
    dv = np.zeros((rows, ld), np.float32)
    dv[:, :cols] = rng.standard_normal((rows, cols))
    d_out = np.full((rows, ld), 7.0, np.float32)
    p0, dv0 = prod.copy(), dv.copy()
    flag = np.array([1], np.int32)
    emu.emu_kl_vaux(ptr(prod), ptr(d_out), ptr(dv), C.c_long(ld), rows, cols, ptr(planes), C.c_long(planes[0].size),
                    rows_p, nv, ptr(flag))
    assert np.array_equal(dv, dv0) and np.all(d_out == 7.0)                       # frozen by the flag
    emu.emu_kl_vaux(ptr(prod), ptr(d_out), ptr(dv), C.c_long(ld), rows, cols, ptr(planes), C.c_long(planes[0].size),
                    rows_p, nv, None)
    P, d0 = p0[:, :cols].astype(np.float64), dv0[:, :cols].astype(np.float64)
    t = P - d0 - 1.0
    root = np.sqrt(t * t + 4 * v_dev)
    va = np.where(t >= 0, 0.5 * (t + root), 2 * v_dev / np.where(root - t > 0, root - t, 1.0))   # == kl_v_aux, stably
    naive = orc.kl_v_aux(v_dev, P - d0)
    ok = t > -50                                                                   # where the naive fp64 form is accurate
    np.testing.assert_allclose(va[ok], naive[ok], rtol=1e-9, atol=1e-12)
    dn = d0 + va - P
    np.testing.assert_allclose(dv[:, :cols], dn, rtol=1e-6, atol=1e-6)
    np.testing.assert_allclose(d_out[:, :cols], va + dn.astype(np.float32), rtol=1e-6, atol=1e-6)
    assert np.all(d_out[:, cols:] == 7.0) and np.array_equal(prod, p0)            # pitch padding / input untouched
    assert np.all(va >= 0) and np.all(np.isfinite(d_out))
    # in place (ADMM passes the product buffer as the output)
    prod2, dv2 = p0.copy(), dv0.copy()
    emu.emu_kl_vaux(ptr(prod2), ptr(prod2), ptr(dv2), C.c_long(ld), rows, cols, ptr(planes),
                    C.c_long(planes[0].size), rows_p, nv, None)
    assert np.array_equal(prod2[:, :cols], d_out[:, :cols]) and np.array_equal(dv2, dv)


def test_fixed_point_ingest_of_signed_data(emu):
    """The KL branch re-ingests D = v_aux + dual_v (signed) every iteration: plane 0 is a multiple of
    2^(E-6) of at most 128 units, E from |max| of the row (v_rows copy) / column (v_cols copy); planes 0 + 1
    reproduce D to 2^-16 of that maximum; the two copies are transposes of each other in the blocked layout."""
    m, n = 70, 130
    rng = np.random.default_rng(3)
    d = (rng.standard_normal((m, n)) * np.exp(rng.standard_normal((m, 1)) * 2)).astype(np.float32)
    d[5] = 0.0
    rows_p, cols_p = 128, 256                               # padded rows of the two copies
    rowmax, colmax = np.zeros(rows_p, np.uint32), np.zeros(cols_p, np.uint32)
    emu.emu_data_absmax(ptr(d), C.c_long(n), 0, m, n, ptr(rowmax), ptr(colmax))
    assert np.array_equal(rowmax[:m].view(np.float32), np.abs(d).max(1))
    assert np.array_equal(colmax[:n].view(np.float32), np.abs(d).max(0))
    nbr, nbc = (n + 63) // 64, (m + 63) // 64
    rowp = np.zeros((2, nbr, rows_p, 64), np.uint16)
    colp = np.zeros((2, nbc, cols_p, 64), np.uint16)
    stats = np.array([0xFFFFFFFF, 0, 0, 0], np.uint32)
    emu.emu_ingest(ptr(d), C.c_long(n), 0, m, n, ptr(rowp), C.c_long(rowp[0].size), rows_p, ptr(colp),
                   C.c_long(colp[0].size), cols_p, 2, ptr(stats), ptr(rowmax), ptr(colmax), 0, m, 0, n)
    assert stats[2] == 0                                     # finite

    def unblock(pl, rows, cols):
        return np.concatenate([pl[b, :rows] for b in range(pl.shape[0])], axis=1)[:, :cols]

    for planes, rows, cols, ref, mx in ((rowp, m, n, d, np.abs(d).max(1)), (colp, n, m, d.T, np.abs(d).max(0))):
        hi = bf16_to_f32(unblock(planes[0], rows, cols))
        lo = bf16_to_f32(unblock(planes[1], rows, cols))
        e = np.frexp(np.maximum(mx, 1e-30))[1] - 1          # exponent E of the row maximum
        unit = np.ldexp(1.0, np.maximum(e, 16 - 127) - 6)[:, None]
        live = mx > 0
        q = hi[live] / unit[live]
        assert np.array_equal(q, np.round(q)) and np.abs(q).max() <= 128
        err = np.abs(hi.astype(np.float64) + lo - ref)
        assert np.all(err[live] <= np.ldexp(mx[live], -15)[:, None])
        assert not hi[~live].any() and not lo[~live].any()


def test_inner_flag_kernel(emu):
    """ao_admm.py:33-43: r = |x-aux|/|x| < 1e-2 and s = |x-x_prev|/|u| < 1e-2; NaN compares false."""
    def run(parts, flags):
        parts = np.ascontiguousarray(parts, dtype=np.float64)
        flags = np.array(flags, np.int32)
        emu.emu_inner_flag(ptr(parts), parts.shape[0], ptr(flags), 2)
        return flags
    small = np.tile([1e-6, 1.0, 1e-6, 1.0], (5, 1))          # r = s = 1e-3
    assert run(small, [0, 0, 0, 0, 0, 0, 0, 0])[1] == 1
    big = small.copy()
    big[:, 0] = 1.0                                          # r = 1
    assert run(big, [0] * 8)[1] == 0
    nan = small.copy()
    nan[:, 3] = 0.0
    nan[:, 2] = 0.0                                          # s = 0/0
    assert run(nan, [0] * 8)[1] == 0


# ----------------------------------------------------------------------------------------------
# tests/softdev.py runs the orchestration tests on NumPy restatements of the three data-path kernels (a host
# thread per CUDA thread is too slow for a data term that is re-ingested every iteration): pin those
# restatements to the kernel text, bit for bit.
# ----------------------------------------------------------------------------------------------
def _data_struct(planes, rows, cols, rows_p, ld, nv, fixed, origin):
    from nmf_b200._lib import Data
    return Data(planes.ctypes.data, rows, cols, rows_p, ld, nv, 1, fixed, origin)


@pytest.mark.parametrize("fixed", [0, 1])
@pytest.mark.parametrize("window", [None, (64, 70, 128, 60)])        # (row_lo, row_n, col_lo, col_n): a shard
def test_numpy_ingest_equals_the_kernel(emu, fixed, window):
    import softdev
    m, n = 150, 200
    rng = np.random.default_rng(17 + fixed)
    v = (rng.standard_normal((m, n)) * np.exp(rng.standard_normal((m, 1)))).astype(np.float32)
    if not fixed:
        v = np.abs(v)
    v[7] = 0
    row_lo, row_n, col_lo, col_n = window or (0, m, 0, n)
    ld_r, ld_c = 256, 256                                   # round_up(n, 256), round_up(m, 256)
    rp_r, rp_c = 256, 256
    nv = 2
    out = []
    for impl in ("kernel", "numpy"):
        rowp = np.zeros((nv, ld_r // 64, rp_r, 64), np.uint16)
        colp = np.zeros((nv, ld_c // 64, rp_c, 64), np.uint16)
        stats = np.array([0xFFFFFFFF, 0, 0, 0], np.uint32)
        rowmax, colmax = np.zeros(256, np.uint32), np.zeros(256, np.uint32)
        for r0, r1 in ((0, 64), (64, 150)):                 # two chunks, like Engine.upload_data
            chunk = np.ascontiguousarray(v[r0:r1])
            if impl == "kernel":
                emu.emu_data_absmax(ptr(chunk), C.c_long(n), r0, r1 - r0, n, ptr(rowmax), ptr(colmax))
            else:
                softdev.np_absmax(chunk, r0, rowmax, colmax)
        for r0, r1 in ((0, 64), (64, 150)):
            chunk = np.ascontiguousarray(v[r0:r1])
            if impl == "kernel":
                emu.emu_ingest(ptr(chunk), C.c_long(n), r0, r1 - r0, n, ptr(rowp), C.c_long(rowp[0].size), rp_r,
                               ptr(colp), C.c_long(colp[0].size), rp_c, nv, ptr(stats),
                               ptr(rowmax) if fixed else None, ptr(colmax) if fixed else None,
                               row_lo, row_n, col_lo, col_n)
            else:
                vr = _data_struct(rowp, row_n, n, rp_r, ld_r, nv, fixed, row_lo)
                vc = _data_struct(colp, col_n, m, rp_c, ld_c, nv, fixed, col_lo)
                softdev.np_ingest(chunk, r0, vr, vc, stats, rowmax if fixed else None, colmax if fixed else None)
        out.append((rowp, colp, stats, rowmax, colmax))
    for a, b in zip(*out):
        assert np.array_equal(a, b)
    rowp, colp, stats = out[1][:3]
    assert rowp.any() and colp.any() and stats[2] == 0
    # and the decode the soft device uses gives back V (to the two planes' precision) for the rank's shard
    got = softdev.SoftLib.decode(_data_struct(rowp, row_n, n, rp_r, ld_r, nv, fixed, row_lo))
    np.testing.assert_allclose(got, v[row_lo:row_lo + row_n], rtol=2 ** -15, atol=np.abs(v).max() * 2 ** -15)
    got = softdev.SoftLib.decode(_data_struct(colp, col_n, m, rp_c, ld_c, nv, fixed, col_lo))
    np.testing.assert_allclose(got, v[:, col_lo:col_lo + col_n].T, rtol=2 ** -15, atol=np.abs(v).max() * 2 ** -15)


def test_numpy_kl_vaux_equals_the_kernel(emu):
    import softdev
    rows, cols, rows_p, ld = 5, 300, 64, 320
    rng = np.random.default_rng(23)
    v = orc.round_to_bf16_grid(rng.random((rows, cols)).astype(np.float32) * 2)
    planes = blocked(f32_to_bf16(v), rows_p)[None]
    prod = np.zeros((rows, ld), np.float32)
    prod[:, :cols] = rng.standard_normal((rows, cols)) * 30
    dv = np.zeros((rows, ld), np.float32)
    dv[:, :cols] = rng.standard_normal((rows, cols))
    a = [prod.copy(), np.zeros_like(prod), dv.copy()]
    b = [prod.copy(), np.zeros_like(prod), dv.copy()]
    emu.emu_kl_vaux(ptr(a[0]), ptr(a[1]), ptr(a[2]), C.c_long(ld), rows, cols, ptr(planes), C.c_long(planes[0].size),
                    rows_p, 1, None)
    softdev.np_kl_vaux(b[0][:, :cols], b[1][:, :cols], b[2][:, :cols], v.astype(np.float64))
    for x, y in zip(a, b):                                   # fma(t, t, 4v) vs t*t + 4v: at most an fp32 ulp apart
        np.testing.assert_allclose(x, y, rtol=3e-7, atol=1e-7)


# ----------------------------------------------------------------------------------------------
# The CUDA-core kernels of the MUR paths (GPU-validated in round 1), through the emulator against the
# reference formulas: the fused multiplicative updates, the row sums, the Gram slot reduction and the
# CUDA-core objective.
# ----------------------------------------------------------------------------------------------
def _factor(KP, k, L, Lp, rng):
    x = np.zeros((KP, Lp), np.float32)
    x[:k, :L] = np.abs(rng.standard_normal((k, L))) + 0.01
    return x


def _check_planes(x, cpl, rpl, KP):
    hi = orc.round_to_bf16_grid(x)
    lo = orc.round_to_bf16_grid(x - hi)
    assert np.array_equal(bf16_to_f32(cpl[0, :KP]), hi) and np.array_equal(bf16_to_f32(cpl[1, :KP]), lo)
    assert np.array_equal(bf16_to_f32(rpl[0]), hi.T) and np.array_equal(bf16_to_f32(rpl[1]), lo.T)


@pytest.mark.parametrize("KP,k", [(64, 10), (128, 128), (256, 200)])
def test_mur_eu_update_kernel_is_the_lee_seung_rule(emu, KP, k):
    """mur.py:29 / :45: x <- x * num / (G x + lambda x + 1e-9), G the Gram of the other factor."""
    L, Lp, slots, lam = 150, 256, 2, 0.05
    rng = np.random.default_rng(KP + k)
    x = _factor(KP, k, L, Lp, rng)
    y = np.zeros((KP, 300), np.float32)
    y[:k] = np.abs(rng.standard_normal((k, 300)))
    gram = (y.astype(np.float64) @ y.T.astype(np.float64)).astype(np.float32)
    part = np.zeros((slots, KP, Lp), np.float32)
    part[:, :k, :L] = np.abs(rng.standard_normal((slots, k, L))) * 20
    x0 = x.copy()
    cpl, rpl = np.zeros((3, 256, Lp), np.uint16), np.zeros((2, Lp, KP), np.uint16)
    emu.emu_mur_eu_update(KP, ptr(part), slots, ptr(gram), ptr(x), ptr(cpl), ptr(rpl), k, L, Lp, Lp, C.c_float(lam))
    xd, num = x0.astype(np.float64), part.astype(np.float64).sum(0)
    want = xd * num / (gram.astype(np.float64) @ xd + lam * xd + 1e-9)
    np.testing.assert_allclose(x[:k, :L], want[:k, :L], rtol=2e-5)
    assert not x[k:].any() and not x[:, L:].any()
    _check_planes(x, cpl, rpl, KP)


@pytest.mark.parametrize("KP,k", [(64, 7), (128, 100)])
@pytest.mark.parametrize("lam", [0.0, 0.3])
def test_mur_kl_update_kernel_is_the_reference_rule(emu, KP, k, lam):
    """mur.py:25-27 / :41-43: a = x * num, b = row sums of the other factor, x <- 2a / (b + sqrt(b^2 + 4 lambda a))."""
    L, Lp, slots = 130, 192, 3
    rng = np.random.default_rng(KP + k)
    x = _factor(KP, k, L, Lp, rng)
    other = _factor(KP, k, 333, 512, rng)
    bsum = np.zeros(KP)
    emu.emu_factor_rowsum(ptr(other), KP, 333, 512, ptr(bsum))
    np.testing.assert_allclose(bsum, other.astype(np.float64).sum(1), rtol=1e-13)
    part = np.zeros((slots, KP, Lp), np.float32)
    part[:, :k, :L] = np.abs(rng.standard_normal((slots, k, L)))
    x0 = x.copy()
    cpl, rpl = np.zeros((3, 256, Lp), np.uint16), np.zeros((2, Lp, KP), np.uint16)
    emu.emu_mur_kl_update(KP, ptr(part), slots, ptr(bsum), ptr(x), ptr(cpl), ptr(rpl), k, L, Lp, Lp, C.c_float(lam))
    a = x0.astype(np.float64) * part.astype(np.float64).sum(0)
    b = bsum[:, None]
    with np.errstate(all="ignore"):
        want = 2 * a / (b + np.sqrt(b * b + 4 * lam * a))
    np.testing.assert_allclose(x[:k, :L], want[:k, :L], rtol=2e-5)
    assert not x[k:].any() and not x[:, L:].any()
    _check_planes(x, cpl, rpl, KP)


def test_gram_reduce_and_cuda_core_objective(emu):
    KP, Lp, slots = 64, 256, 5
    rng = np.random.default_rng(9)
    part = rng.standard_normal((slots, KP, Lp)).astype(np.float32)
    g32, g64 = np.zeros((KP, KP), np.float32), np.zeros((KP, KP))
    emu.emu_gram_reduce(ptr(part), slots, KP, Lp, ptr(g32), ptr(g64))
    want = part[:, :, :KP].astype(np.float64).sum(0)
    assert np.array_equal(g64, want) and np.array_equal(g32, want.astype(np.float32))
    # utils.py:29 on a ragged 100 x 150 problem with two V planes
    m, n, k, Mp, Np = 100, 150, 9, 256, 256
    v = (rng.random((m, n)) * 3).astype(np.float32)
    hi = orc.round_to_bf16_grid(v)
    lo = orc.round_to_bf16_grid(v - hi)
    planes = np.stack([blocked(f32_to_bf16(hi), Mp), blocked(f32_to_bf16(lo), Mp)])
    wt = np.zeros((64, Mp), np.float32)
    wt[:k, :m] = rng.random((k, m))
    ht = np.zeros((64, Np), np.float32)
    ht[:k, :n] = rng.random((k, n))
    nparts = ((n + 63) // 64) * ((m + 63) // 64)
    partial = np.zeros(nparts)
    emu.emu_objective_simt(ptr(planes), C.c_long(planes[0].size), Mp, 2, ptr(wt), Mp, ptr(ht), Np, k, m, n, ptr(partial))
    want = orc.objective((hi + lo).astype(np.float64), wt[:k, :m].T.astype(np.float64) @ ht[:k, :n].astype(np.float64))
    assert abs(0.5 * partial.sum() - want) / want < 1e-5      # sum_partials_kernel applies the 1/2 of utils.py:29


@pytest.mark.parametrize("k,KP", [(10, 64), (64, 64), (100, 128), (200, 256)])
def test_chol_inverse_kernel(emu, k, KP):
    """admm.py:220,230 / ao_admm.py:54-55,59: (G + rho I)^-1 by the fp64 shared-memory Cholesky + substitution
    kernel (1024 threads, warp-shuffle broadcasts), rho given or tr(G)/k, and the not-PD flag."""
    rng = np.random.default_rng(k)
    a = rng.standard_normal((k, 3 * k))
    g = np.zeros((KP, KP))
    g[:k, :k] = a @ a.T
    for rho_mode, rho_in in ((0, 2.5), (1, 0.0)):
        ainv, rho_out, flags = np.zeros((KP, KP)), np.zeros(4), np.zeros(8, np.int32)
        scratch = np.zeros((KP, KP + 1))
        emu.emu_chol_inverse(KP, ptr(g), k, rho_mode, C.c_double(rho_in), ptr(scratch), ptr(ainv), ptr(rho_out),
                             ptr(flags))
        rho = rho_in if rho_mode == 0 else np.trace(g[:k, :k]) / k
        assert rho_out[0] == pytest.approx(rho, rel=1e-15) and flags[0] == 0
        want = np.linalg.inv(g[:k, :k] + rho * np.eye(k))
        np.testing.assert_allclose(ainv[:k, :k], want, rtol=1e-9, atol=1e-12 * np.abs(want).max())
        assert np.allclose(ainv[:k, :k], ainv[:k, :k].T, rtol=0, atol=1e-12 * np.abs(want).max())   # the apply kernels read it as symmetric
    bad = g.copy()
    bad[:k, :k] -= (np.linalg.eigvalsh(g[:k, :k]).max() + 10) * np.eye(k)      # negative definite even with rho
    ainv, rho_out, flags = np.zeros((KP, KP)), np.zeros(4), np.zeros(8, np.int32)
    emu.emu_chol_inverse(KP, ptr(bad), k, 0, C.c_double(1.0), ptr(np.zeros((KP, KP + 1))), ptr(ainv), ptr(rho_out),
                         ptr(flags))
    assert flags[0] == 1                                      # scipy / numpy raise LinAlgError here


def test_fixed_planes_kernel_on_signed_factors(emu):
    """Exact-mode operand planes of a (signed) ADMM auxiliary: plane 0 an integer of at most 128 units of
    2^(E-6), E from the component row's |max|; planes 1-2 floating remainders; together x to ~2^-22 of that max."""
    KP, Lp, k, L = 64, 512, 20, 400
    rng = np.random.default_rng(4)
    x = np.zeros((KP, Lp), np.float32)
    x[:k, :L] = rng.standard_normal((k, L)) * np.exp(rng.standard_normal((k, 1)) * 3)
    cpl = np.zeros((3, 256, Lp), np.uint16)
    emu.emu_fixed_planes(ptr(x), KP, Lp, ptr(cpl), C.c_long(256 * Lp))
    p = bf16_to_f32(cpl[:, :KP]).astype(np.float64)
    mx = np.abs(x).max(1)
    live = mx > 0
    unit = np.ldexp(1.0, np.frexp(np.maximum(mx, 1e-30))[1] - 1 - 6)[:, None]
    q = p[0][live] / unit[live]
    assert np.array_equal(q, np.round(q)) and np.abs(q).max() <= 128
    err = np.abs(p.sum(0) - x)
    assert np.all(err[live] <= np.ldexp(mx[live], -21)[:, None])
    assert not p[:, ~live].any()


# ----------------------------------------------------------------------------------------------
# Round 2: the plane refresh that adopts the exchanged windows of a sharded run.
# ----------------------------------------------------------------------------------------------
@pytest.mark.parametrize("KP,k", [(64, 10), (128, 128)])
def test_fp64_gram_kernels(emu, KP, k):
    """gram_fp64_launch (admm.cu; the DMMA emulated lane by lane): G equals the float64 Gram of the operand the three
    bf16 planes hold -- the operand the numerator GEMM contracts --, symmetric, zero in the padded components."""
    Lp = 256
    rng = np.random.default_rng(KP)
    x = _factor(KP, k, Lp - 60, Lp, rng)
    p0 = orc.round_to_bf16_grid(x)
    p1 = orc.round_to_bf16_grid(x - p0)
    p2 = orc.round_to_bf16_grid(x - p0 - p1)
    cpl = np.zeros((3, 256, Lp), np.uint16)
    cpl[0, :KP], cpl[1, :KP], cpl[2, :KP] = f32_to_bf16(p0), f32_to_bf16(p1), f32_to_bf16(p2)
    operand = p0.astype(np.float64) + p1 + p2
    want = operand @ operand.T
    for rows_per in (96, 256):
        nsplit = (Lp + rows_per - 1) // rows_per
        scratch = np.full((nsplit, KP, KP), np.nan)
        g32, g64 = np.full((KP, KP), np.nan, np.float32), np.full((KP, KP), np.nan)
        emu.emu_gram_fp64(ptr(cpl), KP, Lp, rows_per, nsplit, ptr(scratch), ptr(g32), ptr(g64))
        np.testing.assert_allclose(g64, want, rtol=1e-13, atol=1e-13 * np.abs(want).max())
        assert np.array_equal(g64, g64.T) and not g64[k:].any() and np.array_equal(g32, g64.astype(np.float32))


@pytest.mark.parametrize("KP", [64, 256])
def test_refresh_planes_kernel_adopts_windows_and_skips_the_own_one(emu, KP):
    Lp, lo, hi = 768, 256, 512
    rng = np.random.default_rng(KP)
    xchg = _factor(KP, KP - 3, Lp, Lp, rng)
    master = np.full((KP, Lp), -1.0, np.float32)
    cpl, rpl = np.full((3, 256, Lp), 7, np.uint16), np.full((2, Lp, KP), 7, np.uint16)
    emu.emu_refresh_planes(ptr(xchg), ptr(master), KP, Lp, ptr(cpl), ptr(rpl), lo, hi)
    keep = np.r_[0:lo, hi:Lp]
    assert np.array_equal(master[:, keep], xchg[:, keep]) and (master[:, lo:hi] == -1.0).all()
    h = orc.round_to_bf16_grid(xchg)
    l1 = orc.round_to_bf16_grid(xchg - h)
    l2 = orc.round_to_bf16_grid(xchg - h - l1)
    for pl, want in enumerate((h, l1, l2)):
        assert np.array_equal(bf16_to_f32(cpl[pl, :KP][:, keep]), want[:, keep]) and (cpl[pl, :KP, lo:hi] == 7).all()
    assert np.array_equal(bf16_to_f32(rpl[0][keep]), h.T[keep]) and np.array_equal(bf16_to_f32(rpl[1][keep]), l1.T[keep])
    assert (rpl[:, lo:hi] == 7).all()
    emu.emu_refresh_planes(ptr(xchg), None, KP, Lp, ptr(cpl), ptr(rpl), 0, 0)          # nmfb_refresh_planes
    assert np.array_equal(bf16_to_f32(cpl[0, :KP]), h) and np.array_equal(bf16_to_f32(rpl[1]), l1.T)
