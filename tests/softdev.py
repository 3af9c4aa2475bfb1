"""Test infrastructure: a host stand-in for libnmf_b200.so + the CUDA device, so that the ORCHESTRATION above
the C ABI -- Engine, steppers, host loop, drivers, the sharded exchange (gloo) -- runs on the CPU through the
same code the product runs on a B200.

`SoftLib` implements the C-ABI entry points Engine calls, at the semantic level of include/nmf_b200.h, on the
raw memory of the (CPU) torch tensors Engine allocated:
  * the data path is real kernel code: ingest / |max| / kl_vaux run through tools/host_emu (kernel text verbatim
    from nmf_b200/csrc on host threads), so V and the KL data term really live in K-blocked bf16 planes and
    every contraction decodes them again;
  * the contractions, solves and updates are NumPy fp64 restatements (the oracle's update rules on the rank's
    window), results rounded to the fp32 the device stores;
  * bf16 factor planes are not materialised; instead their FRESHNESS is tracked (which master content they were
    derived from, floating or fixed-point form) and every consumer asserts the planes it would read are current.
    A missing nmfb_refresh_planes / nmfb_fixed_planes / all-gather in the host code fails the test.
It is not a performance model and not a product path (nothing under nmf_b200/ imports it).
"""
from __future__ import annotations

import contextlib
import ctypes as C
import os
import subprocess
import sys
import tempfile

import numpy as np
import scipy.linalg as sla
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

from nmf_b200.engine import Engine  # noqa: E402
from oracle import nmf_oracle as orc  # noqa: E402

EMU_DIR = os.path.join(ROOT, "tools", "host_emu")
CUDA_INC = "/usr/local/cuda/include"
_emu = None


def emu_lib():
    """tools/host_emu built once per process tree (the .so is cached next to the system temp dir)."""
    global _emu
    if _emu is None:
        d = os.path.join(tempfile.gettempdir(), f"nmfb_host_emu_{os.getuid()}")
        os.makedirs(d, exist_ok=True)
        so = os.path.join(d, "libemu.so")
        srcs = [os.path.join(EMU_DIR, f) for f in ("harness.cpp", "cuda_on_host.h", "extract.py")]
        srcs += [os.path.join(ROOT, "nmf_b200", "csrc", f) for f in ("admm.cu", "admm.cuh", "elementwise.cu", "kl.cu",
                                                                     "ptx.cuh", "kernels.cuh")]
        if not os.path.exists(so) or any(os.path.getmtime(s) > os.path.getmtime(so) for s in srcs):
            tmp = so + f".{os.getpid()}"
            subprocess.check_call([sys.executable, os.path.join(EMU_DIR, "extract.py"),
                                   os.path.join(d, f"gen_kernels.h")])
            subprocess.check_call(["g++", "-std=c++20", "-O1", "-fPIC", "-shared", "-pthread", "-I", CUDA_INC,
                                   "-I", EMU_DIR, "-I", d, os.path.join(EMU_DIR, "harness.cpp"), "-o", tmp])
            os.replace(tmp, so)
        _emu = C.CDLL(so)
    return _emu


def _obj(a):
    """The ctypes struct behind C.byref(x)."""
    return a._obj if hasattr(a, "_obj") else a


def mem(ptr, dtype, count):
    """Writable NumPy view of `count` elements at a raw address."""
    if isinstance(ptr, C.c_void_p):
        ptr = ptr.value
    dt = np.dtype(dtype)
    buf = (C.c_char * (dt.itemsize * int(count))).from_address(int(ptr))
    return np.frombuffer(buf, dtype=dt)



# ----------------------------------------------------------------------------------------------
# NumPy restatements of the data-path kernels (elementwise.cu: ingest_data_kernel, data_absmax_kernel,
# kl_vaux_kernel).  The emulated kernels are ~1000x slower than NumPy (a host thread per CUDA thread), too slow
# for the KL branch that re-ingests its data term every iteration, so the soft device uses these; they are
# checked BIT FOR BIT against the emulated kernel text in tests/test_host_emulation.py.
# ----------------------------------------------------------------------------------------------
def bf16_grid(x):
    return orc.round_to_bf16_grid(np.ascontiguousarray(x, dtype=np.float32))


def bf16_bits(x_on_grid):
    return (np.ascontiguousarray(x_on_grid, dtype=np.float32).view(np.uint32) >> 16).astype(np.uint16)


def fixed_units(absmax_bits):
    """ptx.cuh fixed_unit: 2^(E-6), E the (clamped) biased exponent of the row maximum."""
    e = np.clip((absmax_bits >> 23).astype(np.int64), 16, 240)
    return np.ldexp(np.float32(1.0), (e - 6 - 127).astype(np.int32)).astype(np.float32)


def _planes(d):
    nb = d.ld // 64
    return mem(d.planes, np.uint16, d.nv * nb * d.rows_p * 64).reshape(d.nv, nb, d.rows_p, 64)


def _split(x, unit, nv=2):
    """((hi, lo, lo2), inexact) of one copy: fixed-point slice + bf16 remainder(s) when `unit` is given, else hi / lo.
    nv = 3 (the KL data term of ADMM): a second remainder plane, exactness judged after it."""
    zero = np.zeros_like(x, dtype=np.float32)
    if unit is None:
        hi = bf16_grid(x)
        lo = bf16_grid(x - hi)
        return (hi, lo, zero), lo != 0
    hi0 = (np.rint(x * (np.float32(1.0) / unit)) * unit).astype(np.float32)
    rem = (x - hi0).astype(np.float32)
    lo = bf16_grid(rem)
    left = (rem - lo).astype(np.float32)
    lo2 = zero
    if nv > 2:
        lo2 = bf16_grid(left)
        left = (left - lo2).astype(np.float32)
    return (bf16_grid(hi0), lo, lo2), left != 0


def np_ingest(x, row0, vr, vc, stats, rowmax, colmax):
    """x: (nrows, ncols) fp32 chunk of V starting at global row row0."""
    nrows, ncols = x.shape
    fixed = rowmax is not None
    with np.errstate(all="ignore"):
        planes, inex = _split(x, fixed_units(rowmax[row0:row0 + nrows])[:, None] if fixed else None, vr.nv)
        pr = _planes(vr)
        lr = np.arange(nrows) + row0 - vr.origin
        ok = (lr >= 0) & (lr < vr.rows)
        pad = np.zeros((nrows, vr.ld), np.uint16)
        for q, val in enumerate(planes[:vr.nv]):
            pad[:, :ncols] = bf16_bits(val)
            blk = pad.reshape(nrows, vr.ld // 64, 64).transpose(1, 0, 2)       # [nb][nrows][64]
            nbc = (ncols + 63) // 64
            tail = ncols - (nbc - 1) * 64
            pr[q][:nbc - 1, lr[ok]] = blk[:nbc - 1, ok]
            pr[q][nbc - 1, lr[ok], :tail] = blk[nbc - 1, ok, :tail]
        need_lo = bool(inex.any())
        # transposed copy: columns [origin, origin + rows) of V, all of this chunk's rows
        c0, c1 = vc.origin, min(vc.origin + vc.rows, ncols)
        if c1 > c0:
            xt = np.ascontiguousarray(x[:, c0:c1].T)                           # (cols in window, nrows)
            planes, inex = _split(xt, fixed_units(colmax[c0:c1])[:, None] if fixed else None, vc.nv)
            if fixed:
                need_lo = need_lo or bool(inex.any())
            pc = _planes(vc)
            g = np.arange(nrows) + row0                                        # column index in the V^T copy
            lc = np.arange(c1 - c0)[None, :]                                   # local column = row of the V^T copy
            for q, val in enumerate(planes[:vc.nv]):
                pc[q][(g >> 6)[:, None], lc, (g & 63)[:, None]] = bf16_bits(val).T
    vmin = np.float32(x.min()) if x.size else np.float32(np.inf)
    bits = int(np.array([vmin], np.float32).view(np.uint32)[0])
    key = (~bits & 0xFFFFFFFF) if bits & 0x80000000 else (bits | 0x80000000)
    stats[0] = min(int(stats[0]), key)
    if need_lo:
        stats[1] |= 1
    if not np.all(np.isfinite(x)):
        stats[2] |= 1


def np_absmax(x, row0, rowmax, colmax):
    a = np.abs(x).astype(np.float32)
    nrows, ncols = x.shape
    rowmax[row0:row0 + nrows] = np.maximum(rowmax[row0:row0 + nrows], a.max(1).view(np.uint32))
    colmax[:ncols] = np.maximum(colmax[:ncols], a.max(0).view(np.uint32))


def np_kl_vaux(prod, d_out, dual_v, v):
    """In-place on the strided views; fp64 arithmetic like the kernel, results rounded to fp32."""
    P, d0 = prod.astype(np.float64), dual_v.astype(np.float64)
    t = P - d0 - 1.0
    root = np.sqrt(t * t + 4.0 * v)
    with np.errstate(all="ignore"):
        va = np.where(t >= 0.0, 0.5 * (t + root), np.where(root - t > 0.0, 2.0 * v / (root - t), 0.0))
    dn = (d0 + va - P).astype(np.float32)
    dual_v[...] = dn
    d_out[...] = (va + dn.astype(np.float64)).astype(np.float32)

class Fac:
    """View of an nmfb_factor."""

    def __init__(self, f, KP):
        self.f, self.KP = f, KP
        self.L, self.Lp = f.L, f.Lp
        self.ptr = int(f.master) if f.master else 0
        self.master = mem(self.ptr, np.float32, KP * f.Lp).reshape(KP, f.Lp) if self.ptr else None
        self.w0 = f.win0 if f.win_p > 0 else 0
        self.wp = f.win_p if f.win_p > 0 else f.Lp
        self.live = max(0, min(f.L - self.w0, self.wp))
        self.sharded = f.win_p > 0

    def key(self, lo=0, hi=None):
        return hash(self.master[:, lo:hi].tobytes())

    @property
    def window(self):
        return (self.w0, self.w0 + self.wp)


class SoftLib:
    def __init__(self):
        self.emu = emu_lib()
        self.planes = {}            # master ptr -> {"cpl": (form, key), "rpl": key}
        self.cpl_owner = {}         # cplanes ptr -> (master ptr, factor struct copy)
        self.numer = None           # numerator of the running AO-ADMM side (window-local, fp64)
        self.launches = 0
        self.log = []

    # ------------------------------------------------------------------ trivia
    def nmfb_last_error(self):
        return b"softdev"

    def nmfb_launch_count(self):
        return self.launches

    def nmfb_padded_components(self, k):
        return 64 if k <= 64 else (128 if k <= 128 else 256)

    def nmfb_padded_length(self, L):
        return (L + 255) // 256 * 256

    def nmfb_partial_bytes(self, m, n, k, num_sms):
        return 4 * 4 * self.nmfb_padded_components(k) * max(self.nmfb_padded_length(m), self.nmfb_padded_length(n))

    def nmfb_obj_partial_len(self, m, n, num_sms):
        return max(num_sms, ((m + 63) // 64) * ((n + 63) // 64))

    # ------------------------------------------------------------------ plane bookkeeping
    # Freshness is tracked per row range of a factor: an update kernel re-derives the planes of the window it
    # wrote, nmfb_refresh_planes / nmfb_fixed_planes those of the whole factor; a consumer names the range it reads.
    def _mark(self, fac: Fac, cpl=None, rpl=False, rng=None):
        lo, hi = rng or (0, fac.Lp)
        st = self.planes.setdefault(fac.ptr, {"cpl": [], "rpl": []})
        key = fac.key(lo, hi)
        if cpl is not None:
            st["cpl"] = [e for e in st["cpl"] if e[0] == cpl and not (lo <= e[1] and e[2] <= hi)] + [(cpl, lo, hi, key)]
        if rpl:
            st["rpl"] = [e for e in st["rpl"] if not (lo <= e[1] and e[2] <= hi)] + [(None, lo, hi, key)]
        if fac.f.cplanes:
            self.cpl_owner[int(fac.f.cplanes)] = fac

    def _fresh(self, fac: Fac, kind, form, rng):
        lo, hi = rng or (0, fac.Lp)
        for f, a, b, key in self.planes.get(fac.ptr, {}).get(kind, []):
            if f == form and a <= lo and hi <= b and key == fac.key(a, b):
                return True
        return False

    def _need_cpl(self, fac: Fac, form, who, rng=None):
        forms = {e[0] for e in self.planes.get(fac.ptr, {}).get("cpl", [])}
        assert self._fresh(fac, "cpl", form, rng), \
            f"{who}: needs current {form} cplanes; held forms {sorted(map(str, forms))} are stale or of the other form"

    def _need_rpl(self, fac: Fac, who, rng=None):
        assert self._fresh(fac, "rpl", None, rng), f"{who}: rplanes are stale"

    # ------------------------------------------------------------------ data planes (real kernels, emulated)
    @staticmethod
    def decode(d):
        """(rows, cols) float64 matrix held by the K-blocked planes of an nmfb_data."""
        assert d.blocked == 1
        nb = d.ld // 64
        pl = mem(d.planes, np.uint16, d.nv * nb * d.rows_p * 64).reshape(d.nv, nb, d.rows_p, 64)
        val = (pl.astype(np.uint32) << 16).view(np.float32).astype(np.float64).sum(0)
        return val.transpose(1, 0, 2).reshape(d.rows_p, d.ld)[:d.rows, :d.cols]

    def _ingest(self, src, is64, ld, row0, nrows, vr, vc, stats, rowmax, colmax):
        vr, vc = _obj(vr), _obj(vc)
        assert vr.origin + vr.rows <= vc.cols and vc.origin + vc.rows <= vr.cols and vr.nv == vc.nv
        assert 0 <= row0 and row0 + nrows <= vc.cols
        if ld < 0:                                                # K-blocked source [cols / 64][-ld][64]
            nbk = (vr.cols + 63) // 64
            b = mem(src, np.float64 if is64 else np.float32, nbk * (-ld) * 64).reshape(nbk, -ld, 64)
            x = b.transpose(1, 0, 2).reshape(-ld, nbk * 64)[:nrows, :vr.cols]
        else:
            chunk = mem(src, np.float64 if is64 else np.float32, (nrows - 1) * ld + vr.cols)
            x = np.lib.stride_tricks.as_strided(chunk, (nrows, vr.cols), (ld * chunk.itemsize, chunk.itemsize))
        np_ingest(x.astype(np.float32), row0, vr, vc, mem(stats, np.uint32, 4),
                  mem(rowmax, np.uint32, vc.cols) if rowmax else None,
                  mem(colmax, np.uint32, vr.cols) if colmax else None)
        self.launches += 1
        return 0

    def nmfb_ingest_data(self, src, is64, ld, row0, nrows, vr, vc, stats, stream):
        assert not _obj(vr).fixed
        return self._ingest(src, is64, ld, row0, nrows, vr, vc, stats, None, None)

    def nmfb_ingest_data_fixed(self, src, is64, ld, row0, nrows, vr, vc, stats, rowmax, colmax, stream):
        assert _obj(vr).fixed and _obj(vc).fixed and _obj(vr).nv >= 2 and rowmax and colmax
        return self._ingest(src, is64, ld, row0, nrows, vr, vc, stats, rowmax, colmax)

    def nmfb_data_absmax(self, src, is64, ld, row0, nrows, ncols, rowmax, colmax, stream):
        chunk = mem(src, np.float64 if is64 else np.float32, (nrows - 1) * ld + ncols)
        x = np.lib.stride_tricks.as_strided(chunk, (nrows, ncols), (ld * chunk.itemsize, chunk.itemsize))
        np_absmax(x.astype(np.float32), row0, mem(rowmax, np.uint32, row0 + nrows), mem(colmax, np.uint32, ncols))
        self.launches += 1
        return 0

    def nmfb_upload_2d(self, dst, dst_pitch, src_host, src_pitch, width_bytes, height, stream):
        """cudaMemcpy2DAsync stand-in: `height` rows of `width_bytes` bytes between pitched buffers."""
        src = mem(src_host, np.uint8, (height - 1) * src_pitch + width_bytes)
        out = mem(dst, np.uint8, (height - 1) * dst_pitch + width_bytes)
        s2 = np.lib.stride_tricks.as_strided(src, (height, width_bytes), (src_pitch, 1))
        d2 = np.lib.stride_tricks.as_strided(out, (height, width_bytes), (dst_pitch, 1))
        d2[...] = s2
        return 0

    def nmfb_kl_vaux(self, vr, prod, d_out, ld, dual_v, done_flag, stream):
        vr = _obj(vr)
        assert vr.blocked and 1 <= vr.nv <= 2 and ld >= vr.cols
        if done_flag and mem(done_flag, np.int32, 1)[0] != 0:
            return 0
        n = (vr.rows - 1) * ld + vr.cols
        shape, st = (vr.rows, vr.cols), (ld * 4, 4)
        view = lambda ptr: np.lib.stride_tricks.as_strided(mem(ptr, np.float32, n), shape, st)   # noqa: E731
        np_kl_vaux(view(prod), view(d_out), view(dual_v), self.decode(vr))
        self.launches += 1
        return 0

    def nmfb_row_planes3(self, KP, f, planes3, stream):
        """The three 24-bit row planes of a master: the stand-in keeps the fp32 values they represent."""
        fac = Fac(_obj(f), KP)
        if not hasattr(self, "planes3"):
            self.planes3 = {}
        self.planes3[int(planes3)] = fac.master.copy()
        self.launches += 1
        return 0

    def nmfb_kl_data_step(self, vr, a_planes3, a_Lp, a_row0, b_planes3, b_Lp, KP, dual_v, d_out, ld, rowmax, colmax,
                          done_flag, num_sms, stream):
        """kl_data.cu at the semantic level: P = A.B in fp32-grade arithmetic, the v_aux / dual_v step, |max| of D."""
        vr = _obj(vr)
        assert vr.blocked and 1 <= vr.nv <= 2 and ld >= vr.cols and KP <= 128
        if done_flag and mem(done_flag, np.int32, 1)[0] != 0:
            return 0
        a = self.planes3[int(a_planes3)][:, a_row0:a_row0 + vr.rows].T.astype(np.float64)
        b = self.planes3[int(b_planes3)][:, :vr.cols].astype(np.float64)
        prod = (a @ b).astype(np.float32)
        nb = ld // 64                                             # K-blocked fp32 [ld / 64][rows_p][64]
        blk = lambda ptr: mem(ptr, np.float32, nb * vr.rows_p * 64).reshape(nb, vr.rows_p, 64)   # noqa: E731
        rowmajor = lambda b: b.transpose(1, 0, 2).reshape(vr.rows_p, ld)[:vr.rows, :vr.cols].copy()   # noqa: E731

        def put(b, x):
            full = b.transpose(1, 0, 2).reshape(vr.rows_p, ld).copy()
            full[:vr.rows, :vr.cols] = x
            b[...] = full.reshape(vr.rows_p, nb, 64).transpose(1, 0, 2)

        dv, dd = rowmajor(blk(dual_v)), np.zeros((vr.rows, vr.cols), np.float32)
        np_kl_vaux(prod, dd, dv, self.decode(vr))
        put(blk(dual_v), dv)
        put(blk(d_out), dd)
        rm, cm = mem(rowmax, np.uint32, vr.origin + vr.rows), mem(colmax, np.uint32, vr.cols)
        rm[vr.origin:] = 0
        cm[:] = 0
        np_absmax(dd, vr.origin, rm, cm)
        self.launches += 2
        return 0

    # ------------------------------------------------------------------ factors
    def nmfb_import_factor(self, src, is64, stride_l, stride_c, k, KP, f, stream):
        fac = Fac(_obj(f), KP)
        n_src = (fac.L - 1) * stride_l + (k - 1) * stride_c + 1
        s = mem(src, np.float64 if is64 else np.float32, n_src)
        r, c = np.arange(fac.L)[None, :], np.arange(k)[:, None]
        fac.master[:] = 0
        fac.master[:k, :fac.L] = s[r * stride_l + c * stride_c].astype(np.float32)
        self._mark(fac, cpl="float", rpl=True)
        return 0

    def nmfb_export_factor(self, f, k, KP, dst, is64, stride_l, stride_c, stream):
        fac = Fac(_obj(f), KP)
        n_dst = (fac.L - 1) * stride_l + (k - 1) * stride_c + 1
        d = mem(dst, np.float64 if is64 else np.float32, n_dst)
        r, c = np.arange(fac.L)[None, :], np.arange(k)[:, None]
        d[r * stride_l + c * stride_c] = fac.master[:k, :fac.L]
        return 0

    def nmfb_refresh_planes(self, KP, f, stream):
        self._mark(Fac(_obj(f), KP), cpl="float", rpl=True)
        return 0

    def nmfb_fixed_planes(self, KP, f, stream):
        self._mark(Fac(_obj(f), KP), cpl="fixed")
        return 0

    # ------------------------------------------------------------------ objectives
    def _objective(self, p, kind, out, need_planes=True):
        KP, k = p.KP, p.k
        w, h = Fac(p.w, KP), Fac(p.h, KP)
        if need_planes:
            self._need_rpl(w, "objective (W)", w.window)
            self._need_rpl(h, "objective (H)")
        v = self.decode(p.v_rows)                                 # this rank's rows of V
        wl = w.master[:k, w.w0:w.w0 + p.v_rows.rows].astype(np.float64).T
        wh = wl @ h.master[:k, :p.n].astype(np.float64)
        with np.errstate(all="ignore"):
            mem(out, np.float64, 1)[0] = orc.objective(v, wh, kind)
        return 0

    def nmfb_objective_eu(self, p, impl, out, stream):
        return self._objective(_obj(p), "eu", out, need_planes=(impl == 0))

    def nmfb_objective_kl(self, p, out, stream):
        return self._objective(_obj(p), "kl", out)

    # ------------------------------------------------------------------ MUR
    def _sides(self, p, side):
        """(self, other, data) of a MUR half step: side 0 = W with V rows, side 1 = H with V^T rows."""
        KP = p.KP
        if side == 0:
            return Fac(p.w, KP), Fac(p.h, KP), p.v_rows
        return Fac(p.h, KP), Fac(p.w, KP), p.v_cols

    def _after_update(self, fac: Fac, cpl, rpl=True):
        """An update kernel re-derives the planes of the window it wrote (the whole factor when unsharded)."""
        self._mark(fac, cpl=cpl, rpl=rpl, rng=fac.window)

    def nmfb_mur_eu_half_step(self, p, side, lam, stream):
        p = _obj(p)
        k = p.k
        x, y, data = self._sides(p, side)
        self._need_cpl(y, "float", f"mur_eu side {side}: Gram / GEMM operand")
        d = self.decode(data)                                      # (window rows, L_other)
        yv = y.master[:k, :y.L].astype(np.float64)
        xw = x.master[:k, x.w0:x.w0 + x.live].astype(np.float64)
        num = yv @ d[:x.live].T                                    # mur.py:29 / :45
        den = (yv @ yv.T) @ xw + float(lam) * xw + 1e-9
        x.master[:k, x.w0:x.w0 + x.live] = (xw * num / den).astype(np.float32)
        self._after_update(x, "float")
        self.launches += 3
        return 0

    def nmfb_mur_eu_step(self, p, lw, lh, out, stream):
        self.nmfb_mur_eu_half_step(p, 0, lw, stream)
        self.nmfb_mur_eu_half_step(p, 1, lh, stream)
        if out:
            self.nmfb_objective_eu(p, 0, out, stream)
        return 0

    def nmfb_mur_kl_half_step(self, p, side, lam, obj_out, stream):
        p = _obj(p)
        k = p.k
        x, y, data = self._sides(p, side)
        self._need_rpl(x, f"mur_kl side {side}: X operand", x.window)
        self._need_cpl(y, "float", f"mur_kl side {side}: Y operand")
        d = self.decode(data)[:x.live]                             # rows of V (side 0) or of V^T (side 1)
        yv = y.master[:k, :y.L].astype(np.float64)
        xw = x.master[:k, x.w0:x.w0 + x.live].astype(np.float64)
        s = xw.T @ yv                                              # (window, L_other)
        if obj_out:
            with np.errstate(all="ignore"):
                mem(obj_out, np.float64, 1)[0] = orc.objective(d, s, "kl")
        a = xw * (yv @ (d / (s + 1e-9)).T)                         # mur.py:24-27 / :40-43
        b = yv.sum(1)[:, None]
        new = 2 * a / (b + np.sqrt(b * b + 4 * float(lam) * a))
        x.master[:k, x.w0:x.w0 + x.live] = new.astype(np.float32)
        self._after_update(x, "float")
        self.launches += 3
        return 0

    def nmfb_mur_kl_step(self, p, lw, lh, out, stream):
        self.nmfb_mur_kl_half_step(p, 0, lw, None, stream)
        self.nmfb_mur_kl_half_step(p, 1, lh, None, stream)
        if out:
            self.nmfb_objective_kl(p, out, stream)
        return 0

    # ------------------------------------------------------------------ ADMM
    @staticmethod
    def _state(s, KP):
        flags = mem(s.flags, np.int32, 8)
        scal = mem(s.scal, np.float64, 4)
        return flags, scal

    def _apply(self, p, s, side, x: Fac, dual_ptr, aux: Fac, numer, cho, rho, lam, prox, norms):
        """solve / prox / dual on the window (admm.py:230,126-156,320-321; ao_admm.py:59-62)."""
        k, KP = p.k, p.KP
        w0, live = x.w0, x.live
        u = mem(dual_ptr, np.float32, KP * x.Lp).reshape(KP, x.Lp)
        xw = x.master[:k, w0:w0 + live].astype(np.float64)
        uw = u[:k, w0:w0 + live].astype(np.float64)
        b = numer[:, :live] + rho * (xw + uw)
        a = sla.cho_solve(cho, b)
        kind = {0: "nn", 1: "l1n", 2: "l2n"}[int(prox)]
        xn = orc.prox(kind, a, uw, rho, float(lam))
        un = uw + xn - a
        x.master[:k, w0:w0 + live] = xn.astype(np.float32)
        u[:k, w0:w0 + live] = un.astype(np.float32)
        if aux is not None and aux.ptr:
            aux.master[:k, w0:w0 + live] = a.astype(np.float32)
        if norms:
            part = mem(s.norm_partial, np.float64, 4 * (x.Lp // 64)).reshape(-1, 4)
            xs, us = x.master[:k, w0:w0 + x.wp].astype(np.float64), u[:k, w0:w0 + x.wp].astype(np.float64)
            af = np.zeros_like(xs)
            af[:, :live] = a
            xp = np.zeros_like(xs)
            xp[:, :live] = xw
            for blk in range(x.wp // 64):
                sl = slice(blk * 64, blk * 64 + 64)
                part[w0 // 64 + blk] = [((xs - af)[:, sl] ** 2).sum(), (xs[:, sl] ** 2).sum(),
                                        ((xs - xp)[:, sl] ** 2).sum(), (us[:, sl] ** 2).sum()]
        self.launches += 1

    def _system(self, other: Fac, k, rho_mode, rho_in, flags):
        g = other.master[:k, :other.L].astype(np.float64)
        g = g @ g.T
        rho = np.trace(g) / k if rho_mode == 1 else float(rho_in)
        try:
            cho = sla.cho_factor(g + rho * np.eye(k), lower=True)
            if not np.all(np.isfinite(cho[0])):
                raise np.linalg.LinAlgError
        except (np.linalg.LinAlgError, ValueError):
            flags[0] = 1
            cho = (np.eye(k), True)
        self.launches += 2
        return cho, rho

    def nmfb_admm_prepare_l2n(self, p, s, side, lam, stream):
        assert _obj(s).binv and _obj(s).tik_scratch and _obj(p).KP <= 128
        return 0

    def nmfb_admm_half_step(self, p, s, side, rho, lam, prox, stream):
        p, s = _obj(p), _obj(s)
        k, KP = p.k, p.KP
        flags, scal = self._state(s, KP)
        other = Fac(s.w_aux if side == 0 else s.h_aux, KP)
        x = Fac(p.h if side == 0 else p.w, KP)
        aux = Fac(s.h_aux if side == 0 else s.w_aux, KP)
        data = p.v_cols if side == 0 else p.v_rows
        form = "fixed" if s.exact else "float"
        self._need_cpl(other, form, f"admm side {side}: Gram / GEMM operand (auxiliary of the other factor)")
        assert (not s.exact) or data.fixed, "exact mode needs fixed-point data planes"
        cho, rho = self._system(other, k, 0, rho, flags)           # admm.py:220
        scal[side] = rho
        d = self.decode(data)[:x.live]
        numer = other.master[:k, :other.L].astype(np.float64) @ d.T     # admm.py:221 / :224
        self._apply(p, s, side, x, s.dual_h if side == 0 else s.dual_w, aux, numer, cho, rho, lam, prox, False)
        self._after_update(x, None)                                # rplanes only (the objective's operands)
        if not s.exact:
            self._after_update(aux, "float", rpl=False)
        return 0

    def nmfb_admm_publish(self, s, mailbox, stream):
        s = _obj(s)
        flags = mem(s.flags, np.int32, 8)
        mb = mem(mailbox, np.float64, 8)
        mb[1], mb[2], mb[3] = flags[2], flags[3], flags[0]
        return 0

    def nmfb_admm_step(self, p, s, rho, lw, pw, lh, ph, mailbox, stream):
        po, so = _obj(p), _obj(s)
        self.nmfb_admm_half_step(p, s, 0, rho, lh, ph, stream)
        if so.exact:
            self.nmfb_fixed_planes(po.KP, so.h_aux, stream)
        self.nmfb_admm_half_step(p, s, 1, rho, lw, pw, stream)
        if so.exact:
            self.nmfb_fixed_planes(po.KP, so.w_aux, stream)
        self.nmfb_objective_eu(p, 0, mailbox, stream)
        return self.nmfb_admm_publish(s, mailbox, stream)

    # ------------------------------------------------------------------ AO-ADMM
    def _ao_roles(self, p, s, side):
        KP = p.KP
        other = Fac(p.w if side == 0 else p.h, KP)
        x = Fac(p.h if side == 0 else p.w, KP)
        data = p.v_cols if side == 0 else p.v_rows
        dual = s.dual_h if side == 0 else s.dual_w
        aux = Fac(s.h_aux if side == 0 else s.w_aux, KP)
        return other, x, data, dual, aux

    def nmfb_aoadmm_side_begin(self, p, s, side, stream):
        p, s = _obj(p), _obj(s)
        flags, scal = self._state(s, p.KP)
        other, x, data, _dual, _aux = self._ao_roles(p, s, side)
        flags[1] = 0
        flags[2 + side] = 0
        self._need_cpl(other, "fixed" if s.exact else "float", f"ao_admm side {side}: Gram / GEMM operand")
        assert (not s.exact) or data.fixed, "exact mode needs fixed-point data planes"
        self.cho, rho = self._system(other, p.k, 1, 0.0, flags)    # ao_admm.py:53-55
        scal[side] = rho
        d = self.decode(data)[:x.live]
        self.numer = other.master[:p.k, :other.L].astype(np.float64) @ d.T      # ao_admm.py:56
        self.launches += 1
        return 1

    def nmfb_panel_gemm_exact(self, data, cplanes, factor_Lp, KP, partial, partial_bytes, out_Lp, num_sms, stream):
        data = _obj(data)
        fac = self.cpl_owner[int(cplanes)]
        fac = Fac(fac.f, KP)
        self._need_cpl(fac, "fixed", "panel_gemm_exact: factor operand")
        assert data.fixed and data.nv in (2, 3) and factor_Lp == fac.Lp and data.cols == fac.L
        k = self._k
        self.numer = fac.master[:k, :fac.L].astype(np.float64) @ self.decode(data).T
        self.launches += 1
        return 1

    def nmfb_aoadmm_inner(self, p, s, side, first, slots, lam, prox, stream):
        p, s = _obj(p), _obj(s)
        flags, scal = self._state(s, p.KP)
        if flags[1]:
            return 0
        other, x, data, dual, aux = self._ao_roles(p, s, side)
        assert slots == 1
        if first:
            self.c64 = self.numer[:, :x.live].copy()
        self._apply(p, s, side, x, dual, aux, self.c64, self.cho, scal[side], lam, prox, True)
        self._after_update(x, None if s.exact else "float")
        return 0

    def nmfb_aoadmm_inner_flag(self, p, s, side, stream):
        p, s = _obj(p), _obj(s)
        flags, _scal = self._state(s, p.KP)
        if flags[1]:
            return 0
        x = Fac(p.h if side == 0 else p.w, p.KP)
        t = mem(s.norm_partial, np.float64, 4 * (x.Lp // 64)).reshape(-1, 4).sum(0)
        with np.errstate(all="ignore"):
            r, sres = np.sqrt(t[0]) / np.sqrt(t[1]), np.sqrt(t[2]) / np.sqrt(t[3])
        if r < 1e-2 and sres < 1e-2:                               # ao_admm.py:36-43; NaN compares false
            flags[1] = 1
        else:
            flags[2 + side] += 1
        self.launches += 1
        return 0

    def nmfb_aoadmm_step(self, p, s, admm_iter, lw, pw, lh, ph, mailbox, stream):
        po, so = _obj(p), _obj(s)
        for side in range(2):
            lam, prox = (lh, ph) if side == 0 else (lw, pw)
            slots = self.nmfb_aoadmm_side_begin(p, s, side, stream)
            for it in range(admm_iter):
                self.nmfb_aoadmm_inner(p, s, side, 1 if it == 0 else 0, slots, lam, prox, stream)
                self.nmfb_aoadmm_inner_flag(p, s, side, stream)
            if so.exact:
                self.nmfb_fixed_planes(po.KP, po.h if side == 0 else po.w, stream)
        self.nmfb_objective_eu(p, 0, mailbox, stream)
        return self.nmfb_admm_publish(s, mailbox, stream)


class SoftEngine(Engine):
    """Engine whose device is host memory and whose library is SoftLib."""

    def _open_device(self, device):
        lib = SoftLib()
        return lib, torch.device("cpu"), 148

    def __init__(self, *a, **kw):
        super().__init__(*a, **kw)
        self.lib._k = self.k

    def _device_ctx(self):
        return contextlib.nullcontext()

    def _sync_stream(self):
        pass

    def _hbm_free(self) -> int:
        return 1 << 40

    @property
    def stream(self) -> int:
        return 0


@contextlib.contextmanager
def soft_device():
    """Route the public drivers to SoftEngine for the duration of the block."""
    import nmf_b200.admm as admm_mod
    import nmf_b200.ao_admm as ao_mod
    import nmf_b200.mur as mur_mod
    saved = [(m, m.Engine) for m in (mur_mod, admm_mod, ao_mod)]
    for m, _ in saved:
        m.Engine = SoftEngine
    try:
        yield
    finally:
        for m, e in saved:
            m.Engine = e
