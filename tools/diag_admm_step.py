"""Component-wise accuracy of ONE ADMM half step on the GPU against fp64 NumPy, from an oracle state: Gram, inverse,
numerator (exact-hi GEMM), auxiliary, prox'd factor, dual -- relative error AND mean signed relative error (a bias
accumulates over iterations, noise does not).  1 GPU."""
import ctypes as C
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch

from oracle import nmf_oracle as orc
from nmf_b200._lib import check
from nmf_b200.engine import Engine


def stats(name, got, want):
    got, want = np.asarray(got, np.float64), np.asarray(want, np.float64)
    rel = np.linalg.norm(got - want) / np.linalg.norm(want)
    mask = np.abs(want) > 1e-3 * np.abs(want).max()
    signed = np.mean((got[mask] - want[mask]) / want[mask])
    bias = (got - want).sum() / np.abs(want).sum()
    print(f"{name:28s} rel {rel:.2e}   mean signed rel (large entries) {signed:+.2e}   sum(diff)/sum|want| {bias:+.2e}", flush=True)


m, n, k, rho, lam = 2048, 1536, 64, 100.0, 0.1
v = orc.planted(m, n, k, 0).astype(np.float64)
w, h = orc.seeded_init(m, n, k, 1)
w_aux, h_aux, dw, dh = w.copy(), h.copy(), np.zeros_like(w), np.zeros_like(h)
for i in range(int(sys.argv[1]) if len(sys.argv) > 1 else 8):        # oracle iterations (admm.py:294-301,320-321)
    h_aux = orc.admm_aux(h, dh, w_aux, v, rho)
    w_aux = orc.admm_aux(w.T, dw.T, h_aux.T, v.T, rho).T
    hn = orc.prox("l1n", h_aux, dh, rho, lam)
    wn = orc.prox("nn", w_aux, dw, rho, 0.0)
    dh, dw, h, w = dh + hn - h_aux, dw + wn - w_aux, hn, wn

f32 = lambda a: np.asarray(a, np.float32)
eng = Engine(m, n, k)
eng.upload_data(v, fixed=True)
eng.set_factors(f32(w), f32(h))
eng.alloc_admm(with_aux=True)
KP = eng.KP


def put(fac_master, arr_kl):          # arr [k][L] -> master [KP][Lp]
    fac_master.zero_()
    fac_master[: arr_kl.shape[0], : arr_kl.shape[1]] = torch.from_numpy(f32(arr_kl)).to(eng.device)


put(eng.w_aux.master, w_aux.T); put(eng.h_aux.master, h_aux)
put(eng.dual_w, dw.T); put(eng.dual_h, dh)
for fac in (eng.w_aux, eng.h_aux):
    fs = fac.struct()
    check(eng.lib.nmfb_refresh_planes(KP, C.byref(fs), eng.stream))
    check(eng.lib.nmfb_fixed_planes(KP, C.byref(fs), eng.stream))
# what the GPU actually holds (fp32 roundings of the oracle state)
W_aux, H, DH = (f32(x).astype(np.float64) for x in (w_aux, h, dh))
torch.cuda.synchronize()
W_planes = eng.w_aux.cplanes[:, :k, :m].float().double().sum(0).cpu().numpy().T       # what the contractions see
want_gram = W_aux.T @ W_aux
want_c = W_aux.T @ v                                     # [k][n]
want_aux = np.linalg.solve(want_gram + rho * np.eye(k), want_c + rho * (H + DH))
want_h = orc.prox("l1n", want_aux, DH, rho, lam)
want_dh = DH + want_h - want_aux

# numerator alone through the C ABI
vr, vc = eng._data_structs()
slots = check(eng.lib.nmfb_panel_gemm_exact(C.byref(vc), eng.w_aux.cplanes.data_ptr(), eng.w_aux.Lp, KP,
                                            eng.partial.data_ptr(), eng.partial.numel() * 4, eng.h.Lp, eng.num_sms,
                                            eng.stream))
torch.cuda.synchronize()
got_c = eng.partial[: slots * KP * eng.h.Lp].view(slots, KP, eng.h.Lp).double().sum(0)[:k, :n].cpu().numpy()
stats("numerator W_aux^T V", got_c, want_c)
stats("numerator vs planes^T V", got_c, W_planes.T @ v)
stats("planes vs master", W_planes, W_aux)
want_gram_p = W_planes.T @ W_planes
want_aux_p = np.linalg.solve(want_gram_p + rho * np.eye(k), W_planes.T @ v + rho * (H + DH))

eng.admm_half_step(0, rho, lam, 1)
torch.cuda.synchronize()
stats("Gram W_aux^T W_aux", eng.gram64[0][:k, :k].cpu().numpy(), want_gram)
stats("inverse (G + rho I)^-1", eng.ainv[0][:k, :k].cpu().numpy(), np.linalg.inv(want_gram + rho * np.eye(k)))
stats("Gram vs planes^T planes", eng.gram64[0][:k, :k].cpu().numpy(), want_gram_p)
stats("h_aux", eng.h_aux.master[:k, :n].cpu().numpy(), want_aux)
stats("h_aux vs solve with planes", eng.h_aux.master[:k, :n].cpu().numpy(), want_aux_p)
stats("ridge(planes) vs ridge(master)", want_aux_p, want_aux)
stats("H = prox", eng.h.master[:k, :n].cpu().numpy(), want_h)
stats("dual_h", eng.dual_h[:k, :n].cpu().numpy(), want_dh)
