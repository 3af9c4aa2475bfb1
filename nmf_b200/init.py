"""Initial factors.  Random init draws from the GLOBAL NumPy RNG in the reference's order
(|randn(m,k)| then |randn(k,n)|: mur.py:108-109, admm.py:23-24, ao_admm.py:22-23) so that a
user script seeded with np.random.seed(s) starts from the same point as the reference."""
from __future__ import annotations

import numpy as np


def initial_factors(x, k, nndsvd_init=(False, 'zero')):
    m, n = x.shape
    if nndsvd_init[0]:
        return nndsvd(x, k, nndsvd_init[1])
    w = np.abs(np.random.randn(m, k))
    h = np.abs(np.random.randn(k, n))
    return w, h


def nndsvd(x, rank, variant='zero'):
    """NNDSVD (Boutsidis & Gallopoulos), the reference's utils.nndsvd (utils.py:36-93), with the
    SVD on the GPU (cuSOLVER through torch.linalg.svd -- one-off setup, not the solver loop)."""
    import torch
    if not torch.cuda.is_available():
        raise RuntimeError("nmf_b200 needs a CUDA device")
    xd = torch.as_tensor(np.ascontiguousarray(x, dtype=np.float64), device='cuda')
    u, s, vt = torch.linalg.svd(xd, full_matrices=False)
    u, s, vt = u[:, :rank], s[:rank], vt[:rank, :]
    up, un = u.clamp(min=0), (-u).clamp(min=0)
    vp, vn = vt.clamp(min=0), (-vt).clamp(min=0)
    nup, nun = up.norm(dim=0), un.norm(dim=0)
    nvp, nvn = vp.norm(dim=1), vn.norm(dim=1)
    pos, neg = nup * nvp, nun * nvn
    use_pos = pos >= neg
    use_pos[0] = True                                  # first triplet: |u0|, |v0| (utils.py:55-56)
    uu = torch.where(use_pos[None, :], up, un)
    vv = torch.where(use_pos[:, None], vp, vn)
    nu = torch.where(use_pos, nup, nun)
    nv = torch.where(use_pos, nvp, nvn)
    sc = torch.sqrt(s * torch.where(use_pos, pos, neg))
    w = uu * (sc / nu)[None, :]
    h = vv * (sc / nv)[:, None]
    w[:, 0] = torch.sqrt(s[0]) * u[:, 0].abs()
    h[0, :] = torch.sqrt(s[0]) * vt[0, :].abs()
    w, h = w.cpu().numpy(), h.cpu().numpy()
    if variant == 'mean':
        mu = np.mean(x)
        w[w == 0] = mu
        h[h == 0] = mu
    elif variant == 'random':
        mu = np.mean(x)
        fill = mu * np.random.random_sample(w.shape) / 100
        w = np.where(w == 0, fill, w)
        fill = mu * np.random.random_sample(h.shape) / 100
        h = np.where(h == 0, fill, h)
    return w, h
