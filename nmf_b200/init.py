"""Initial factors.  Random init draws from the GLOBAL NumPy RNG in the reference's order
(|randn(m,k)| then |randn(k,n)|: mur.py:108-109, admm.py:23-24, ao_admm.py:22-23) so that a
user script seeded with np.random.seed(s) starts from the same point as the reference."""
from __future__ import annotations

import numpy as np


def initial_factors(x, k, nndsvd_init=(False, 'zero'), engine=None):
    """engine: the Engine of a sharded run (world > 1).  Rank 0 alone then draws from its global RNG (or runs
    NNDSVD) and the result is broadcast, so that every rank starts from the SAME replicated factors whatever
    the other ranks' RNG states are -- the start of a sharded run equals that of the single-GPU run seeded
    like rank 0."""
    m, n = x.shape
    if engine is not None and getattr(engine, "world", 1) > 1:
        from .sharding import broadcast_from_rank0
        pair = _draw(x, k, nndsvd_init) if engine.rank == 0 else (None, None)
        w = broadcast_from_rank0(pair[0], (m, k), engine.device, engine.group)
        h = broadcast_from_rank0(pair[1], (k, n), engine.device, engine.group)
        return w, h
    return _draw(x, k, nndsvd_init)


def _draw(x, k, nndsvd_init):
    m, n = x.shape
    if nndsvd_init[0]:
        return nndsvd(x, k, nndsvd_init[1])
    w = np.abs(np.random.randn(m, k))
    h = np.abs(np.random.randn(k, n))
    return w, h


GRAM_LIMIT = 32768      # min(m, n) up to which the k leading triplets come from the Gram matrix's eigenvectors


def leading_triplets(xd, rank):
    """(u [m, rank], s [rank], vt [rank, n]) of the `rank` largest singular values, fp64 on xd's device.
    NNDSVD needs nothing else of the SVD (utils.py:44 computes all min(m, n) triplets and uses `rank` of
    them).  Up to GRAM_LIMIT the small Gram matrix (X^T X or X X^T, one fp64 GEMM) is diagonalised by
    cuSOLVER's symmetric eigensolver and the other side follows from u_i = X v_i / s_i -- O(mn min(m,n))
    GEMM work + one min(m,n)^3 eigenproblem instead of the SVD of the whole m x n matrix; the leading
    vectors carry a relative error of about eps * (s_1 / s_i)^2, far inside the 1e-4 parity bar for the
    well-separated leading part NNDSVD uses.  Larger problems take the full SVD."""
    import torch
    m, n = xd.shape
    if min(m, n) > GRAM_LIMIT:
        u, s, vt = torch.linalg.svd(xd, full_matrices=False)
        return u[:, :rank], s[:rank], vt[:rank, :]
    small_is_n = n <= m
    g = xd.T @ xd if small_is_n else xd @ xd.T
    try:
        lam, vec = torch.linalg.eigh(g)                   # ascending
    except RuntimeError:                                  # eigensolver did not converge: the SVD route (same device)
        u, s, vt = torch.linalg.svd(xd, full_matrices=False)
        return u[:, :rank], s[:rank], vt[:rank, :]
    lam, vec = lam.flip(0)[:rank], vec.flip(1)[:, :rank]
    s = lam.clamp(min=0).sqrt()
    safe = torch.where(s > 0, s, torch.ones_like(s))
    if small_is_n:
        vt = vec.T.contiguous()
        u = (xd @ vec) / safe[None, :]
    else:
        u = vec.contiguous()
        vt = (vec.T @ xd) / safe[:, None]
    return u, s, vt


def nndsvd(x, rank, variant='zero', _device='cuda'):
    """NNDSVD (Boutsidis & Gallopoulos), the reference's utils.nndsvd (utils.py:36-93), on the GPU
    (one-off setup, not the solver loop): leading singular triplets by `leading_triplets`, the sign
    split and scaling as torch ops.  (_device is for the CPU tests of this arithmetic only.)"""
    import torch
    if _device == 'cuda' and not torch.cuda.is_available():
        raise RuntimeError("nmf_b200 needs a CUDA device")
    xd = torch.as_tensor(np.ascontiguousarray(x, dtype=np.float64), device=_device)
    u, s, vt = leading_triplets(xd, rank)
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
