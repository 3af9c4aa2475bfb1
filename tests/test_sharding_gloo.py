"""CPU (gloo, world_size 2): host-side logic of the sharded runs -- shard bounds, the window
all-gather, and the MPI-FAUN exchange pattern itself, emulated with the oracle's update rules on
each rank's shard: two ranks that only ever see their rows / columns of V must reproduce the
unsharded oracle iteration exactly."""
import os
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

from nmf_b200.sharding import allgather_windows, allreduce_sum_, live_rows, shard_bounds  # noqa: E402


def test_shard_bounds_cover_and_align():
    for L, world in [(1000, 2), (65536, 8), (300, 4), (1, 2), (20000, 8)]:
        Lp, per, wins = shard_bounds(L, world)
        assert Lp == per * world and Lp >= L and per % 256 == 0
        assert wins[0][0] == 0 and wins[-1][1] == Lp
        assert all(wins[i][1] == wins[i + 1][0] for i in range(world - 1))
        assert sum(live_rows(L, lo, hi) for lo, hi in wins) == L
    assert shard_bounds(1000, 1) == (1024, 1024, [(0, 1024)])


def _worker(rank, world, port, ret):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from oracle import nmf_oracle as orc
        m, n, k, iters = 300, 200, 6, 5
        v = orc.planted(m, n, k, 0).astype(np.float64)
        w0, h0 = orc.seeded_init(m, n, k, 1)
        Mp, per_m, mb = shard_bounds(m, world)
        Np, per_n, nb = shard_bounds(n, world)
        r0, r1 = mb[rank][0], mb[rank][0] + live_rows(m, *mb[rank])
        c0, c1 = nb[rank][0], nb[rank][0] + live_rows(n, *nb[rank])
        v_rows, v_cols = v[r0:r1], v[:, c0:c1]                 # all this rank ever reads of V
        # component-major padded masters, like the device layout
        wt = torch.zeros((8, Mp), dtype=torch.float64); wt[:k, :m] = torch.from_numpy(w0.T)
        ht = torch.zeros((8, Np), dtype=torch.float64); ht[:k, :n] = torch.from_numpy(h0)
        objs = []
        for _ in range(iters):
            w, h = wt[:k, :m].numpy().T.copy(), ht[:k, :n].numpy().copy()
            # W window: rows r0:r1 need V[r0:r1, :], full H and H H^T            mur.py:29
            wl = w[r0:r1]
            wl = wl * (v_rows @ h.T) / (wl @ (h @ h.T) + 1e-9)
            wt[:k, r0:r1] = torch.from_numpy(wl.T)
            allgather_windows(wt, per_m)
            w = wt[:k, :m].numpy().T.copy()
            # H window: columns c0:c1 need V[:, c0:c1], full (new) W and W^T W    mur.py:45
            hl = h[:, c0:c1]
            hl = hl * (w.T @ v_cols) / ((w.T @ w) @ hl + 1e-9)
            ht[:k, c0:c1] = torch.from_numpy(hl)
            allgather_windows(ht, per_n)
            h = ht[:k, :n].numpy()
            part = torch.tensor([0.5 * np.sum((v_rows - w[r0:r1] @ h) ** 2)])
            allreduce_sum_(part)
            objs.append(float(part))
        tr = orc.mur_run(v, k, w0, h0, distance_type="eu", min_iter=10**9, max_iter=iters)
        ew = np.linalg.norm(wt[:k, :m].numpy().T - tr.w) / np.linalg.norm(tr.w)
        eh = np.linalg.norm(ht[:k, :n].numpy() - tr.h) / np.linalg.norm(tr.h)
        eo = max(abs(a - b) / b for a, b in zip(objs, tr.obj_history[1:]))
        pads_zero = float(wt[:, m:].abs().sum() + ht[:, n:].abs().sum() + wt[k:].abs().sum()) == 0.0
        ret[rank] = (ew, eh, eo, pads_zero)
    finally:
        dist.destroy_process_group()


def test_two_rank_exchange_reproduces_unsharded_oracle():
    world = 2
    ctx = mp.get_context("spawn")
    ret = ctx.Manager().dict()
    port = 29600 + (os.getpid() % 300)
    procs = [ctx.Process(target=_worker, args=(r, world, port, ret)) for r in range(world)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(120)
        assert p.exitcode == 0
    for r in range(world):
        ew, eh, eo, pads_zero = ret[r]
        assert ew < 1e-12 and eh < 1e-12 and eo < 1e-12 and pads_zero


# ----------------------------------------------------------------------------------------------
# ADMM / AO-ADMM: the exchange pattern of solvers.py (AdmmStepper / AoAdmmStepper, sharded branch)
# emulated with the oracle's arithmetic on each rank's shard
# ----------------------------------------------------------------------------------------------
def _admm_worker(rank, world, port, ret):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        import scipy.linalg as sla
        from oracle import nmf_oracle as orc
        from nmf_b200.sharding import allgather_flat
        m, n, k, iters, KP = 300, 280, 6, 6, 8
        v = orc.planted(m, n, k, 0).astype(np.float64)
        w0, h0 = orc.seeded_init(m, n, k, 1)
        Mp, per_m, mb = shard_bounds(m, world)
        Np, per_n, nb = shard_bounds(n, world)
        r0, r1 = mb[rank][0], mb[rank][0] + live_rows(m, *mb[rank])
        c0, c1 = nb[rank][0], nb[rank][0] + live_rows(n, *nb[rank])
        v_rows, v_cols = v[r0:r1], v[:, c0:c1]

        def padded(a_t, Lp):                              # component-major [KP][Lp] like the device masters
            t = torch.zeros((KP, Lp), dtype=torch.float64)
            t[:a_t.shape[0], :a_t.shape[1]] = torch.from_numpy(np.ascontiguousarray(a_t))
            return t

        # ---------------- ADMM (admm.py:294-301, 320-324), rho fixed, l1n on H
        rho, lam_h = 50.0, 0.1
        wt, ht = padded(w0.T, Mp), padded(h0, Np)
        wat, hat = wt.clone(), ht.clone()                 # auxiliaries (admm.py:27-28)
        uw, uh = np.zeros((k, r1 - r0)), np.zeros((k, c1 - c0))        # duals stay local
        objs = []
        for _ in range(iters):
            w_aux = wat[:k, :m].numpy().T.copy()
            h_loc, hx = ht[:k, c0:c1].numpy().copy(), None
            h_aux_loc = orc.admm_aux(h_loc, uh, w_aux, v_cols, rho)                 # my columns
            h_new = orc.prox("l1n", h_aux_loc, uh, rho, lam_h)
            uh = uh + h_new - h_aux_loc
            hat[:k, c0:c1] = torch.from_numpy(h_aux_loc); ht[:k, c0:c1] = torch.from_numpy(h_new)
            allgather_windows(hat, per_n); allgather_windows(ht, per_n)
            h_aux = hat[:k, :n].numpy().copy()
            w_loc_t = wt[:k, r0:r1].numpy().copy()
            w_aux_loc_t = orc.admm_aux(w_loc_t, uw, h_aux.T, v_rows.T, rho)         # my rows (transposed problem)
            w_new_t = orc.prox("nn", w_aux_loc_t, uw, rho, 0.0)
            uw = uw + w_new_t - w_aux_loc_t
            wat[:k, r0:r1] = torch.from_numpy(w_aux_loc_t); wt[:k, r0:r1] = torch.from_numpy(w_new_t)
            allgather_windows(wat, per_m); allgather_windows(wt, per_m)
            part = torch.tensor([0.5 * np.sum((v_rows - w_new_t.T @ ht[:k, :n].numpy()) ** 2)])
            allreduce_sum_(part)
            objs.append(float(part))
        tr = orc.admm_run(v, k, w0, h0, rho=rho, reg_w=(0, "nn"), reg_h=(lam_h, "l1n"), min_iter=10**9, max_iter=iters)
        e_admm = (np.linalg.norm(wt[:k, :m].numpy().T - tr.w) / np.linalg.norm(tr.w),
                  np.linalg.norm(ht[:k, :n].numpy() - tr.h) / np.linalg.norm(tr.h),
                  max(abs(a - b) / b for a, b in zip(objs, tr.obj_history[1:])))

        # ---------------- AO-ADMM (ao_admm.py:46-68, 261-270): norms of the residual test all-gathered
        admm_iter, lam_h = 6, 0.1
        wt, ht = padded(w0.T, Mp), padded(h0, Np)
        uw, uh = np.zeros((k, r1 - r0)), np.zeros((k, c1 - c0))
        used_all = []

        def side(a_full, y_loc, x_loc, dual, kind, lam):
            g = a_full.T @ a_full
            rho_ = np.trace(g) / k
            cho = sla.cholesky(g + rho_ * np.eye(k), lower=True)
            aty = a_full.T @ y_loc
            used = admm_iter
            for it in range(admm_iter):
                aux = sla.cho_solve((cho, True), aty + rho_ * (x_loc + dual))
                x_prev = x_loc
                x_loc = orc.prox(kind, aux, dual, rho_, lam)
                dual = dual + x_loc - aux
                mine = torch.tensor([np.sum((x_loc - aux) ** 2), np.sum(x_loc ** 2),
                                     np.sum((x_loc - x_prev) ** 2), np.sum(dual ** 2)])
                full = torch.zeros(4 * world, dtype=torch.float64)
                allgather_flat(full, mine)
                t = full.view(world, 4).sum(0).numpy()
                with np.errstate(divide="ignore", invalid="ignore"):
                    r_, s_ = np.sqrt(t[0]) / np.sqrt(t[1]), np.sqrt(t[2]) / np.sqrt(t[3])
                if r_ < 1e-2 and s_ < 1e-2:               # the same decision on every rank
                    used = it
                    break
            return x_loc, dual, used

        objs = []
        for _ in range(iters):
            w = wt[:k, :m].numpy().T.copy()
            h_loc, uh, used_h = side(w, v_cols, ht[:k, c0:c1].numpy().copy(), uh, "l1n", lam_h)
            ht[:k, c0:c1] = torch.from_numpy(h_loc)
            allgather_windows(ht, per_n)
            h = ht[:k, :n].numpy().copy()
            w_loc_t, uw, used_w = side(h.T, v_rows.T, wt[:k, r0:r1].numpy().copy(), uw, "nn", 0.0)
            wt[:k, r0:r1] = torch.from_numpy(w_loc_t)
            allgather_windows(wt, per_m)
            used_all.append((used_h, used_w))
            part = torch.tensor([0.5 * np.sum((v_rows - w_loc_t.T @ h) ** 2)])
            allreduce_sum_(part)
            objs.append(float(part))
        tr = orc.ao_admm_run(v, k, w0, h0, reg_w=(0, "nn"), reg_h=(lam_h, "l1n"), admm_iter=admm_iter,
                             min_iter=10**9, max_iter=iters)
        e_ao = (np.linalg.norm(wt[:k, :m].numpy().T - tr.w) / np.linalg.norm(tr.w),
                np.linalg.norm(ht[:k, :n].numpy() - tr.h) / np.linalg.norm(tr.h),
                max(abs(a - b) / b for a, b in zip(objs, tr.obj_history[1:])))
        ret[rank] = (e_admm, e_ao, used_all == [tuple(u) for u in tr.inner_breaks])
    finally:
        dist.destroy_process_group()


def test_two_rank_admm_and_ao_admm_exchange_reproduce_unsharded_oracle():
    world = 2
    ctx = mp.get_context("spawn")
    ret = ctx.Manager().dict()
    port = 29950 + (os.getpid() % 40)
    procs = [ctx.Process(target=_admm_worker, args=(r, world, port, ret)) for r in range(world)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(180)
        assert p.exitcode == 0
    for r in range(world):
        e_admm, e_ao, same_breaks = ret[r]
        assert max(e_admm) < 1e-10, e_admm
        assert max(e_ao) < 1e-10, e_ao
        assert same_breaks
