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
