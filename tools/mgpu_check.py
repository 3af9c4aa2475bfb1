"""Run under torchrun: every sharded solver path through the public drivers vs the oracle (rank 0
checks), plus the agreement of the stop iteration across ranks.

    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 tools/mgpu_check.py
"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
import torch.distributed as dist

rank, world, lr = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(lr)
dist.init_process_group("nccl", device_id=torch.device("cuda", lr))
from oracle import nmf_oracle as orc
from nmf_b200.admm import admm
from nmf_b200.ao_admm import ao_admm
from nmf_b200.mur import mur

TOL = 1e-4
STOP_MAX = 500 if world <= 2 else 60      # the fp64 oracle on rank 0 is the slow part at the larger shapes
ok = True
rel = lambda a, b: float(np.linalg.norm(a - b) / np.linalg.norm(b))


def report(tag, res, tr, shape):
    global ok
    if rank != 0:
        return
    ew, eh = rel(res.w, tr.w), rel(res.h, tr.h)
    eo = max(abs(a - b) / b for a, b in zip(res.obj_history, tr.obj_history))
    good = max(ew, eh, eo) < TOL and res.i == tr.i and len(res.obj_history) == len(tr.obj_history)
    ok &= good
    print(f"sharded x{world} {tag} {'OK ' if good else 'BAD'} {shape} relW={ew:.2e} relH={eh:.2e} relObj={eo:.2e} "
          f"i={res.i}/{tr.i}", flush=True)


def same_on_all_ranks(value):
    t = torch.tensor([value], device="cuda")
    parts = [torch.zeros_like(t) for _ in range(world)]
    dist.all_gather(parts, t)
    return [int(p) for p in parts]


SHAPES = [(1000, 700, 10, 30), (2048, 1536, 64, 40), (1500, 2500, 128, 30)] if world <= 2 else \
    [(3700, 3900, 10, 30), (4096, 6144, 64, 40), (7500, 3800, 128, 30)]      # every rank keeps live rows / columns
for (m, n, k, iters) in SHAPES:
    v = orc.planted(m, n, k, 0).astype(np.float64)
    w0, h0 = orc.seeded_init(m, n, k, 1)
    shape = f"m={m} n={n} k={k}"
    res = mur(v, k, distance_type="eu", min_iter=10**9, max_iter=iters, w0=w0, h0=h0, verbose=False, comm=True)
    report("mur-eu ", res, orc.mur_run(v, k, w0, h0, distance_type="eu", min_iter=10**9, max_iter=iters), shape)
    # stop test must fire at the same iteration on every rank
    res2 = mur(v, k, distance_type="eu", min_iter=5, max_iter=STOP_MAX, tol1=1.0, tol2=1e-1, w0=w0, h0=h0, verbose=False, comm=True)
    its = same_on_all_ranks(res2.i)
    if rank == 0:
        tr2 = orc.mur_run(v, k, w0, h0, distance_type="eu", min_iter=5, max_iter=STOP_MAX, tol1=1.0, tol2=1e-1)
        same = all(g == tr2.i for g in its)
        ok &= same
        print(f"   stop iteration ranks={its} oracle={tr2.i} {'OK' if same else 'BAD'}", flush=True)

    res = mur(v, k, distance_type="kl", min_iter=10**9, max_iter=iters, w0=w0, h0=h0, verbose=False, comm=True)
    report("mur-kl ", res, orc.mur_run(v, k, w0, h0, distance_type="kl", min_iter=10**9, max_iter=iters), shape)

    kw = dict(rho=100, reg_w=(0, "nn"), reg_h=(0.1, "l1n"), min_iter=10**9, max_iter=iters)
    res = admm(v, k, w0=w0, h0=h0, verbose=False, comm=True, nndsvd_init=(False, "zero"), **kw)
    report("admm   ", res, orc.admm_run(v, k, w0, h0, **kw), shape)

    kw = dict(reg_w=(0, "nn"), reg_h=(0.1, "l1n"), admm_iter=5, min_iter=10**9, max_iter=iters)
    res = ao_admm(v, k, w0=w0, h0=h0, verbose=False, comm=True, nndsvd_init=(False, "zero"), **kw)
    tr = orc.ao_admm_run(v, k, w0, h0, **kw)
    report("ao-admm", res, tr, shape)
    inner = ao_admm.last_inner_iterations
    if rank == 0 and True:
        same = [tuple(x) for x in inner] == [tuple(x) for x in tr.inner_breaks]
        ok &= same
        print(f"   inner-iteration pattern {'OK' if same else 'BAD'}", flush=True)
    if os.environ.get("MGPU_CHECK_PENDING", "0") == "1":
        # paths whose first GPU run was pending at the end of round 1: l2n prox, Kullback-Leibler branch
        kw = dict(rho=50, reg_w=(0.3, "l2n"), reg_h=(0.5, "l2n"), min_iter=10**9, max_iter=iters)
        if k <= 128:
            res = admm(v, k, w0=w0, h0=h0, verbose=False, comm=True, nndsvd_init=(False, "zero"), **kw)
            report("admm-l2n", res, orc.admm_run(v, k, w0, h0, **kw), shape)
        kw = dict(rho=100, distance_type="kl", reg_w=(0, "nn"), reg_h=(0.05, "l1n"), min_iter=10**9, max_iter=iters)
        res = admm(v, k, w0=w0, h0=h0, verbose=False, comm=True, nndsvd_init=(False, "zero"), **kw)
        report("admm-kl", res, orc.admm_run(v, k, w0, h0, **kw), shape)
        kw = dict(distance_type="kl", reg_w=(0, "nn"), reg_h=(0.05, "l1n"), admm_iter=4, min_iter=10**9,
                  max_iter=iters)
        res = ao_admm(v, k, w0=w0, h0=h0, verbose=False, comm=True, nndsvd_init=(False, "zero"), **kw)
        report("ao-admm-kl", res, orc.ao_admm_run(v, k, w0, h0, **kw), shape)
if rank == 0:
    print("MGPU CHECK", "PASS" if ok else "FAIL", flush=True)
dist.barrier()
dist.destroy_process_group()
