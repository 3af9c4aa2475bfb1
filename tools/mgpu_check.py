"""Run under torchrun: sharded MUR-eu through the public driver vs the oracle (rank 0 checks)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
import torch.distributed as dist

rank, world, lr = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(lr)
dist.init_process_group("nccl", device_id=torch.device("cuda", lr))
from oracle import nmf_oracle as orc
from nmf_b200.mur import mur

ok = True
for (m, n, k, iters) in [(1000, 700, 10, 30), (2048, 1536, 64, 40), (1500, 2500, 128, 30)]:
    v = orc.planted(m, n, k, 0).astype(np.float64)
    w0, h0 = orc.seeded_init(m, n, k, 1)
    res = mur(v, k, distance_type="eu", min_iter=10**9, max_iter=iters, w0=w0, h0=h0, verbose=False, comm=True)
    if rank == 0:
        tr = orc.mur_run(v, k, w0, h0, distance_type="eu", min_iter=10**9, max_iter=iters)
        rel = lambda a, b: float(np.linalg.norm(a - b) / np.linalg.norm(b))
        ew, eh = rel(res.w, tr.w), rel(res.h, tr.h)
        eo = max(abs(a - b) / b for a, b in zip(res.obj_history, tr.obj_history))
        good = max(ew, eh, eo) < 1e-4
        ok &= good
        print(f"sharded x{world} mur-eu {'OK ' if good else 'BAD'} m={m} n={n} k={k} relW={ew:.2e} relH={eh:.2e} relObj={eo:.2e}", flush=True)
    # stop test must fire at the same iteration on every rank
    res2 = mur(v, k, distance_type="eu", min_iter=5, max_iter=500, tol1=1.0, tol2=1e-1, w0=w0, h0=h0, verbose=False, comm=True)
    its = torch.tensor([res2.i], device="cuda")
    gathered = [torch.zeros_like(its) for _ in range(world)]
    dist.all_gather(gathered, its)
    if rank == 0:
        tr2 = orc.mur_run(v, k, w0, h0, distance_type="eu", min_iter=5, max_iter=500, tol1=1.0, tol2=1e-1)
        same = all(int(g) == tr2.i for g in gathered)
        ok &= same
        print(f"   stop iteration ranks={[int(g) for g in gathered]} oracle={tr2.i} {'OK' if same else 'BAD'}", flush=True)
if rank == 0:
    print("MGPU CHECK", "PASS" if ok else "FAIL", flush=True)
dist.barrier()
dist.destroy_process_group()
