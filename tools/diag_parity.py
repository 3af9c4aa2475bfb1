"""Where does an objective-history mismatch come from?  For a few (solver, shape) pairs: the GPU run's objective history
against the fp64 oracle per iteration, and -- at a few iteration counts -- the objective of the GPU's W, H re-evaluated
in fp64 on the host (trajectory error) next to the value the GPU reported (evaluation error).  1 GPU."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np

from oracle import nmf_oracle as orc
from nmf_b200.admm import admm
from nmf_b200.ao_admm import ao_admm

rel = lambda a, b: float(np.linalg.norm(a - b) / np.linalg.norm(b))
CASES = [
    ("admm-kl k=64", admm, orc.admm_run, 2048, 1536, 64, 40,
     dict(rho=100, distance_type="kl", reg_w=(0, "nn"), reg_h=(0.05, "l1n"))),
    ("ao-admm-kl k=128", ao_admm, orc.ao_admm_run, 1500, 2500, 128, 30,
     dict(distance_type="kl", reg_w=(0, "nn"), reg_h=(0.05, "l1n"), admm_iter=4)),
    ("ao-admm eu k=64", ao_admm, orc.ao_admm_run, 2048, 1536, 64, 40,
     dict(reg_w=(0, "nn"), reg_h=(0.1, "l1n"), admm_iter=5)),
]
only = sys.argv[1:]
for name, gpu_fn, orc_fn, m, n, k, iters, kw in CASES:
    if only and not any(o in name for o in only):
        continue
    v = orc.planted(m, n, k, 0).astype(np.float64)
    w0, h0 = orc.seeded_init(m, n, k, 1)
    kind = kw.get("distance_type", "eu")
    tr = orc_fn(v, k, w0, h0, min_iter=10**9, max_iter=iters, snap=lambda i: True, **kw)
    res = gpu_fn(v, k, min_iter=10**9, max_iter=iters, w0=w0, h0=h0, verbose=False, nndsvd_init=(False, "zero"), **kw)
    err = np.abs(np.array(res.obj_history) - np.array(tr.obj_history)) / np.array(tr.obj_history)
    print(f"{name}: max relObj {err.max():.2e} at iteration {int(err.argmax()) - 1}; per iteration:",
          " ".join(f"{e:.1e}" for e in err), flush=True)
    for stop in sorted({max(1, int(err.argmax()) - 1), int(err.argmax()), iters // 2, iters}):
        stop = min(max(stop, 1), iters)
        r = gpu_fn(v, k, min_iter=10**9, max_iter=stop, w0=w0, h0=h0, verbose=False, nndsvd_init=(False, "zero"), **kw)
        _, wo, ho = tr.snaps[stop - 1]
        host = orc.objective(v, r.w @ r.h, kind)
        print(f"   after {stop:3d} iterations: relW {rel(r.w, wo):.2e} relH {rel(r.h, ho):.2e}  oracle obj {tr.obj_history[stop]:.6e}  "
              f"GPU reported {r.obj_history[-1]:.6e} (rel {abs(r.obj_history[-1] - tr.obj_history[stop]) / tr.obj_history[stop]:.2e})  "
              f"GPU W,H evaluated in fp64 {host:.6e} (rel {abs(host - tr.obj_history[stop]) / tr.obj_history[stop]:.2e})", flush=True)
