#!/usr/bin/env python3
"""Mid-size and full-size goldens from the UNMODIFIED reference (/root/reference), SURVEY.md 8d sizes.

Run in the build container only (the GPU box has no /root/reference; minutes of CPU per case):

    python tests/golden/make_golden_mid.py [case ...]

`mid_*`: 100 iterations of `NMF(V, k).factorize(...)` at 8192^2 k=64 (C2 stand-in), 4000x10000 k=128 (C3),
16384x4096 k=128 (C4, ADMM and AO-ADMM).  `full_*`: 5 iterations at the BASELINE shapes the container's
62 GB hold in float64 (C2 32768^2 k=64, C3 20000x50000 k=128, C4 65536x16384 k=128; the 65536^2 headline and
C5 do not fit: 34 / 275 GB of V alone).  Frozen per case: obj_history (every iteration, float64) and the final
W, H as float32 on a fixed row / column subset (every `stride`-th line plus the last 256: contraction tails,
last tiles) so that the fixtures stay a few MB; the subset indices are rebuilt from the recipe.
"""
import contextlib
import io
import json
import os
import sys
import time

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, "/root/reference")

from nmf import NMF  # noqa: E402  (the reference)
from oracle.nmf_oracle import planted  # noqa: E402

ADMM_KW = dict(rho=100, reg_w=(0, "nn"), reg_h=(0.1, "l1n"), tol1=1e-3, tol2=1e-3, nndsvd_init=(False, "zero"))
AO_KW = dict(reg_w=(0, "nn"), reg_h=(0.1, "l1n"), admm_iter=5, tol1=1e-3, tol2=1e-3, nndsvd_init=(False, "zero"))
MUR_KW = dict(tol1=1e-5, tol2=1e-5)

CASES = {
    "mid_mur_eu": dict(m=8192, n=8192, k=64, seed=0, init_seed=1, method="mur", stride=4,
                       kw=dict(distance_type="eu", min_iter=10**9, max_iter=100, **MUR_KW)),
    "mid_admm": dict(m=16384, n=4096, k=128, seed=0, init_seed=1, method="admm", stride=8,
                     kw=dict(min_iter=10**9, max_iter=100, **ADMM_KW)),
    "mid_ao_admm": dict(m=16384, n=4096, k=128, seed=0, init_seed=1, method="ao_admm", stride=8,
                        kw=dict(min_iter=10**9, max_iter=100, **AO_KW)),
    "mid_mur_kl": dict(m=4000, n=10000, k=128, seed=0, init_seed=1, method="mur", stride=4,
                       kw=dict(distance_type="kl", min_iter=10**9, max_iter=100, **MUR_KW)),
    "mid_admm_stop": dict(m=16384, n=4096, k=128, seed=0, init_seed=1, method="admm", stride=8,
                          kw=dict(min_iter=10, max_iter=100, **ADMM_KW)),
    "full_c2_mur_eu": dict(m=32768, n=32768, k=64, seed=0, init_seed=1, method="mur", stride=32,
                           kw=dict(distance_type="eu", min_iter=10**9, max_iter=5, **MUR_KW)),
    "full_c4_admm": dict(m=65536, n=16384, k=128, seed=0, init_seed=1, method="admm", stride=32,
                         kw=dict(min_iter=10**9, max_iter=5, **ADMM_KW)),
    "full_c4_ao_admm": dict(m=65536, n=16384, k=128, seed=0, init_seed=1, method="ao_admm", stride=32,
                            kw=dict(min_iter=10**9, max_iter=5, **AO_KW)),
    "full_c3_mur_kl": dict(m=20000, n=50000, k=128, seed=0, init_seed=1, method="mur", stride=32,
                           kw=dict(distance_type="kl", min_iter=10**9, max_iter=5, **MUR_KW)),
}


def subset(length, stride):
    """The frozen lines of a factor: every stride-th plus the last 256."""
    return np.union1d(np.arange(0, length, stride), np.arange(max(0, length - 256), length))


def main():
    only = set(sys.argv[1:])
    for name, case in CASES.items():
        if only and name not in only:
            continue
        t0 = time.time()
        m, n, k = case["m"], case["n"], case["k"]
        v = planted(m, n, k, case["seed"]).astype(np.float64)
        np.random.seed(case["init_seed"])
        model = NMF(v, k)
        sink = io.StringIO()
        with contextlib.redirect_stdout(sink):
            model.factorize(method=case["method"], **case["kw"])
        res = model.results
        n_inner_breaks = sink.getvalue().count("ADMM break after")
        rows, cols = subset(m, case["stride"]), subset(n, case["stride"])
        recipe = dict(v=dict(kind="planted", m=m, n=n, k=k, seed=case["seed"]), k=k, init_seed=case["init_seed"],
                      method=case["method"], kw=case["kw"], stride=case["stride"])
        np.savez_compressed(
            os.path.join(HERE, name + ".npz"),
            w_rows=res.w[rows].astype(np.float32), h_cols=res.h[:, cols].astype(np.float32),
            w_norm=np.float64(np.linalg.norm(res.w)), h_norm=np.float64(np.linalg.norm(res.h)),
            i=np.int64(res.i), obj_history=np.asarray(res.obj_history, dtype=np.float64),
            n_inner_breaks=np.int64(n_inner_breaks), recipe=json.dumps(recipe))
        print(f"{name:18s} i={res.i:5d} len(obj)={len(res.obj_history):5d} obj0={res.obj_history[0]:.6f} "
              f"objN={res.obj_history[-1]:.9f} inner_breaks={n_inner_breaks} ({time.time() - t0:.0f} s)", flush=True)
        del model, v, res


if __name__ == "__main__":
    main()
