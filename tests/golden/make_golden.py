#!/usr/bin/env python3
"""Freeze outputs of the UNMODIFIED reference (/root/reference) as golden fixtures.

Run in the build container only (the GPU box has no /root/reference):

    python tests/golden/make_golden.py

For every case below the reference's own public API is driven
(`NMF(V, k).factorize(method=..., **kw)`) from a fixed global-RNG seed, and the
returned `Results` (w, h, i, obj_history) are written to tests/golden/<case>.npz
together with the exact recipe (data generator + seed + kwargs), so that the
oracle (oracle/nmf_oracle.py) and the CUDA path can replay it without the reference.
"""
import contextlib
import io
import json
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, "/root/reference")

from nmf import NMF  # noqa: E402  (the reference)
from oracle.nmf_oracle import planted, planted_with_zeros  # noqa: E402


def make_v(spec):
    kind = spec["kind"]
    if kind == "uniform":            # un-rounded fp64 uniform data
        return np.random.default_rng(spec["seed"]).random((spec["m"], spec["n"]))
    if kind == "planted":
        return planted(spec["m"], spec["n"], spec["k"], spec["seed"]).astype(np.float64)
    if kind == "zeros":              # exact zeros (30 %), an all-zero row and an all-zero column
        return planted_with_zeros(spec["m"], spec["n"], spec["k"], spec["seed"]).astype(np.float64)
    raise ValueError(kind)


CASES = {
    # config C1 of BASELINE.json (SURVEY 8c golden values: i=199, obj[0]=9542736.41..., obj[-1]=19939.29...)
    "c1_mur_eu": dict(
        v=dict(kind="uniform", m=1000, n=500, seed=0), k=10, init_seed=0, method="mur",
        kw=dict(distance_type="eu", min_iter=100, max_iter=200, tol1=1e-5, tol2=1e-5)),
    "mur_eu_reg": dict(
        v=dict(kind="planted", m=384, n=320, k=12, seed=3), k=12, init_seed=4, method="mur",
        kw=dict(distance_type="eu", min_iter=20, max_iter=80, tol1=1e-5, tol2=1e-5,
                lambda_w=0.05, lambda_h=0.1)),
    "mur_eu_stop": dict(
        v=dict(kind="planted", m=256, n=192, k=8, seed=5), k=8, init_seed=6, method="mur",
        kw=dict(distance_type="eu", min_iter=10, max_iter=5000, tol1=1.0, tol2=1e-1)),
    "mur_kl": dict(
        v=dict(kind="planted", m=300, n=200, k=8, seed=7), k=8, init_seed=8, method="mur",
        kw=dict(distance_type="kl", min_iter=100, max_iter=60, tol1=1e-5, tol2=1e-5)),
    "mur_kl_reg": dict(
        v=dict(kind="uniform", m=200, n=260, seed=9), k=6, init_seed=10, method="mur",
        kw=dict(distance_type="kl", min_iter=100, max_iter=40, tol1=1e-5, tol2=1e-5,
                lambda_w=0.2, lambda_h=0.3)),
    "admm_l1n_rho100": dict(
        v=dict(kind="planted", m=512, n=384, k=16, seed=11), k=16, init_seed=12, method="admm",
        kw=dict(rho=100, reg_w=(0, "nn"), reg_h=(0.1, "l1n"), min_iter=100, max_iter=60,
                tol1=1e-3, tol2=1e-3, nndsvd_init=(False, "zero"))),
    "admm_stop": dict(
        v=dict(kind="planted", m=320, n=256, k=8, seed=13), k=8, init_seed=14, method="admm",
        kw=dict(rho=1, reg_w=(0, "nn"), reg_h=(0, "l2n"), min_iter=10, max_iter=500,
                tol1=1e-3, tol2=1e-3, nndsvd_init=(False, "zero"))),
    "admm_nndsvd": dict(
        v=dict(kind="planted", m=200, n=160, k=6, seed=15), k=6, init_seed=16, method="admm",
        kw=dict(rho=10, reg_w=(0.05, "l1n"), reg_h=(0, "nn"), min_iter=100, max_iter=30,
                tol1=1e-3, tol2=1e-3, nndsvd_init=(True, "zero"))),
    "admm_l2n": dict(
        v=dict(kind="planted", m=320, n=224, k=12, seed=21), k=12, init_seed=22, method="admm",
        kw=dict(rho=50, reg_w=(0.3, "l2n"), reg_h=(0.5, "l2n"), min_iter=100, max_iter=40,
                tol1=1e-3, tol2=1e-3, nndsvd_init=(False, "zero"))),
    "admm_kl": dict(
        v=dict(kind="planted", m=240, n=200, k=8, seed=23), k=8, init_seed=24, method="admm",
        kw=dict(rho=20, distance_type="kl", reg_w=(0, "nn"), reg_h=(0.05, "l1n"), min_iter=100, max_iter=40,
                tol1=1e-3, tol2=1e-3, nndsvd_init=(False, "zero"))),
    "ao_admm_kl": dict(
        v=dict(kind="planted", m=200, n=240, k=8, seed=25), k=8, init_seed=26, method="ao_admm",
        kw=dict(distance_type="kl", reg_w=(0, "nn"), reg_h=(0.05, "l1n"), min_iter=100, max_iter=25,
                admm_iter=6, tol1=1e-3, tol2=1e-3, nndsvd_init=(False, "zero"))),
    # ---- edge cases: exact zeros in V, a single component, the other NNDSVD fill variants
    "mur_kl_zeros": dict(
        v=dict(kind="zeros", m=180, n=150, k=5, seed=27), k=5, init_seed=28, method="mur",
        kw=dict(distance_type="kl", min_iter=100, max_iter=40, tol1=1e-5, tol2=1e-5)),
    "mur_eu_zeros": dict(
        v=dict(kind="zeros", m=150, n=180, k=5, seed=29), k=5, init_seed=30, method="mur",
        kw=dict(distance_type="eu", min_iter=100, max_iter=40, tol1=1e-5, tol2=1e-5)),
    "mur_eu_k1": dict(
        v=dict(kind="uniform", m=37, n=23, seed=31), k=1, init_seed=32, method="mur",
        kw=dict(distance_type="eu", min_iter=100, max_iter=30, tol1=1e-5, tol2=1e-5)),
    "mur_kl_k1": dict(
        v=dict(kind="uniform", m=23, n=37, seed=33), k=1, init_seed=34, method="mur",
        kw=dict(distance_type="kl", min_iter=100, max_iter=30, tol1=1e-5, tol2=1e-5)),
    "admm_nndsvd_mean": dict(
        v=dict(kind="planted", m=160, n=120, k=5, seed=35), k=5, init_seed=36, method="admm",
        kw=dict(rho=10, reg_w=(0, "nn"), reg_h=(0.05, "l1n"), min_iter=100, max_iter=20,
                tol1=1e-3, tol2=1e-3, nndsvd_init=(True, "mean"))),
    "ao_admm_nndsvd_random": dict(
        v=dict(kind="planted", m=120, n=160, k=5, seed=37), k=5, init_seed=38, method="ao_admm",
        kw=dict(reg_w=(0, "nn"), reg_h=(0.05, "l1n"), min_iter=100, max_iter=15, admm_iter=5,
                tol1=1e-3, tol2=1e-3, nndsvd_init=(True, "random"))),
    "ao_admm_l1n": dict(
        v=dict(kind="planted", m=512, n=384, k=16, seed=17), k=16, init_seed=18, method="ao_admm",
        kw=dict(reg_w=(0, "nn"), reg_h=(0.1, "l1n"), min_iter=100, max_iter=40, admm_iter=5,
                tol1=1e-3, tol2=1e-3, nndsvd_init=(False, "zero"))),
    "ao_admm_nn_stop": dict(
        v=dict(kind="planted", m=256, n=320, k=8, seed=19), k=8, init_seed=20, method="ao_admm",
        kw=dict(reg_w=(0, "nn"), reg_h=(0, "nn"), min_iter=10, max_iter=400, admm_iter=10,
                tol1=1e-3, tol2=1e-3, nndsvd_init=(False, "zero"))),
}


def main():
    only = set(sys.argv[1:])
    for name, case in CASES.items():
        if only and name not in only:
            continue
        v = make_v(case["v"])
        np.random.seed(case["init_seed"])
        model = NMF(v.copy(), case["k"])
        sink = io.StringIO()
        with contextlib.redirect_stdout(sink):
            model.factorize(method=case["method"], **case["kw"])
        res = model.results
        n_inner_breaks = sink.getvalue().count("ADMM break after")
        recipe = dict(v=case["v"], k=case["k"], init_seed=case["init_seed"],
                      method=case["method"], kw=case["kw"])
        np.savez_compressed(
            os.path.join(HERE, name + ".npz"),
            w=res.w, h=res.h, i=np.int64(res.i),
            obj_history=np.asarray(res.obj_history, dtype=np.float64),
            n_inner_breaks=np.int64(n_inner_breaks),
            recipe=json.dumps(recipe))
        print(f"{name:18s} i={res.i:5d} len(obj)={len(res.obj_history):5d} "
              f"obj0={res.obj_history[0]:.6f} objN={res.obj_history[-1]:.9f} "
              f"inner_breaks={n_inner_breaks}")


if __name__ == "__main__":
    main()
