#!/usr/bin/env python3
"""A few iterations of every solver and branch at a small ragged shape through the public API: the target of
tools/sanitize.sh (compute-sanitizer memcheck / racecheck / synccheck / initcheck).  Prints one line per run."""
import contextlib
import io
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from nmf_b200 import NMF  # noqa: E402

ITERS = int(os.environ.get("SANITIZE_ITERS", "3"))
rng = np.random.default_rng(0)
for m, n, k in ((600, 500, 10), (300, 700, 100)):
    v = rng.random((m, n)).astype(np.float32) + 0.01
    runs = [("mur", dict(distance_type="eu")), ("mur", dict(distance_type="kl")),
            ("admm", dict(rho=10.0, reg_w=(0, "nn"), reg_h=(0.1, "l1n"))),
            ("admm", dict(rho=10.0, reg_w=(0.1, "l2n"), reg_h=(0.1, "l2n"))),
            ("admm", dict(rho=10.0, distance_type="kl")),
            ("ao_admm", dict(reg_w=(0, "nn"), reg_h=(0.1, "l1n"), admm_iter=3)),
            ("ao_admm", dict(distance_type="kl", admm_iter=2))]
    for method, kw in runs:
        np.random.seed(1)
        model = NMF(v, k)
        with contextlib.redirect_stdout(io.StringIO()):
            model.factorize(method=method, min_iter=ITERS, max_iter=ITERS, **kw)
        res = model.results
        print(f"{m}x{n} k={k} {method} {kw.get('distance_type', 'eu')}: i={res.i} obj={res.obj_history[-1]:.6e}",
              flush=True)
