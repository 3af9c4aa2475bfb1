"""GPU parity: MUR Kullback-Leibler (fused GEMM -> quotient -> GEMM kernel) vs the frozen reference
outputs and the oracle.  Bar: relative Frobenius error of W, H and the objective <= 1e-4."""
import contextlib
import io

import numpy as np
import pytest

from conftest import golden_v, load_golden, rel_fro

pytestmark = pytest.mark.gpu

TOL = 1e-4


@pytest.mark.parametrize("name", ["mur_kl", "mur_kl_reg"])
def test_mur_kl_matches_reference_golden(name):
    from nmf_b200 import NMF
    g = load_golden(name)
    r = g["recipe"]
    v = golden_v(r["v"])
    np.random.seed(r["init_seed"])
    model = NMF(v.copy(), r["k"])
    with contextlib.redirect_stdout(io.StringIO()):
        model.factorize(method="mur", **r["kw"])
    res = model.results
    assert res.i == int(g["i"]) and len(res.obj_history) == len(g["obj_history"])
    obj_err = np.max(np.abs(np.array(res.obj_history) - g["obj_history"]) / np.abs(g["obj_history"]))
    assert obj_err < TOL
    assert rel_fro(res.w, g["w"]) < TOL and rel_fro(res.h, g["h"]) < TOL


@pytest.mark.parametrize("shape", [(640, 512, 24), (1536, 1280, 100)])
def test_single_kl_step_vs_oracle(shape):
    """One outer iteration from a given start: isolates the fused kernel from trajectory effects."""
    from oracle import nmf_oracle as orc
    from nmf_b200.mur import mur
    m, n, k = shape
    v = orc.planted(m, n, k, 0).astype(np.float64)
    w0, h0 = orc.seeded_init(m, n, k, 1)
    res = mur(v, k, distance_type="kl", min_iter=10**9, max_iter=1, w0=w0, h0=h0, verbose=False)
    tr = orc.mur_run(v, k, w0, h0, distance_type="kl", min_iter=10**9, max_iter=1)
    assert rel_fro(res.w, tr.w) < 2e-6 and rel_fro(res.h, tr.h) < 2e-6
    assert abs(res.obj_history[0] - tr.obj_history[0]) / tr.obj_history[0] < 1e-5
    assert abs(res.obj_history[1] - tr.obj_history[1]) / tr.obj_history[1] < 1e-5


def test_mur_kl_trajectory_vs_oracle_k100():
    from oracle import nmf_oracle as orc
    from nmf_b200.mur import mur
    m, n, k, iters = 1536, 1280, 100, 100
    v = orc.planted(m, n, k, 0).astype(np.float64)
    w0, h0 = orc.seeded_init(m, n, k, 1)
    tr = orc.mur_run(v, k, w0, h0, distance_type="kl", min_iter=10**9, max_iter=iters,
                     snap=lambda i: i in (9, 49, 99))
    snaps = {i: (w, h) for i, w, h in tr.snaps}
    for upto in (10, 50, 100):
        res = mur(v, k, distance_type="kl", min_iter=10**9, max_iter=upto, w0=w0, h0=h0, verbose=False)
        w_ref, h_ref = snaps[upto - 1]
        assert rel_fro(res.w, w_ref) < TOL and rel_fro(res.h, h_ref) < TOL
        err = np.max(np.abs(np.array(res.obj_history) - tr.obj_history[: upto + 1]) / np.array(tr.obj_history[: upto + 1]))
        assert err < TOL


def test_kl_rejects_k_above_128():
    from nmf_b200._lib import NmfbError
    from nmf_b200.mur import mur
    v = np.random.default_rng(0).random((512, 512))
    with pytest.raises(NmfbError):
        mur(v, 200, distance_type="kl", min_iter=100, max_iter=2, verbose=False)
