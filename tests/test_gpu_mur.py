"""GPU parity: MUR solvers through the public API / C ABI vs the oracle and the frozen reference
outputs.  Bar (BASELINE.json north_star): relative Frobenius error of W, H and of the objective
<= 1e-4 per iteration over 100 iterations; stop iteration equal to the reference's."""
import contextlib
import io

import numpy as np
import pytest

from conftest import golden_v, load_golden, rel_fro

pytestmark = pytest.mark.gpu

TOL = 1e-4


def _run_api(v, k, method, init_seed, kw):
    from nmf_b200 import NMF
    np.random.seed(init_seed)
    model = NMF(v, k)
    sink = io.StringIO()
    with contextlib.redirect_stdout(sink):
        model.factorize(method=method, **kw)
    return model, sink.getvalue()


@pytest.mark.parametrize("name", ["c1_mur_eu", "mur_eu_reg", "mur_eu_stop"])
def test_mur_eu_matches_reference_golden(name):
    g = load_golden(name)
    r = g["recipe"]
    v = golden_v(r["v"])
    model, out = _run_api(v.copy(), r["k"], "mur", r["init_seed"], r["kw"])
    res = model.results
    assert res.i == int(g["i"]), "stop iteration differs from the reference"
    assert len(res.obj_history) == len(g["obj_history"])
    obj_err = np.max(np.abs(np.array(res.obj_history) - g["obj_history"]) / g["obj_history"])
    assert obj_err < TOL
    assert rel_fro(res.w, g["w"]) < TOL
    assert rel_fro(res.h, g["h"]) < TOL
    assert res.w.dtype == np.float64 and res.h.dtype == np.float64
    assert "Factorization done." in out and "[0]: " in out
    assert model.w is res.w and model.h is res.h


def test_mur_eu_trajectory_vs_oracle():
    """Per-iteration check (every 10th iteration) on planted data at k=64."""
    from oracle import nmf_oracle as orc
    from nmf_b200.mur import mur
    m, n, k, iters = 1536, 1280, 64, 100
    v = orc.planted(m, n, k, 0).astype(np.float64)
    w0, h0 = orc.seeded_init(m, n, k, 1)
    tr = orc.mur_run(v, k, w0, h0, distance_type="eu", min_iter=10**9, max_iter=iters,
                     snap=lambda i: i % 10 == 9)
    snaps = {i: (w, h) for i, w, h in tr.snaps}
    for upto in (10, 50, 100):
        res = mur(v, k, distance_type="eu", min_iter=10**9, max_iter=upto, w0=w0, h0=h0, verbose=False)
        w_ref, h_ref = snaps[upto - 1]
        assert rel_fro(res.w, w_ref) < TOL and rel_fro(res.h, h_ref) < TOL
        # ||V - WH|| (not squared) within 1e-4
        nrm = np.sqrt(2 * np.array(res.obj_history))
        nrm_ref = np.sqrt(2 * np.array(tr.obj_history[: upto + 1]))
        assert np.max(np.abs(nrm - nrm_ref) / nrm_ref) < TOL


def test_objective_kernels_agree():
    """tcgen05 residual kernel vs the CUDA-core kernel vs numpy."""
    from oracle import nmf_oracle as orc
    from nmf_b200.engine import Engine
    m, n, k = 900, 1100, 48
    v = orc.planted(m, n, k, 2)
    w0, h0 = orc.seeded_init(m, n, k, 3)
    eng = Engine(m, n, k)
    eng.upload_data(v)
    eng.set_factors(w0, h0)
    eng.objective_eu(slot=0, impl=0)
    eng.objective_eu(slot=1, impl=1)
    a, b = eng.read_scalar(0), eng.read_scalar(1)
    want = orc.objective(v.astype(np.float64), w0.astype(np.float32).astype(np.float64)
                         @ h0.astype(np.float32).astype(np.float64))
    assert abs(a - want) / want < 1e-5 and abs(b - want) / want < 1e-5


def test_negative_data_is_shifted_in_place():
    from nmf_b200.mur import mur
    rng = np.random.default_rng(0)
    v = rng.random((200, 150)) - 0.25
    vmin = v.min()
    v_in = v.copy()
    res = mur(v_in, 5, distance_type="eu", min_iter=100, max_iter=5, verbose=False)
    assert np.isclose(v_in.min(), 0.0) and np.allclose(v_in, v - vmin)     # mur.py:99-102
    assert np.all(res.w >= 0) and np.all(res.h >= 0)


def test_checkpoint_and_resume(tmp_path):
    """A checkpoint holds the factors of its iteration; MUR resumed from it retraces the original run."""
    from oracle import nmf_oracle as orc
    from nmf_b200.mur import mur
    m, n, k = 512, 384, 16
    v = orc.planted(m, n, k, 0).astype(np.float64)
    w0, h0 = orc.seeded_init(m, n, k, 1)
    path = str(tmp_path / "ckpt")
    full = mur(v, k, distance_type="eu", min_iter=10**9, max_iter=20, w0=w0, h0=h0, verbose=False,
               checkpoint_every=8, checkpoint_path=path)
    z = np.load(path + ".npz", allow_pickle=True)
    assert int(z["i"]) == 16 and len(z["obj_history"]) == 18            # last multiple of 8 below 20
    assert sorted(z.files) == ["experiment", "h", "i", "obj_history", "w"]
    tr = orc.mur_run(v, k, w0, h0, distance_type="eu", min_iter=10**9, max_iter=17)
    assert rel_fro(z["w"], tr.w) < TOL and rel_fro(z["h"], tr.h) < TOL
    rest = mur(v, k, distance_type="eu", min_iter=10**9, max_iter=3, w0=z["w"], h0=z["h"], verbose=False)
    assert rel_fro(rest.w, full.w) < 1e-5 and rel_fro(rest.h, full.h) < 1e-5
