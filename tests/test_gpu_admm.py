"""GPU parity: ADMM and AO-ADMM through the public API / C ABI vs the frozen reference outputs and
the oracle.  Bar: relative Frobenius error of W, H and the objective <= 1e-4, equal stop iteration,
equal inner-loop break pattern (AO-ADMM)."""
import contextlib
import ctypes as C
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


@pytest.mark.parametrize("name", ["admm_l1n_rho100", "admm_stop", "admm_nndsvd", "ao_admm_l1n", "ao_admm_nn_stop"])
def test_matches_reference_golden(name):
    g = load_golden(name)
    r = g["recipe"]
    v = golden_v(r["v"])
    model, out = _run_api(v.copy(), r["k"], r["method"], r["init_seed"], r["kw"])
    res = model.results
    assert res.i == int(g["i"]), "stop iteration differs from the reference"
    assert len(res.obj_history) == len(g["obj_history"])
    obj_err = np.max(np.abs(np.array(res.obj_history) - g["obj_history"]) / g["obj_history"])
    assert obj_err < TOL
    assert rel_fro(res.w, g["w"]) < TOL
    assert rel_fro(res.h, g["h"]) < TOL
    assert np.all(res.w >= 0) and np.all(res.h >= 0)
    if r["method"] == "ao_admm":
        assert out.count("ADMM break after") == int(g["n_inner_breaks"])
    e = res.experiment
    assert e.method == r["method"] and e.prox_h == r["kw"]["reg_h"][1] and e.lambda_h == r["kw"]["reg_h"][0]


@pytest.mark.parametrize("method", ["admm", "ao_admm"])
def test_trajectory_vs_oracle_k64(method):
    """100 iterations at k=64 on planted data from the same init; checked at 10 / 50 / 100."""
    from oracle import nmf_oracle as orc
    from nmf_b200.admm import admm
    from nmf_b200.ao_admm import ao_admm
    m, n, k, iters = 1280, 1024, 64, 100
    v = orc.planted(m, n, k, 0).astype(np.float64)
    w0, h0 = orc.seeded_init(m, n, k, 1)
    snap = lambda i: i in (9, 49, 99)
    if method == "admm":
        kw = dict(rho=100, reg_w=(0, "nn"), reg_h=(0.1, "l1n"))
        tr = orc.admm_run(v, k, w0, h0, min_iter=10**9, max_iter=iters, snap=snap, **kw)
        run = lambda upto: admm(v, k, min_iter=10**9, max_iter=upto, w0=w0, h0=h0, verbose=False,
                                nndsvd_init=(False, "zero"), **kw)
    else:
        kw = dict(reg_w=(0, "nn"), reg_h=(0.1, "l1n"), admm_iter=5)
        tr = orc.ao_admm_run(v, k, w0, h0, min_iter=10**9, max_iter=iters, snap=snap, **kw)
        run = lambda upto: ao_admm(v, k, min_iter=10**9, max_iter=upto, w0=w0, h0=h0, verbose=False,
                                   nndsvd_init=(False, "zero"), **kw)
    snaps = {i: (w, h) for i, w, h in tr.snaps}
    for upto in (10, 50, 100):
        res = run(upto)
        w_ref, h_ref = snaps[upto - 1]
        assert rel_fro(res.w, w_ref) < TOL and rel_fro(res.h, h_ref) < TOL
        nrm = np.sqrt(2 * np.array(res.obj_history))
        nrm_ref = np.sqrt(2 * np.array(tr.obj_history[: upto + 1]))
        assert np.max(np.abs(nrm - nrm_ref) / nrm_ref) < TOL
    if method == "ao_admm":
        assert [tuple(u) for u in ao_admm.last_inner_iterations] == [tuple(u) for u in tr.inner_breaks]


@pytest.mark.parametrize("k", [10, 100, 200])
def test_chol_inverse(k):
    import torch
    from nmf_b200 import _lib
    lib = _lib.load()
    KP = lib.nmfb_padded_components(k)
    rng = np.random.default_rng(k)
    a = rng.random((500, k))
    g = np.zeros((KP, KP))
    g[:k, :k] = a.T @ a
    gd = torch.tensor(g, device="cuda")
    ainv = torch.zeros((KP, KP), dtype=torch.float64, device="cuda")
    scratch = torch.zeros((KP, KP + 1), dtype=torch.float64, device="cuda")
    rho_out = torch.zeros(1, dtype=torch.float64, device="cuda")
    flags = torch.zeros(8, dtype=torch.int32, device="cuda")
    st = torch.cuda.current_stream().cuda_stream
    _lib.check(lib.nmfb_chol_inverse(gd.data_ptr(), k, KP, 1, 0.0, scratch.data_ptr(), ainv.data_ptr(),
                                     rho_out.data_ptr(), flags.data_ptr(), st))
    torch.cuda.synchronize()
    rho = np.trace(g[:k, :k]) / k
    assert abs(rho_out.item() - rho) < 1e-12 * rho
    want = np.linalg.inv(g[:k, :k] + rho * np.eye(k))
    got = ainv.cpu().numpy()
    assert np.linalg.norm(got[:k, :k] - want) / np.linalg.norm(want) < 1e-10
    assert flags[0].item() == 0
    # not positive definite -> flag (scipy.linalg.cholesky raises LinAlgError, ao_admm.py:55)
    gd.zero_()
    _lib.check(lib.nmfb_chol_inverse(gd.data_ptr(), k, KP, 1, 0.0, scratch.data_ptr(), ainv.data_ptr(),
                                     rho_out.data_ptr(), flags.data_ptr(), st))
    torch.cuda.synchronize()
    assert flags[0].item() == 1


def test_ao_admm_collapse_raises_linalgerror():
    """Uniform-random V + l1n drives H to 0, rho = tr(0)/k = 0: the reference raises LinAlgError."""
    from nmf_b200.ao_admm import ao_admm
    v = np.random.default_rng(0).random((512, 384))
    np.random.seed(1)
    with pytest.raises(np.linalg.LinAlgError):
        ao_admm(v, 16, reg_w=(0, "nn"), reg_h=(500.0, "l1n"), admm_iter=5, min_iter=100, max_iter=50,
                nndsvd_init=(False, "zero"), verbose=False)


def test_unknown_prox_and_distance():
    from nmf_b200.admm import admm
    v = np.ones((64, 64))
    with pytest.raises(TypeError):
        admm(v, 4, reg_h=(0.1, "bogus"), nndsvd_init=(False, "zero"), max_iter=2, verbose=False)
    with pytest.raises(TypeError):
        admm(v, 4, distance_type="xx", nndsvd_init=(False, "zero"), max_iter=2, verbose=False)


@pytest.mark.parametrize("k,KP,L", [(10, 64, 700), (128, 128, 20000), (200, 256, 5000)])
def test_fp64_gram_of_the_planes_is_exact(k, KP, L):
    """nmfb_gram_fp64 (FP64 tensor cores, admm.cu): G == float64 Gram of the operand the three bf16 planes hold --
    the operand of the numerator contraction --, symmetric, zero in the padded components."""
    import ctypes as C
    import torch
    from nmf_b200._lib import Factor, check, load
    lib = load()
    dev = torch.device("cuda:0")
    sms = torch.cuda.get_device_properties(dev).multi_processor_count
    Lp = (L + 255) // 256 * 256
    rng = np.random.default_rng(k)
    host = (rng.random((L, k)) * np.exp(rng.normal(0, 3, (1, k)))).astype(np.float32)       # columns of mixed scale
    with torch.cuda.device(dev):
        master = torch.zeros((KP, Lp), dtype=torch.float32, device=dev)
        cpl = torch.zeros((3, 256, Lp), dtype=torch.bfloat16, device=dev)
        rpl = torch.zeros((2, Lp, KP), dtype=torch.bfloat16, device=dev)
        fs = Factor(master.data_ptr(), cpl.data_ptr(), rpl.data_ptr(), L, Lp, 0, 0)
        src = torch.from_numpy(host).to(dev)
        check(lib.nmfb_import_factor(src.data_ptr(), 0, k, 1, k, KP, C.byref(fs), None))
        nbytes = int(lib.nmfb_gram_partial_bytes(k, L, sms))
        scratch = torch.empty(nbytes // 8 + 1, dtype=torch.float64, device=dev)
        g32 = torch.zeros((KP, KP), dtype=torch.float32, device=dev)
        g64 = torch.zeros((KP, KP), dtype=torch.float64, device=dev)
        check(lib.nmfb_gram_fp64(C.byref(fs), KP, scratch.data_ptr(), nbytes, g32.data_ptr(), g64.data_ptr(), None))
        torch.cuda.synchronize(dev)
        operand = cpl[:, :KP].to(torch.float64).sum(0).cpu().numpy()                          # [KP][Lp]
    got, want = g64.cpu().numpy(), operand @ operand.T
    err = np.abs(got - want).max() / np.abs(want).max()
    print(f"k={k} L={L}: max |G - G64| / max |G64| = {err:.1e}")
    assert err < 1e-13 and np.array_equal(got, got.T) and not got[k:].any()
    assert np.array_equal(g32.cpu().numpy(), got.astype(np.float32))
    ref = host.astype(np.float64).T @ host.astype(np.float64)
    assert np.linalg.norm(got[:k, :k] - ref) / np.linalg.norm(ref) < 1e-6                     # planes = 24-bit operand
