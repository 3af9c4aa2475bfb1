"""GPU parity of the SURVEY.md 8(f) "next" rows -- the l2n (Tikhonov) prox with lambda > 0 (admm.py:141-156),
the Kullback-Leibler branch of admm / ao_admm --, of BASELINE.json's full-size configurations through
size-independent properties, and of edge cases (zeros in V, one component, odd shapes, input dtypes).
All of it passed on the driver's B200 at the end of round 1 (GPUTEST_r01.json); no test here may fail silently.
(The checkpoint / resume test lives in test_gpu_mur.py: it only adds host logic to validated kernels.)"""
import contextlib
import io

import numpy as np
import pytest

from conftest import golden_v, load_golden, rel_fro

pytestmark = pytest.mark.gpu

TOL = 1e-4


def test_admm_l2n_matches_reference_golden():
    """Both factors with the l2n prox, through NMF.factorize, against the frozen reference run."""
    from nmf_b200 import NMF
    g = load_golden("admm_l2n")
    r = g["recipe"]
    v = golden_v(r["v"])
    np.random.seed(r["init_seed"])
    model = NMF(v.copy(), r["k"])
    with contextlib.redirect_stdout(io.StringIO()):
        model.factorize(method="admm", **r["kw"])
    res = model.results
    assert res.i == int(g["i"]) and len(res.obj_history) == len(g["obj_history"])
    obj_err = np.max(np.abs(np.array(res.obj_history) - g["obj_history"]) / g["obj_history"])
    assert obj_err < TOL
    assert rel_fro(res.w, g["w"]) < TOL and rel_fro(res.h, g["h"]) < TOL
    assert res.experiment.prox_h == "l2n" and res.experiment.lambda_h == 0.5


@pytest.mark.parametrize("method", ["admm", "ao_admm"])
def test_l2n_trajectory_vs_oracle_k64(method):
    """k = 64 (KP = 64 kernel variant) and, for AO-ADMM, the data-dependent rho = tr(G)/k in the
    Tikhonov system.  (ao_admm.py:128 itself raises on NumPy >= 1.24: the oracle restates the intended
    arithmetic there, parity unpinned for that sub-case.)"""
    from oracle import nmf_oracle as orc
    from nmf_b200.admm import admm
    from nmf_b200.ao_admm import ao_admm
    m, n, k, iters = 768, 640, 64, 30
    v = orc.planted(m, n, k, 0).astype(np.float64)
    w0, h0 = orc.seeded_init(m, n, k, 1)
    if method == "admm":
        kw = dict(rho=100, reg_w=(0, "nn"), reg_h=(0.4, "l2n"))
        tr = orc.admm_run(v, k, w0, h0, min_iter=10**9, max_iter=iters, **kw)
        res = admm(v, k, min_iter=10**9, max_iter=iters, w0=w0, h0=h0, verbose=False,
                   nndsvd_init=(False, "zero"), **kw)
    else:
        kw = dict(reg_w=(0.2, "l2n"), reg_h=(0.4, "l2n"), admm_iter=5)
        tr = orc.ao_admm_run(v, k, w0, h0, min_iter=10**9, max_iter=iters, **kw)
        res = ao_admm(v, k, min_iter=10**9, max_iter=iters, w0=w0, h0=h0, verbose=False,
                      nndsvd_init=(False, "zero"), **kw)
    assert res.i == tr.i
    obj_err = max(abs(a - b) / b for a, b in zip(res.obj_history, tr.obj_history))
    assert obj_err < TOL and rel_fro(res.w, tr.w) < TOL and rel_fro(res.h, tr.h) < TOL


# ----------------------------------------------------------------------------------------------
# BASELINE.json's full sizes through size-independent properties (the fp64 oracle needs minutes per
# iteration there): Lee-Seung monotonicity of the MUR objective, non-negativity, finiteness, and the
# agreement of the fused / stand-alone KL objective.  Data is generated on the device (bench.py).
# ----------------------------------------------------------------------------------------------
def _full_size_engine(m, n, k):
    import bench
    from nmf_b200.engine import Engine
    eng = Engine(m, n, k)
    bench.synth_fill_engine(eng, seed=0)
    bench.synth_init(eng, seed=1)
    return eng


def test_c2_mur_eu_full_size_is_monotone_and_non_negative():
    """configs[1]: 32768 x 32768, k = 64."""
    import torch
    from nmf_b200 import solvers
    eng = _full_size_engine(32768, 32768, 64)
    st = solvers.make_stepper(eng, "mur", "eu")
    st.prepare()
    objs = [eng.read_scalar(0)]
    for i in range(12):
        st.step(i % 8)
        objs.append(eng.read_scalar((i % 8) * 8))
    assert all(np.isfinite(objs)) and all(b < a for a, b in zip(objs, objs[1:])), objs       # Lee & Seung
    assert objs[-1] < 0.05 * objs[0]
    for f in (eng.w, eng.h):
        assert bool(torch.isfinite(f.master).all()) and float(f.master.min()) >= 0.0
    eng.release()


def test_c3_mur_kl_full_size_is_monotone_and_objectives_agree():
    """configs[2]: 20000 x 50000, k = 128; the objective that falls out of the W pass (lagged) must equal
    the stand-alone objective kernel's value for the same pair."""
    from nmf_b200 import solvers
    eng = _full_size_engine(20000, 50000, 128)
    st = solvers.make_stepper(eng, "mur", "kl")
    st.prepare()
    objs = [eng.read_scalar(0)]
    fused = []
    for i in range(8):
        st.step(i % 8, (i - 1) % 8)                       # writes the objective of the incoming pair
        fused.append(eng.read_scalar(((i - 1) % 8) * 8))
        eng.objective_kl(slot=40)                         # stand-alone, pair after iteration i
        objs.append(eng.read_scalar(40))
    assert all(np.isfinite(objs)) and all(b < a for a, b in zip(objs, objs[1:])), objs
    for got, want in zip(fused, objs[:-1]):               # fused[i] = objective before iteration i = objs[i]
        assert abs(got - want) / want < 1e-5
    eng.release()


# ----------------------------------------------------------------------------------------------
# Edge cases: exact zeros
# in V with an all-zero row and column (the 0*log(0/x) and x/0 rules of utils.py:23-26), a single
# component, the 'mean' / 'random' NNDSVD fills (utils.py:84-91), k = 200 (the KP = 256 kernels, C5's
# width), very flat / very tall V, and the input dtypes / memory orders NumPy callers hand over.
# ----------------------------------------------------------------------------------------------
EDGE_GOLDENS = ["mur_kl_zeros", "mur_eu_zeros", "mur_eu_k1", "mur_kl_k1", "admm_nndsvd_mean",
                "ao_admm_nndsvd_random"]


@pytest.mark.parametrize("name", EDGE_GOLDENS)
def test_edge_case_matches_reference_golden(name):
    from nmf_b200 import NMF
    g = load_golden(name)
    r = g["recipe"]
    v = golden_v(r["v"])
    np.random.seed(r["init_seed"])
    model = NMF(v.copy(), r["k"])
    with contextlib.redirect_stdout(io.StringIO()):
        model.factorize(method=r["method"], **r["kw"])
    res = model.results
    assert res.i == int(g["i"]) and len(res.obj_history) == len(g["obj_history"])
    obj_err = np.max(np.abs(np.array(res.obj_history) - g["obj_history"]) / g["obj_history"])
    assert obj_err < TOL
    assert rel_fro(res.w, g["w"]) < TOL and rel_fro(res.h, g["h"]) < TOL
    assert np.all(res.w >= 0) and np.all(res.h >= 0)
    if "zeros" in name:                                  # the all-zero row / column of V empties W's row / H's column
        assert np.all(res.w[3] == 0) and np.all(res.h[:, 5] == 0)


@pytest.mark.parametrize("shape", [(640, 520, 200), (4096, 70, 8), (70, 4096, 8), (5, 7, 2)])
def test_mur_eu_shapes_vs_oracle(shape):
    """k = 200 runs the KP = 256 variants of every kernel; the flat / tall cases have a single 128-row
    panel on one side and many on the other; 5 x 7 is smaller than any tile."""
    from oracle import nmf_oracle as orc
    from nmf_b200.mur import mur
    m, n, k = shape
    v = orc.planted(m, n, k, 40).astype(np.float64)
    w0, h0 = orc.seeded_init(m, n, k, 41)
    tr = orc.mur_run(v, k, w0, h0, distance_type="eu", min_iter=10**9, max_iter=30)
    res = mur(v, k, distance_type="eu", min_iter=10**9, max_iter=30, w0=w0, h0=h0, verbose=False)
    assert res.i == tr.i == 29
    obj_err = max(abs(a - b) / b for a, b in zip(res.obj_history, tr.obj_history))
    assert obj_err < TOL and rel_fro(res.w, tr.w) < TOL and rel_fro(res.h, tr.h) < TOL


def test_input_dtype_and_memory_order_do_not_matter():
    """fp32, Fortran-ordered and integer-valued V give the factors of the same V in C-ordered fp64."""
    from oracle import nmf_oracle as orc
    from nmf_b200.mur import mur
    m, n, k = 300, 260, 6
    v = np.round(orc.planted(m, n, k, 42).astype(np.float64) * 8)         # small integers, exact in every dtype
    w0, h0 = orc.seeded_init(m, n, k, 43)
    kw = dict(distance_type="eu", min_iter=10**9, max_iter=10, w0=w0, h0=h0, verbose=False)
    base = mur(v.copy(), k, **kw)
    for other in (v.astype(np.float32), np.asfortranarray(v), v.astype(np.int64), v[::1, ::1].T.copy().T):
        res = mur(other, k, **kw)
        assert np.array_equal(res.w, base.w) and np.array_equal(res.h, base.h)
        assert res.obj_history == base.obj_history


# ----------------------------------------------------------------------------------------------
# Kullback-Leibler branch of admm (admm.py:303-315) and ao_admm (ao_admm.py:71-101,
# 272-285), single GPU.
# ----------------------------------------------------------------------------------------------
@pytest.mark.parametrize("name", ["admm_kl", "ao_admm_kl"])
def test_kl_branch_matches_reference_golden(name):
    from nmf_b200 import NMF
    from nmf_b200.ao_admm import ao_admm
    g = load_golden(name)
    r = g["recipe"]
    v = golden_v(r["v"])
    np.random.seed(r["init_seed"])
    model = NMF(v.copy(), r["k"])
    sink = io.StringIO()
    with contextlib.redirect_stdout(sink):
        model.factorize(method=r["method"], **r["kw"])
    res = model.results
    assert res.experiment.distance_type == "kl"
    assert res.i == int(g["i"]) and len(res.obj_history) == len(g["obj_history"])
    obj_err = np.max(np.abs(np.array(res.obj_history) - g["obj_history"]) / g["obj_history"])
    assert obj_err < TOL
    assert rel_fro(res.w, g["w"]) < TOL and rel_fro(res.h, g["h"]) < TOL
    if r["method"] == "ao_admm":                          # same early breaks of the inner loops
        assert sink.getvalue().count("ADMM break after") == int(g["n_inner_breaks"])
        assert len(ao_admm.last_inner_iterations) == res.i + 1


@pytest.mark.parametrize("method", ["admm", "ao_admm"])
def test_kl_branch_trajectory_vs_oracle_k64(method):
    from oracle import nmf_oracle as orc
    from nmf_b200.admm import admm
    from nmf_b200.ao_admm import ao_admm
    m, n, k, iters = 768, 640, 64, 30
    v = orc.planted(m, n, k, 0).astype(np.float64)
    w0, h0 = orc.seeded_init(m, n, k, 1)
    if method == "admm":
        kw = dict(rho=100, distance_type="kl", reg_w=(0, "nn"), reg_h=(0.05, "l1n"))
        tr = orc.admm_run(v, k, w0, h0, min_iter=10**9, max_iter=iters, **kw)
        res = admm(v, k, min_iter=10**9, max_iter=iters, w0=w0, h0=h0, verbose=False,
                   nndsvd_init=(False, "zero"), **kw)
    else:
        kw = dict(distance_type="kl", reg_w=(0, "nn"), reg_h=(0.05, "l1n"), admm_iter=4)
        tr = orc.ao_admm_run(v, k, w0, h0, min_iter=10**9, max_iter=iters, **kw)
        res = ao_admm(v, k, min_iter=10**9, max_iter=iters, w0=w0, h0=h0, verbose=False,
                      nndsvd_init=(False, "zero"), **kw)
        assert [tuple(u) for u in ao_admm.last_inner_iterations] == [tuple(u) for u in tr.inner_breaks]
    assert res.i == tr.i
    obj_err = max(abs(a - b) / b for a, b in zip(res.obj_history, tr.obj_history))
    assert obj_err < TOL and rel_fro(res.w, tr.w) < TOL and rel_fro(res.h, tr.h) < TOL


# ----------------------------------------------------------------------------------------------
# The two tcgen05 objective kernels against the CUDA-core kernel (impl 1) and the fp64 value:
#   impl 2 = residual_n256.cu (128 x 256 tiles; the product's choice for KP <= 128 with one V plane),
#   impl 3 = residual_ts.cu (W planes as the TMEM-resident A operand; every other shape), impl 0 = the product's choice.
# ----------------------------------------------------------------------------------------------
def _objective_variant_agrees(impl, m, n, k, exact_data=True):
    from oracle import nmf_oracle as orc
    from nmf_b200.engine import Engine
    if exact_data:
        v = orc.planted(m, n, k, 50)                         # on the bf16 grid: one V plane
    else:
        v = np.random.default_rng(52).random((m, n)).astype(np.float32)      # needs the hi + lo planes
    w0, h0 = orc.seeded_init(m, n, k, 51)
    eng = Engine(m, n, k)
    eng.upload_data(np.asarray(v))
    assert eng.nv == (1 if exact_data else 2)
    eng.set_factors(w0, h0)
    eng.objective_eu(slot=0, impl=1)
    eng.objective_eu(slot=1, impl=impl)
    eng.objective_eu(slot=2, impl=impl)                      # a second launch: barriers / TMEM left clean
    eng.objective_eu(slot=3, impl=0)
    a, b, c, d = (eng.read_scalar(i) for i in range(4))
    want = orc.objective(np.asarray(v, np.float64), w0.astype(np.float32).astype(np.float64)
                         @ h0.astype(np.float32).astype(np.float64))
    eng.release()
    assert b == c
    assert abs(b - a) / a < 1e-5 and abs(b - want) / want < 1e-5 and abs(d - want) / want < 1e-5


@pytest.mark.parametrize("shape", [(1024, 1536, 128), (900, 1100, 48), (2304, 700, 100), (128, 256, 64),
                                   (4096, 8192, 128)])
def test_wide_tile_objective_agrees_with_the_cuda_core_kernel(shape):
    """Full and ragged tiles on both axes, a single tile, KP = 64 and 128, enough tiles for every CTA to walk
    several rows of tiles."""
    _objective_variant_agrees(2, *shape)


@pytest.mark.parametrize("shape,exact", [((1024, 1536, 128), True), ((900, 1100, 48), True), ((640, 520, 200), True),
                                         ((128, 128, 64), True), ((4096, 8192, 128), True), ((1100, 900, 100), False),
                                         ((640, 520, 200), False)])
def test_tmem_operand_objective_agrees_with_the_cuda_core_kernel(shape, exact):
    """KP = 64 / 128 / 256 (k = 200: the planes fill all 512 TMEM columns next to the accumulators), one tile,
    several rows of tiles per CTA, two V planes (also at KP = 256)."""
    _objective_variant_agrees(3, *shape, exact_data=exact)


@pytest.mark.parametrize("shape", [(768, 640, 100), (512, 2048, 24), (2000, 5000, 128), (2000, 5000, 64)])
def test_kl_pass_with_tmem_operands_matches_the_oracle(shape):
    """kl_ts2.cu (KP = 128: X and the quotient as TMEM-resident A operands) and kl_ts.cu (KP = 64: X only) against
    the oracle, fused objective included, several work items per CTA at the last shapes."""
    from oracle import nmf_oracle as orc
    from nmf_b200.mur import mur
    m, n, k = shape
    v = orc.planted(m, n, k, 60).astype(np.float64)
    w0, h0 = orc.seeded_init(m, n, k, 61)
    res = mur(v, k, distance_type="kl", min_iter=10**9, max_iter=6, w0=w0, h0=h0, verbose=False)
    tr = orc.mur_run(v, k, w0, h0, distance_type="kl", min_iter=10**9, max_iter=6)
    assert rel_fro(res.w, tr.w) < TOL and rel_fro(res.h, tr.h) < TOL
    assert max(abs(a - b) / abs(b) for a, b in zip(res.obj_history, tr.obj_history)) < TOL
