"""CPU: the C-ABI library builds / loads and exports every symbol declared in include/nmf_b200.h
(no compute calls -- there is no GPU here), and the host-side mirror of the reference API behaves
like the reference (names, defaults, file naming, stop test, error types)."""
import ctypes as C
import inspect
import os
import re
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def _header_symbols():
    text = open(os.path.join(ROOT, "include", "nmf_b200.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(nmfb_[a-z0-9_]+)\s*\(", text)))


def test_library_exports_every_declared_symbol():
    from nmf_b200 import _build, _lib
    _build.build()
    lib = C.CDLL(_build.LIB_PATH)
    declared = _header_symbols()
    assert len(declared) >= 20
    missing = [s for s in declared if not hasattr(lib, s)]
    assert not missing, f"declared in the header but not exported: {missing}"
    bound = set(_lib.exported_symbols())
    assert set(declared) == bound, f"ctypes binding and header disagree: {set(declared) ^ bound}"


def test_pure_host_entry_points():
    from nmf_b200 import _lib
    lib = _lib.load()
    assert lib.nmfb_abi_version() >= 1
    assert [lib.nmfb_padded_components(k) for k in (1, 64, 65, 128, 129, 256)] == [64, 64, 128, 128, 256, 256]
    assert lib.nmfb_padded_components(257) < 0 and b"256" in lib.nmfb_last_error()
    assert lib.nmfb_padded_length(1000) == 1024 and lib.nmfb_padded_length(1024) == 1024


def test_struct_layouts_match_header():
    """sizeof and every field's offsetof of the ctypes mirrors == what a C compiler lays out for the
    header (guards against drift: same names, same order, same padding)."""
    import subprocess, tempfile
    from nmf_b200 import _lib
    pairs = [("nmfb_data", _lib.Data), ("nmfb_factor", _lib.Factor), ("nmfb_problem", _lib.Problem),
             ("nmfb_admm_state", _lib.AdmmState), ("nmfb_peer_ctx", _lib.PeerCtx)]
    lines, want = [], []
    for cname, cls in pairs:
        lines.append(f'printf("%zu\\n", sizeof({cname}));')
        want.append(C.sizeof(cls))
        for fname, _ in cls._fields_:
            lines.append(f'printf("%zu\\n", offsetof({cname}, {fname}));')      # fails to compile on a renamed field
            want.append(getattr(cls, fname).offset)
    src = '#include <stdio.h>\n#include <stddef.h>\n#include "nmf_b200.h"\nint main(){' + "".join(lines) + 'return 0;}\n'
    with tempfile.TemporaryDirectory() as d:
        c = os.path.join(d, "t.c")
        open(c, "w").write(src)
        exe = os.path.join(d, "t")
        subprocess.check_call(["gcc", "-I", os.path.join(ROOT, "include"), c, "-o", exe])
        got = [int(x) for x in subprocess.check_output([exe]).split()]
    assert got == want


def test_ctypes_signatures_match_header_prototypes():
    """Every prototype of the header, argument by argument, against the ctypes argtypes / restype."""
    from nmf_b200 import _lib
    text = open(os.path.join(ROOT, "include", "nmf_b200.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    text = re.sub(r"//[^\n]*", "", text)
    protos = re.findall(r"([A-Za-z_][\w \*]*?)\b(nmfb_[a-z0-9_]+)\s*\(([^)]*)\)\s*;", text)
    assert len(protos) == len(_lib.exported_symbols())
    structs = {"nmfb_data": _lib.Data, "nmfb_factor": _lib.Factor, "nmfb_problem": _lib.Problem,
               "nmfb_admm_state": _lib.AdmmState, "nmfb_peer_ctx": _lib.PeerCtx}
    scalars = {"int": C.c_int, "long": C.c_long, "float": C.c_float, "double": C.c_double, "size_t": C.c_size_t,
               "unsigned long long": C.c_ulonglong, "unsigned int": C.c_uint}

    def to_ctypes(ctype):
        ctype = ctype.replace("const", "").strip()
        if ctype.endswith("*"):
            base = ctype[:-1].strip()
            if base == "char":
                return C.c_char_p
            return C.POINTER(structs[base]) if base in structs else C.c_void_p
        return scalars[ctype]

    for ret, name, args in protos:
        res, argtypes = _lib._SIGNATURES[name]
        want = []
        for a in args.split(","):
            a = a.strip()
            if a and a != "void":
                want.append(to_ctypes(re.match(r"(.*?)(\w+)$", a).group(1)))
        assert want == list(argtypes), name
        assert to_ctypes(ret) == res, name


def test_no_cpu_fallback():
    """The product path must fail loudly without a CUDA device."""
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    from nmf_b200 import NMF
    with pytest.raises(RuntimeError, match="CUDA"):
        NMF(np.ones((8, 8)), 2).factorize(method="mur", distance_type="eu", max_iter=1)


def test_product_never_imports_the_oracle():
    pkg = os.path.join(ROOT, "nmf_b200")
    for fn in os.listdir(pkg):
        if fn.endswith(".py"):
            assert "oracle" not in open(os.path.join(pkg, fn)).read(), fn


def test_solver_signatures_mirror_the_reference():
    """Keyword names and defaults of mur.py:52-53, admm.py:233-235, ao_admm.py:201-203."""
    from nmf_b200.admm import admm
    from nmf_b200.ao_admm import ao_admm
    from nmf_b200.mur import mur
    want = {
        mur: dict(distance_type='kl', min_iter=100, max_iter=100000, tol1=1e-5, tol2=1e-5, lambda_w=0.0,
                  lambda_h=0.0, nndsvd_init=(False, 'zero'), save_dir='./results/'),
        admm: dict(rho=1, distance_type='eu', reg_w=(0, 'nn'), reg_h=(0, 'l2n'), min_iter=10, max_iter=100000,
                   tol1=1e-3, tol2=1e-3, nndsvd_init=(True, 'zero'), save_dir='./results/'),
        ao_admm: dict(distance_type='eu', reg_w=(0, 'nn'), reg_h=(0, 'l2n'), min_iter=10, max_iter=100000,
                      admm_iter=10, tol1=1e-3, tol2=1e-3, nndsvd_init=(True, 'zero'), save_dir='./results/'),
    }
    for fn, kw in want.items():
        params = inspect.signature(fn).parameters
        for name, default in kw.items():
            assert params[name].kind is inspect.Parameter.KEYWORD_ONLY
            assert params[name].default == default, (fn.__name__, name)


def test_stop_test_and_precision():
    from nmf_b200.utils import convergence_check, tol_precision
    assert convergence_check(0.5e-3, 1.0, 1e-3, 1e-3) is True          # (1) new < tol1
    assert convergence_check(1.0, 1.0005, 1e-3, 1e-3) is True          # (2) new >= old - tol2
    assert convergence_check(1.0, 1.1, 1e-3, 1e-3) is False
    assert tol_precision(1e-5, 1e-3) == 5 and tol_precision(2.0, 3.0) == 2


def test_prox_names():
    from nmf_b200.solvers import prox_code
    assert prox_code((0, 'nn')) == (0.0, 0) and prox_code((0.1, 'l1n')) == (0.1, 1)
    assert prox_code((0, 'l2n')) == (0.0, 0)                           # admm.py:150: identity system
    with pytest.raises(TypeError):
        prox_code((0.1, 'l1'))                                         # not a reference name (admm.py:213)
    assert prox_code((0.5, 'l2n')) == (0.5, 2)                         # admm.py:141-156
    with pytest.raises(NotImplementedError):
        prox_code((0.5, 'l2n'), k=200)
    with pytest.raises(NotImplementedError):
        prox_code((0.5, 'l1inf'))


def test_facade_dispatch_and_save_names(tmp_path, capsys):
    from collections import namedtuple
    from nmf_b200 import NMF
    model = NMF(np.ones((4, 3)), 2)
    with pytest.raises(Exception, match="Method not known"):
        model.factorize(method="bogus")                                # nmf.py:76
    with pytest.raises(NotImplementedError):
        model.factorize(method="anls")
    assert model.w is None and model.h is None
    # file naming of nmf.py:95-126 for the three solvers
    R = namedtuple('Results', 'w h i obj_history experiment')
    E1 = namedtuple('Experiment', 'method components distance_type nndsvd_init max_iter tol1 tol2 lambda_w lambda_h')
    model.results = R(np.ones((4, 2)), np.ones((2, 3)), 3, [1.0, 0.5, 0.4, 0.3, 0.2],
                      E1('mur', 2, 'eu', (False, 'zero'), 10, 1e-5, 1e-5, 0.0, 0.1))
    assert model.default_save_name() == 'nmf_mur_2_eu_0.0_0.1_random'
    E2 = namedtuple('Experiment', 'method components rho distance_type nndsvd_init min_iter max_iter tol1 tol2 lambda_w prox_w lambda_h prox_h')
    model.results = model.results._replace(experiment=E2('admm', 2, 100, 'eu', (True, 'mean'), 1, 2, 1e-3, 1e-3, 0, 'nn', 0.1, 'l1n'))
    assert model.default_save_name() == 'nmf_admm_2_eu_100_0:nn_0.1:l1n_nndsvdm'
    E3 = namedtuple('Experiment', 'method components distance_type nndsvd_init min_iter max_iter admm_iter tol1 tol2 lambda_w prox_w lambda_h prox_h')
    model.results = model.results._replace(experiment=E3('ao_admm', 2, 'eu', (False, 'zero'), 1, 2, 5, 1e-3, 1e-3, 0, 'nn', 0.1, 'l1n'))
    assert model.default_save_name() == 'nmf_ao_admm_2_eu_0:nn_0.1:l1n_random'
    model.save_factorization(save_dir=str(tmp_path))
    z = np.load(tmp_path / 'nmf_ao_admm_2_eu_0:nn_0.1:l1n_random.npz', allow_pickle=True)
    assert sorted(z.files) == ['experiment', 'h', 'i', 'obj_history', 'w']      # utils.py:103
    assert z['experiment'].item()['admm_iter'] == 5 and int(z['i']) == 3
    assert 'Results saved in' in capsys.readouterr().out


def test_reference_order_initial_factors():
    from nmf_b200.init import initial_factors
    np.random.seed(7)
    w, h = initial_factors(np.ones((5, 4)), 3, (False, 'zero'))
    np.random.seed(7)
    w_ref = np.abs(np.random.randn(5, 3)); h_ref = np.abs(np.random.randn(3, 4))   # mur.py:108-109
    assert np.array_equal(w, w_ref) and np.array_equal(h, h_ref)


def test_nndsvd_from_the_gram_eigenvectors_matches_the_reference_svd():
    """init.nndsvd takes the k leading triplets from the eigenvectors of the small Gram matrix (utils.py:44 takes
    them from the full SVD): same W0 / H0 to ~1e-13 for tall, wide, planted and uniform data and every fill
    variant, and through the full-SVD branch used above GRAM_LIMIT."""
    from nmf_b200 import init
    from oracle import nmf_oracle as orc
    rng = np.random.default_rng(0)
    rel = lambda a, b: np.linalg.norm(a - b) / np.linalg.norm(b)   # noqa: E731
    cases = [(orc.planted(200, 160, 6, 15).astype(np.float64), 6), (rng.random((300, 120)), 10),
             (rng.random((90, 400)), 8), (orc.planted(120, 160, 5, 37).astype(np.float64), 5)]
    for x, k in cases:
        for variant in ("zero", "mean", "random"):
            np.random.seed(3)
            w, h = init.nndsvd(x, k, variant, _device="cpu")
            np.random.seed(3)
            w_ref, h_ref = orc.nndsvd(x, k, variant)
            assert rel(w, w_ref) < 1e-11 and rel(h, h_ref) < 1e-11
            assert w.shape == (x.shape[0], k) and h.shape == (k, x.shape[1]) and w.min() >= 0 and h.min() >= 0
    x, k = cases[1]
    old, init.GRAM_LIMIT = init.GRAM_LIMIT, 64              # force the full-SVD branch
    try:
        w, h = init.nndsvd(x, k, "zero", _device="cpu")
    finally:
        init.GRAM_LIMIT = old
    w_ref, h_ref = orc.nndsvd(x, k, "zero")
    assert rel(w, w_ref) < 1e-11 and rel(h, h_ref) < 1e-11


SHAPES_FOR_PLANS = [(5, 7, 2), (37, 23, 1), (23, 37, 1), (4096, 70, 8), (70, 4096, 8), (640, 520, 200), (1000, 500, 10),
                    (32768, 32768, 64), (20000, 50000, 128), (65536, 16384, 128), (65536, 65536, 128),
                    (262144, 131072, 256)]


@pytest.mark.parametrize("shape", SHAPES_FOR_PLANS)
def test_argument_and_plan_stage_accepts_the_shape(shape):
    """Without a driver every solver entry point must get as far as encoding its first TMA descriptor: tile plans,
    workspace sizes and padded-extent checks (all host code of the library) accept the edge-case shapes of the GPU
    tests and BASELINE.json's configurations.  (Dummy pointers: never run where a GPU could dereference them.)"""
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present: the calls would launch on dummy pointers")
    from nmf_b200 import _lib
    from nmf_b200._lib import Data, Factor, Problem
    lib = _lib.load()
    m, n, k = shape
    KP = lib.nmfb_padded_components(k)
    Mp, Np = lib.nmfb_padded_length(m), lib.nmfb_padded_length(n)
    buf = np.zeros(1 << 16, np.uint8)
    base = buf.ctypes.data
    pbytes = lib.nmfb_partial_bytes(Mp, Np, k, 148)
    assert pbytes > 0

    def problem(nv, fixed):
        vr = Data(base, m, n, Mp, Np, nv, 1, fixed, 0)
        vc = Data(base, n, m, Np, Mp, nv, 1, fixed, 0)
        w, h = Factor(base, base, base, m, Mp, 0, 0), Factor(base, base, base, n, Np, 0, 0)
        return Problem(m, n, k, KP, vr, vc, w, h, base, base, base, pbytes, base,
                       lib.nmfb_obj_partial_len(m, n, 148), 148, None, 0, base)

    def reached_the_driver(rc, what):
        msg = lib.nmfb_last_error().decode()
        assert rc == -2 and "cuTensorMapEncodeTiled" in msg, f"{what} {shape}: rc={rc} {msg}"

    for nv in (1, 2):
        p = problem(nv, 0)
        for side in (0, 1):
            reached_the_driver(lib.nmfb_mur_eu_half_step(C.byref(p), side, 0.0, None), f"mur_eu side {side} nv={nv}")
            rc = lib.nmfb_mur_kl_half_step(C.byref(p), side, 0.0, None, None)
            if KP <= 128:
                reached_the_driver(rc, f"mur_kl side {side} nv={nv}")
            else:
                assert rc == -1 and b"k <= 128" in lib.nmfb_last_error()
        # every (KP, planes) pair has a tcgen05 objective kernel: wide tiles for KP <= 128 with one plane, the TS
        # form (W planes in tensor memory, so that two V planes fit next to KP = 256) for the rest
        reached_the_driver(lib.nmfb_objective_eu(C.byref(p), 0, base, None), "objective_eu")
        reached_the_driver(lib.nmfb_objective_kl(C.byref(p), base, None), "objective_kl")
    # the ADMM / AO-ADMM contractions, floating (3 factor planes) and exact-hi, on both copies of V
    w, h = Factor(base, base, base, m, Mp, 0, 0), Factor(base, base, base, n, Np, 0, 0)
    for fixed in (0, 1):
        p = problem(2, fixed)
        for data, other, self_ in ((p.v_cols, w, h), (p.v_rows, h, w)):
            if fixed:
                reached_the_driver(lib.nmfb_gram_exact(C.byref(other), KP, base, pbytes, base, base, 148, None),
                                   "gram_exact")
                reached_the_driver(lib.nmfb_panel_gemm_exact(C.byref(data), base, other.Lp, KP, base, pbytes,
                                                             self_.Lp, 148, None), "panel_gemm_exact")
            else:
                reached_the_driver(lib.nmfb_gram(C.byref(other), KP, 3, base, pbytes, base, base, 148, None), "gram")
                reached_the_driver(lib.nmfb_panel_gemm(C.byref(data), base, other.Lp, KP, 3, base, pbytes, self_.Lp,
                                                       148, 0, 0, None), "panel_gemm nb=3")


def test_reference_package_name_resolves_to_the_b200_build():
    """`from nmf import NMF` (reference nmf/__init__.py:1, README.md:18-22) is the drop-in facade."""
    import nmf
    import nmf_b200
    from nmf import NMF
    from nmf.mur import mur
    assert NMF is nmf_b200.NMF and mur is nmf_b200.mur.mur and nmf.admm.admm is nmf_b200.admm.admm


def test_round2_entry_points_reject_bad_arguments_before_touching_the_gpu():
    """Argument stage of the entry points added in round 2 (exchange, fp64 Gram, no-allocation Gram): clear status
    codes and messages, no CUDA call needed."""
    from nmf_b200 import _lib
    lib = _lib.load()
    err = lambda: lib.nmfb_last_error().decode()
    fac = _lib.Factor(0x1000, 0x1000, 0x1000, 1000, 1024, 0, 0)                       # no window
    ctx = _lib.PeerCtx()                                                              # all NULL / zero
    assert lib.nmfb_exchange_factor(None, 128, C.byref(fac), 0, 1, 0, 0, 0, None, None, 1, None, None) == -1
    assert lib.nmfb_exchange_factor(C.byref(ctx), 128, C.byref(fac), 0, 1, 0, 0, 0, None, None, 1, None, None) == -1
    peers = (C.c_void_p * 2)(0x1000, 0x2000)
    ctx = _lib.PeerCtx(C.cast(peers, C.c_void_p), None, 5, 2, None, None, None, None, 0, None, None, 148)
    assert lib.nmfb_exchange_factor(C.byref(ctx), 128, C.byref(fac), 0, 1, 0, 0, 0, None, None, 1, None, None) == -1
    assert "rank 5 of 2" in err()
    ctx.rank = 1
    assert lib.nmfb_exchange_factor(C.byref(ctx), 128, C.byref(fac), 0, 1, 0, 0, 0, None, None, 1, None, None) == -1
    assert "no window" in err()
    fac.win0, fac.win_p = 512, 512
    assert lib.nmfb_exchange_factor(C.byref(ctx), 128, C.byref(fac), 0, 1, 0, 2, 0, 0x1000, 0x1000, 1, None, None) == -1
    assert "Gram workspaces" in err()
    assert lib.nmfb_exchange_scalar(None, 0x1000, 0, 1, 0x1000, None) == -1
    assert lib.nmfb_gram_fp64(C.byref(fac), 100, 0x1000, 1 << 30, 0x1000, 0x1000, None) == -1 and "64 / 128 / 256" in err()
    assert lib.nmfb_gram_fp64(C.byref(fac), 128, 0x1000, 16, 0x1000, 0x1000, None) == -3 and "scratch" in err()
    big = _lib.Factor(0x1000, 0x1000, 0x1000, 262144, 262144, 0, 0)
    assert lib.nmfb_gram_static(C.byref(big), 256, 2, 0x1000, 0x1000, 148, None) == -1 and "built-in scratch" in err()
    assert lib.nmfb_gram_static(C.byref(big), 128, 5, 0x1000, 0x1000, 148, None) == -1 and "nplanes" in err()
    assert int(lib.nmfb_gram_static_max_bytes()) == 8 << 20
