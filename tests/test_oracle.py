"""CPU: the oracle (oracle/nmf_oracle.py) must reproduce the frozen outputs of the
unmodified reference (tests/golden/*.npz, made by tests/golden/make_golden.py)."""
import numpy as np
import pytest

from conftest import GOLDEN_NAMES, MID_GOLDEN_NAMES, golden_v, load_golden, rel_fro
from oracle import nmf_oracle as orc


def replay(g):
    r = g["recipe"]
    v = golden_v(r["v"])
    kw = dict(r["kw"])
    nnd = kw.pop("nndsvd_init", (False, "zero"))
    m, n = v.shape
    if nnd[0]:
        np.random.seed(r["init_seed"])               # the 'random' fill draws from the global RNG (utils.py:88-91)
        w0, h0 = orc.nndsvd(v, r["k"], nnd[1])
    else:
        w0, h0 = orc.seeded_init(m, n, r["k"], r["init_seed"])
    fn = {"mur": orc.mur_run, "admm": orc.admm_run, "ao_admm": orc.ao_admm_run}[r["method"]]
    return fn(v, r["k"], w0, h0, **kw)


@pytest.mark.parametrize("name", GOLDEN_NAMES)
def test_oracle_matches_reference_golden(name):
    g = load_golden(name)
    t = replay(g)
    assert t.i == int(g["i"])
    assert len(t.obj_history) == len(g["obj_history"]) == t.i + 2
    np.testing.assert_allclose(t.obj_history, g["obj_history"], rtol=1e-9, atol=1e-9)
    assert rel_fro(t.w, g["w"]) < 1e-9
    assert rel_fro(t.h, g["h"]) < 1e-9
    if g["recipe"]["method"] == "ao_admm":
        used = sum((a < g["recipe"]["kw"]["admm_iter"]) + (b < g["recipe"]["kw"]["admm_iter"])
                   for a, b in t.inner_breaks)
        assert used == int(g["n_inner_breaks"])


@pytest.mark.parametrize("name", MID_GOLDEN_NAMES)
def test_oracle_first_iterations_match_mid_size_golden(name):
    """SURVEY 8d sizes (8192^2 k=64, 4000x10000 k=128, 16384x4096 k=128): the first three iterations of the oracle
    against the reference's frozen objective history (the whole 100-iteration runs are the GPU tests' business;
    three iterations keep the CPU suite short)."""
    g = load_golden(name)
    g["recipe"]["kw"] = dict(g["recipe"]["kw"], max_iter=3, min_iter=10**9)
    t = replay(g)
    np.testing.assert_allclose(t.obj_history, g["obj_history"][:4], rtol=1e-9)


def test_c1_known_answers():
    """values measured on the reference at survey time."""
    g = load_golden("c1_mur_eu")
    assert int(g["i"]) == 199 and len(g["obj_history"]) == 201
    assert abs(g["obj_history"][0] - 9542736.416689271) < 1e-6
    assert abs(g["obj_history"][-1] - 19939.291063649223) < 1e-6


def test_stop_test_semantics():
    assert orc.stop_test(0.5e-3, 1.0, 1e-3, 1e-3) == 1
    assert orc.stop_test(1.0, 1.0005, 1e-3, 1e-3) == 2
    assert orc.stop_test(1.0, 1.1, 1e-3, 1e-3) == 0


def test_kl_objective_special_values():
    v = np.array([[0.0, 1.0], [2.0, 0.0]])
    wh = np.array([[1.0, 1.0], [0.0, 3.0]])          # v/wh = inf at [1,0] -> term mapped to 0
    val = orc.objective(v, wh, "kl")
    expect = (0 - 0 + 1) + (0 - 1 + 1) + (0 - 2 + 0) + (0 - 0 + 3)
    assert abs(val - expect) < 1e-12


def test_bf16_grid_is_idempotent_and_exact():
    v = orc.planted(64, 48, 4, 1)
    assert v.dtype == np.float32
    assert np.array_equal(orc.round_to_bf16_grid(v), v)
    assert np.all((v.view(np.uint32) & 0xFFFF) == 0)
