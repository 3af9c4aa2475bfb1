"""CPU: the whole host side of the product -- drivers, host loop, steppers, Engine, upload passes -- on the host
stand-in of the device (tests/softdev.py): every frozen reference run is replayed through the PUBLIC API
(`NMF(v, k).factorize(...)`), with V really living in K-blocked bf16 planes written by the (emulated) ingest
kernels, the solver arithmetic restated in NumPy behind the C-ABI entry points, and the freshness of every bf16
factor plane asserted at each consumer.  What this pins down is the orchestration: which buffer goes where, in
which order, with which planes re-derived -- including the paths whose first GPU run is still pending (prox l2n,
the Kullback-Leibler branch of admm / ao_admm, checkpoints)."""
import contextlib
import io
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

from conftest import GOLDEN_NAMES, golden_v, load_golden, rel_fro  # noqa: E402
from oracle import nmf_oracle as orc  # noqa: E402

if not os.path.exists("/usr/local/cuda/include/cuda_bf16.h"):
    pytest.skip("CUDA headers not found (host emulation of the ingest kernels)", allow_module_level=True)

import softdev  # noqa: E402

TOL = 1e-4


@pytest.fixture
def soft(monkeypatch):
    import nmf_b200.init as init_mod
    real = init_mod.nndsvd                                       # the product's NNDSVD arithmetic, torch on the CPU
    monkeypatch.setattr(init_mod, "nndsvd", lambda x, k, variant='zero': real(x, k, variant, _device='cpu'))
    with softdev.soft_device():
        yield


@pytest.mark.parametrize("name", GOLDEN_NAMES)
def test_public_api_on_the_soft_device_matches_reference_golden(soft, name):
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
    assert res.i == int(g["i"]), "stop iteration differs from the reference"
    assert len(res.obj_history) == len(g["obj_history"])
    obj_err = np.max(np.abs(np.array(res.obj_history) - g["obj_history"]) / g["obj_history"])
    assert obj_err < TOL
    assert rel_fro(res.w, g["w"]) < TOL and rel_fro(res.h, g["h"]) < TOL
    assert res.w.dtype == np.float64 and res.w.shape == g["w"].shape
    if r["method"] == "ao_admm":
        assert sink.getvalue().count("ADMM break after") == int(g["n_inner_breaks"])
        assert len(ao_admm.last_inner_iterations) == res.i + 1
    assert "Factorization done." in sink.getvalue()


# ----------------------------------------------------------------------------------------------
# N > 1: the same drivers with comm=True on two gloo ranks, each with its own soft device holding only its rows
# and columns of V (the sharded ingest, the window half-steps, the all-gathers of masters / norm partials, the
# plane refreshes after every exchange, the all-reduced objective) against the unsharded oracle.
# ----------------------------------------------------------------------------------------------
def _sharded_worker(rank, world, port, ret):
    import torch.distributed as dist
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from nmf_b200.admm import admm
        from nmf_b200.ao_admm import ao_admm
        from nmf_b200.mur import mur
        m, n, k, iters = 600, 520, 6, 12                     # 2 ranks: windows of 512 rows / columns, the second one ragged
        v = orc.planted(m, n, k, 0).astype(np.float64)
        w0, h0 = orc.seeded_init(m, n, k, 1)
        common = dict(min_iter=10**9, max_iter=iters, w0=w0, h0=h0, verbose=False, comm=True)
        out = {}
        with softdev.soft_device():
            for name, run, ref in (
                ("mur_eu", lambda: mur(v, k, distance_type="eu", lambda_w=0.05, **common),
                 lambda: orc.mur_run(v, k, w0, h0, distance_type="eu", lambda_w=0.05, min_iter=10**9, max_iter=iters)),
                ("mur_kl", lambda: mur(v, k, distance_type="kl", **common),
                 lambda: orc.mur_run(v, k, w0, h0, distance_type="kl", min_iter=10**9, max_iter=iters)),
                ("admm", lambda: admm(v, k, rho=20, reg_w=(0, "nn"), reg_h=(0.05, "l1n"),
                                      nndsvd_init=(False, "zero"), **common),
                 lambda: orc.admm_run(v, k, w0, h0, rho=20, reg_w=(0, "nn"), reg_h=(0.05, "l1n"), min_iter=10**9,
                                      max_iter=iters)),
                ("admm_kl", lambda: admm(v, k, rho=20, distance_type="kl", reg_w=(0, "nn"), reg_h=(0.05, "l1n"),
                                         nndsvd_init=(False, "zero"), **common),
                 lambda: orc.admm_run(v, k, w0, h0, rho=20, distance_type="kl", reg_w=(0, "nn"), reg_h=(0.05, "l1n"),
                                      min_iter=10**9, max_iter=iters)),
                ("ao_admm_kl", lambda: ao_admm(v, k, distance_type="kl", reg_w=(0, "nn"), reg_h=(0.05, "l1n"),
                                               admm_iter=3, nndsvd_init=(False, "zero"), **common),
                 lambda: orc.ao_admm_run(v, k, w0, h0, distance_type="kl", reg_w=(0, "nn"), reg_h=(0.05, "l1n"),
                                         admm_iter=3, min_iter=10**9, max_iter=iters)),
                ("ao_admm", lambda: ao_admm(v, k, reg_w=(0.1, "l2n"), reg_h=(0.05, "l1n"), admm_iter=4,
                                            nndsvd_init=(False, "zero"), **common),
                 lambda: orc.ao_admm_run(v, k, w0, h0, reg_w=(0.1, "l2n"), reg_h=(0.05, "l1n"), admm_iter=4,
                                         min_iter=10**9, max_iter=iters)),
            ):
                res, tr = run(), ref()
                oe = max(abs(a - b) / b for a, b in zip(res.obj_history, tr.obj_history))
                out[name] = (res.i == tr.i, oe, rel_fro(res.w, tr.w), rel_fro(res.h, tr.h),
                             float(np.abs(res.w).sum() + np.abs(res.h).sum()))
                if name == "ao_admm":
                    out["ao_breaks"] = [tuple(u) for u in ao_admm.last_inner_iterations] == \
                        [tuple(u) for u in tr.inner_breaks]
        ret[rank] = out
    finally:
        dist.destroy_process_group()


def test_two_ranks_through_the_public_drivers_reproduce_the_unsharded_oracle():
    import torch.multiprocessing as mp
    softdev.emu_lib()                                        # build once, before the ranks race for it
    world = 2
    ctx = mp.get_context("spawn")
    ret = ctx.Manager().dict()
    port = 29900 + (os.getpid() % 90)
    procs = [ctx.Process(target=_sharded_worker, args=(r, world, port, ret)) for r in range(world)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(600)
    assert all(p.exitcode == 0 for p in procs), [p.exitcode for p in procs]
    for name in ("mur_eu", "mur_kl", "admm", "admm_kl", "ao_admm_kl", "ao_admm"):
        for rank in range(world):
            same_i, oe, ew, eh, _ = ret[rank][name]
            assert same_i and oe < TOL and ew < TOL and eh < TOL, (name, rank, ret[rank][name])
        assert ret[0][name][4] == ret[1][name][4], f"{name}: the ranks hold different factors"
    assert ret[0]["ao_breaks"] and ret[1]["ao_breaks"]


# ----------------------------------------------------------------------------------------------
# Host behaviours around the solver loop, through the public API on the soft device
# ----------------------------------------------------------------------------------------------
@pytest.mark.parametrize("kind", ["eu", "kl"])
def test_stop_iteration_and_returned_factors_follow_the_reference_loop(soft, kind):
    """A stop test that fires: same stop iteration, the factors OF that iteration (the device runs one iteration
    ahead of the host; MUR-KL additionally reads its objective one iteration late), same history."""
    from nmf_b200.mur import mur
    m, n, k = 90, 70, 4
    v = orc.planted(m, n, k, 0).astype(np.float64)
    w0, h0 = orc.seeded_init(m, n, k, 1)
    kw = dict(distance_type=kind, min_iter=3, max_iter=400, tol1=1e-3, tol2=0.05)
    want = orc.mur_run(v, k, w0, h0, **kw)
    assert 3 < want.i < 399
    with contextlib.redirect_stdout(io.StringIO()) as sink:
        res = mur(v, k, w0=w0, h0=h0, **kw)
    assert res.i == want.i and len(res.obj_history) == want.i + 2
    assert rel_fro(res.w, want.w) < 1e-5 and rel_fro(res.h, want.h) < 1e-5
    out = sink.getvalue().splitlines()
    assert out[0].startswith("[0]: ") and out[-1] == "Algorithm converged (2)."


def test_checkpoints_and_resume_through_the_public_api(soft, tmp_path):
    from nmf_b200.mur import mur
    m, n, k = 120, 100, 5
    v = orc.planted(m, n, k, 0).astype(np.float64)
    w0, h0 = orc.seeded_init(m, n, k, 1)
    path = str(tmp_path / "ckpt")
    full = mur(v, k, distance_type="kl", min_iter=10**9, max_iter=20, w0=w0, h0=h0, verbose=False,
               checkpoint_every=8, checkpoint_path=path)
    z = np.load(path + ".npz", allow_pickle=True)
    assert int(z["i"]) == 16 and len(z["obj_history"]) == 18
    assert sorted(z.files) == ["experiment", "h", "i", "obj_history", "w"]
    tr = orc.mur_run(v, k, w0, h0, distance_type="kl", min_iter=10**9, max_iter=17)
    assert rel_fro(z["w"], tr.w) < 1e-5 and rel_fro(z["h"], tr.h) < 1e-5
    rest = mur(v, k, distance_type="kl", min_iter=10**9, max_iter=3, w0=z["w"], h0=z["h"], verbose=False)
    assert np.array_equal(rest.w, full.w) and np.array_equal(rest.h, full.h)


def test_negative_data_is_shifted_in_place_and_reuploaded(soft):
    from nmf_b200.mur import mur
    rng = np.random.default_rng(0)
    v = rng.random((70, 50)) - 0.25
    vmin, v_in = v.min(), v.copy()
    w0, h0 = orc.seeded_init(70, 50, 3, 1)
    res = mur(v_in, 3, distance_type="eu", min_iter=100, max_iter=5, w0=w0, h0=h0, verbose=False)
    assert np.isclose(v_in.min(), 0.0) and np.allclose(v_in, v - vmin)              # mur.py:99-102
    want = orc.mur_run(v - vmin, 3, w0, h0, distance_type="eu", min_iter=100, max_iter=5)
    assert rel_fro(res.w, want.w) < TOL and rel_fro(res.h, want.h) < TOL


@pytest.mark.parametrize("shape", [(5, 7, 2), (300, 70, 3), (70, 300, 3), (130, 260, 70)])
def test_odd_shapes_pad_and_unpad_correctly(soft, shape):
    """Smaller than a tile, flat, tall, and k = 70 (KP = 128): padding to 256 / KP must never leak into results."""
    from nmf_b200.admm import admm
    from nmf_b200.mur import mur
    m, n, k = shape
    v = orc.planted(m, n, k, 4).astype(np.float64)
    w0, h0 = orc.seeded_init(m, n, k, 5)
    res = mur(v, k, distance_type="eu", min_iter=10**9, max_iter=6, w0=w0, h0=h0, verbose=False)
    want = orc.mur_run(v, k, w0, h0, distance_type="eu", min_iter=10**9, max_iter=6)
    assert res.w.shape == (m, k) and res.h.shape == (k, n)
    assert rel_fro(res.w, want.w) < TOL and rel_fro(res.h, want.h) < TOL
    kw = dict(rho=5, reg_w=(0, "nn"), reg_h=(0.01, "l1n"))
    res = admm(v, k, min_iter=10**9, max_iter=4, w0=w0, h0=h0, verbose=False, nndsvd_init=(False, "zero"), **kw)
    want = orc.admm_run(v, k, w0, h0, min_iter=10**9, max_iter=4, **kw)
    assert rel_fro(res.w, want.w) < TOL and rel_fro(res.h, want.h) < TOL


def test_errors_surface_like_the_reference(soft):
    from nmf_b200.admm import admm
    from nmf_b200.ao_admm import ao_admm
    v = np.random.default_rng(0).random((128, 96))
    np.random.seed(1)
    with pytest.raises(np.linalg.LinAlgError):             # H collapses to 0: rho = tr(0)/k = 0 (ao_admm.py:55)
        ao_admm(v, 4, reg_w=(0, "nn"), reg_h=(500.0, "l1n"), admm_iter=5, min_iter=100, max_iter=50,
                nndsvd_init=(False, "zero"), verbose=False)
    with pytest.raises(TypeError):
        admm(v, 4, reg_h=(0.1, "bogus"), nndsvd_init=(False, "zero"), max_iter=2, verbose=False)
    with pytest.raises(TypeError):
        admm(v, 4, distance_type="xx", nndsvd_init=(False, "zero"), max_iter=2, verbose=False)
    with pytest.raises(NotImplementedError):
        admm(v, 4, reg_h=(0.1, "l1inf"), nndsvd_init=(False, "zero"), max_iter=2, verbose=False)


def _upload_worker(rank, world, port, ret):
    import torch
    import torch.distributed as dist
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        m, n, k = 700, 530, 4                                # ragged last windows in both directions
        rng = np.random.default_rng(3)
        cases = {
            "bf16_exact": (orc.planted(m, n, k, 0), False),                       # one lossless plane
            "two_planes": (rng.random((m, n)), False),                            # fp64 input, hi + lo
            "negative": (rng.random((m, n)).astype(np.float32) - 0.5, False),     # min(V) < 0 must reach every rank
            "fixed": (orc.planted(m, n, k, 1), True),                             # exact in two fixed-point planes
            "fixed_fallback": (rng.random((m, n)), True),                         # not exact: floating planes again
        }
        out = {}
        for name, (v, fixed) in cases.items():
            got = []
            for full_walk in ("1", "0"):
                os.environ["NMFB_FULL_WALK_UPLOAD"] = full_walk
                eng = softdev.SoftEngine.for_comm(m, n, k, True)
                eng.ROW_CHUNK_BYTES = 64 * n * 8                                  # several chunks per shard
                vmin = eng.upload_data(np.array(v), fixed=fixed)
                got.append((vmin, eng.nv, eng.fixed, eng.v_rows.clone(), eng.v_cols.clone()))
            a, b = got
            out[name] = (a[0] == b[0], a[1] == b[1], a[2] == b[2], bool(torch.equal(a[3], b[3])),
                         bool(torch.equal(a[4], b[4])), float(a[0]), int(a[1]), bool(a[2]),
                         bool(b[3].any() and b[4].any()))
        os.environ.pop("NMFB_FULL_WALK_UPLOAD", None)
        ret[rank] = out
    finally:
        dist.destroy_process_group()


def test_sharded_upload_reads_only_the_ranks_rows_and_builds_identical_planes():
    """Engine._ingest_pass_sharded (each rank uploads its rows, the column copy comes from an all-to-all of the
    row shards) against the full walk of V that ran on the GPUs in round 1: same planes bit for bit, same
    min(V), same plane count and fixed / floating decision, on 3 ranks with ragged windows."""
    import torch.multiprocessing as mp
    softdev.emu_lib()
    world = 3
    ctx = mp.get_context("spawn")
    ret = ctx.Manager().dict()
    port = 29800 + (os.getpid() % 90)
    procs = [ctx.Process(target=_upload_worker, args=(r, world, port, ret)) for r in range(world)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(600)
    assert all(p.exitcode == 0 for p in procs), [p.exitcode for p in procs]
    for rank in range(world):
        for name, res in ret[rank].items():
            assert all(res[:5]) and res[8], (rank, name, res)
    r0 = ret[0]
    assert r0["bf16_exact"][6] == 1 and r0["two_planes"][6] == 2 and r0["negative"][5] < 0
    assert r0["fixed"][7] is True and r0["fixed_fallback"][7] is False
    assert all(ret[r][nm][5] == r0[nm][5] for r in range(world) for nm in r0)     # every rank agrees on min(V)


def test_bench_time_to_tol_leg_runs_the_public_api_with_the_stop_test_on(soft, monkeypatch):
    """bench.run_time_to_tol (the north-star's end-to-end time-to-tol): reference stop rule with tol2 = 1e-5 x the
    initial objective, seeded default init, bounded iteration count; all four bench methods."""
    import torch
    import bench
    monkeypatch.setattr(torch.cuda, "synchronize", lambda *a, **k: None)
    monkeypatch.setattr(bench, "TTT_MAX_ITER", 40)
    m, n, k = 96, 80, 4
    v = orc.planted(m, n, k, 0)
    for method, dist_, extra in (("mur", "eu", {}), ("mur", "kl", {}),
                                 ("admm", "eu", dict(rho=100, reg_w=(0, "nn"), reg_h=(0.1, "l1n"))),
                                 ("ao_admm", "eu", dict(reg_w=(0, "nn"), reg_h=(0.1, "l1n"), admm_iter=5))):
        np.random.seed(1)
        w0, h0 = orc.reference_init(m, n, k)
        obj0 = orc.objective(v.astype(np.float64), w0 @ h0, dist_)
        t = bench.run_time_to_tol(v, method, dist_, k, extra, obj0)
        assert t["seconds"] > 0 and 11 <= t["iterations"] <= 40 and t["tol2"] == pytest.approx(1e-5 * obj0)
        assert t["converged"] == (t["iterations"] < 40)
        assert t["objective_last"] < t["objective_first"] == pytest.approx(obj0, rel=1e-4)
        fn = {"mur": orc.mur_run, "admm": orc.admm_run, "ao_admm": orc.ao_admm_run}[method]
        kw = dict(extra, distance_type=dist_) if method == "mur" else dict(extra)
        want = fn(v.astype(np.float64), k, w0, h0, min_iter=10, max_iter=40, tol1=0.0, tol2=1e-5 * obj0, **kw)
        assert t["iterations"] == want.i + 1                 # stops where the reference stops


def _bench_e2e_worker(rank, world, port, ret):
    import torch
    import torch.distributed as dist
    import bench
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        torch.cuda.synchronize = lambda *a, **k: None
        torch.cuda.empty_cache = lambda *a, **k: None
        m, n, k, steps = 600, 520, 6, 5
        out = {}
        with softdev.soft_device():
            for method, dist_, extra in (("mur", "eu", {}),
                                         ("admm", "eu", dict(rho=100, reg_w=(0, "nn"), reg_h=(0.1, "l1n")))):
                eng = softdev.SoftEngine.for_comm(m, n, k, True)
                bench.synth_fill_engine(eng, seed=0, fixed=(method == "admm"))
                bench.synth_init(eng, seed=1)
                e = bench.run_e2e_sharded(eng, method, dist_, m, n, k, extra, steps, world, rank)
                out[method] = e
        ret[rank] = out
    finally:
        dist.destroy_process_group()


def test_bench_sharded_end_to_end_leg_on_two_ranks():
    """bench.run_e2e_sharded (the N > 1 `e2e` figure): host V regenerated from the per-chunk counters, the public
    call with comm=True on every rank, agreement protocol, byte accounting."""
    import torch.multiprocessing as mp
    softdev.emu_lib()
    world = 2
    ctx = mp.get_context("spawn")
    ret = ctx.Manager().dict()
    port = 29700 + (os.getpid() % 90)
    procs = [ctx.Process(target=_bench_e2e_worker, args=(r, world, port, ret)) for r in range(world)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(600)
    assert all(p.exitcode == 0 for p in procs), [p.exitcode for p in procs]
    for method in ("mur", "admm"):
        a, b = ret[0][method], ret[1][method]
        assert "error" not in a and a["value"] > 0 and a["value"] == b["value"]          # max over ranks, agreed
        assert a["h2d_bytes_per_step"] == (600 * 520 * 4 + 2 * (600 * 6 + 6 * 520) * 4) // 5
        assert set(a["breakdown"]) == {"engine_and_peer_arena_s", "upload_s", "iterations_s", "download_s"}
        assert a["unit"] == "it/s" and a["host_v_pinned"] is False


def test_too_many_ranks_for_the_problem_is_refused_on_every_rank():
    """Shards are multiples of 256 rows / columns: a rank without rows or columns would leave the others waiting
    in their first collective, so every rank judges every rank's window and all of them raise."""
    for rank in range(4):
        with pytest.raises(ValueError, match="without rows or columns"):
            softdev.SoftEngine(600, 5000, 4, rank=rank, world=4)       # 600 rows: ranks 3 has none
    eng = softdev.SoftEngine(600, 5000, 4, rank=2, world=3)
    assert eng.row_range == (512, 600) and eng.col_range[0] == 2 * eng.per_n
