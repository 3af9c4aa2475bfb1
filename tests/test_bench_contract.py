"""CPU: the parts of bench.py's contract that do not need a GPU -- the reference arm prints exactly one
JSON line with the agreed keys, workload names follow BASELINE.json's configs, algorithmic FLOP counts
match SURVEY.md 8(d), and the committed ncu traffic figures are keyed by existing workloads."""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

import bench  # noqa: E402


def test_reference_arm_prints_one_json_line():
    out = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--workload",
                          "mur_eu_4096x4096_k64", "--steps", "1", "--warmup", "0"],
                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True, timeout=300, check=True).stdout
    lines = [ln for ln in out.splitlines() if ln.strip()]
    assert len(lines) == 1
    d = json.loads(lines[0])
    assert d["impl"] == "reference" and d["metric"] == "nmf_outer_iterations_per_second" and d["unit"] == "it/s"
    assert d["higher_is_better"] is True and d["value"] > 0 and d["gpu_launches"] == 0
    # "reference" where __graft_entry__.build() has vendored the reference into oracle/_ref, else the oracle port
    want_kind = "reference" if os.path.exists(os.path.join(ROOT, "oracle", "_ref", "nmf", "__init__.py")) else "port"
    assert d["cpu_baseline"]["kind"] == want_kind and d["cpu_baseline"]["cores"] >= 1 and d["cpu_baseline"]["sample"]
    assert d["cpu_baseline"]["scaled_from_sample"] is True
    assert d["e2e"] == {"value": d["value"], "unit": "it/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}
    assert d["config"]["workload"] == "mur_eu_4096x4096_k64"


def test_vendored_reference_is_the_unmodified_package():
    """oracle/_ref/nmf (made by __graft_entry__.vendor_reference) is byte for byte /root/reference/nmf where both
    exist, and bench.load_reference() imports it under a private name -- not the drop-in shim `nmf/`."""
    import filecmp
    src, dst = "/root/reference/nmf", os.path.join(ROOT, "oracle", "_ref", "nmf")
    if not (os.path.isdir(src) and os.path.isdir(dst)):
        import pytest
        pytest.skip("needs the build container's /root/reference and a built oracle/_ref")
    names = [f for f in os.listdir(src) if f.endswith(".py")]
    match, mismatch, errors = filecmp.cmpfiles(src, dst, names, shallow=False)
    assert sorted(match) == sorted(names) and not mismatch and not errors
    ref = bench.load_reference()
    assert ref.__name__ == "_nmf_reference" and ref.NMF.__module__.startswith("_nmf_reference")


def test_workloads_cover_baseline_configs():
    w = bench.WORKLOADS
    assert bench.DEFAULT_WORKLOAD == "mur_eu_65536x65536_k128" and bench.DEFAULT_WORKLOAD in w
    assert w["mur_eu_32768x32768_k64"][:5] == ("mur", "eu", 32768, 32768, 64)                  # configs[1]
    assert w["mur_kl_20000x50000_k128"][:5] == ("mur", "kl", 20000, 50000, 128)                # configs[2]
    assert w["admm_65536x16384_k128"][5]["reg_h"] == (0.1, "l1n")                              # configs[3]
    assert w["ao_admm_65536x16384_k128"][5]["admm_iter"] == 5
    assert w["mur_eu_262144x131072_k256"][2:5] == (262144, 131072, 256)                        # configs[4]
    for (kernel, workload) in bench.NCU_TRAFFIC:
        assert workload in w


def test_algorithmic_flops_follow_the_survey():
    m, n, k = 65536, 65536, 128
    assert bench.alg_flops_per_iter("mur", "eu", m, n, k, {}) == 4.0 * m * n * k + 4.0 * (m + n) * k * k
    assert bench.alg_flops_per_iter("mur", "kl", m, n, k, {}) == 8.0 * m * n * k
    f = bench.alg_flops_per_iter("ao_admm", "eu", m, n, k, {"admm_iter": 5})
    assert f == 4.0 * m * n * k + 12.0 * (m + n) * k * k + 4.0 / 3.0 * k ** 3


def test_cpu_sample_is_a_block_of_the_workload():
    ms, ns = bench.cpu_sample_shape(65536, 16384, 128, 1 << 26)
    assert ms * ns <= 1 << 26 and 65536 % ms == 0 and 16384 % ns == 0


def test_roofline_reports_are_per_rank_and_pick_the_binding_roof():
    """solvers.*Stepper.roofline on a stand-in engine: algorithmic work of ONE launch on ONE rank's shard
    (a sharded run must not report the whole job's FLOPs against one GPU's peak)."""
    from types import SimpleNamespace
    from nmf_b200 import solvers
    peaks = dict(hbm_gbs=6500.0, bf16_tflops=1600.0, bf16_tflops_sustained=1400.0, source="measured")
    m, n, k = 65536, 65536, 128
    for world in (1, 8):
        eng = SimpleNamespace(m=m, n=n, k=k, KP=128, world=world, nv=1, fixed=False)
        ms = 1.6 / world
        r = solvers.MurEuStepper(eng).roofline({"gemm_w": ms, "gemm_h": ms}, peaks)
        assert r["kernel"] == "panel_gemm_kernel" and r["bound"] == "hbm"          # 2-byte V read once: HBM side
        assert abs(r["achieved"] - 2.0 * m * n / 1.6e-3 / 1e9) < 1e-6 * r["achieved"]
        assert r["frac"] == r["achieved"] / r["peak"] and r["unit"] == "GB/s" and r["peak"] == 6500.0
        assert 0 < r["frac"] < 1 and 0 < r["tensor"]["frac"] < 1
        q = solvers.MurKlStepper(eng).roofline({"kl_w": ms, "kl_h": ms}, peaks)
        assert q["kernel"] == "kl_panel_ts2_kernel" and q["unit"] == "TFLOP/s"
        assert abs(q["achieved"] - 4.0 * m * n * k / 1.6e-3 / 1e12) < 1e-6 * q["achieved"]
        assert 0 < q["frac"] < 1 and 0 < q["hbm"]["frac"] < 1
    # exact-hi ADMM contraction: 5 executed MMAs per algorithmic product
    eng = SimpleNamespace(m=m, n=16384, k=k, world=1, nv=2, fixed=True)
    st = solvers.AdmmStepper(eng, rho=100, reg_w=(0, "nn"), reg_h=(0.1, "l1n"))
    r = st.roofline({"gemm_w": 1.0, "gemm_h": 1.0}, peaks)
    assert r["executed_mma_per_product"] == 5 and (r["kernel"], "admm_65536x16384_k128") in bench.NCU_TRAFFIC


def test_cpu_arm_step_bodies_are_the_oracle_loops():
    """bench.cpu_step_fn (the `--impl reference` / cpu_baseline arm) advances exactly like the oracle drivers, for
    every workload family incl. the Kullback-Leibler branch of admm / ao_admm."""
    import numpy as np
    from oracle import nmf_oracle as orc
    m, n, k = 60, 50, 4
    v = orc.planted(m, n, k, 0).astype(np.float64)
    w0, h0 = orc.seeded_init(m, n, k, 1)
    admm_kw = dict(rho=100, reg_w=(0, "nn"), reg_h=(0.1, "l1n"))
    ao_kw = dict(reg_w=(0, "nn"), reg_h=(0.1, "l1n"), admm_iter=5)
    for method, dist, extra, ref in (
        ("mur", "eu", {}, lambda: orc.mur_run(v, k, w0, h0, distance_type="eu", min_iter=10**9, max_iter=4)),
        ("mur", "kl", {}, lambda: orc.mur_run(v, k, w0, h0, distance_type="kl", min_iter=10**9, max_iter=4)),
        ("admm", "eu", admm_kw, lambda: orc.admm_run(v, k, w0, h0, min_iter=10**9, max_iter=4, **admm_kw)),
        ("admm", "kl", admm_kw, lambda: orc.admm_run(v, k, w0, h0, distance_type="kl", min_iter=10**9, max_iter=4, **admm_kw)),
        ("ao_admm", "eu", ao_kw, lambda: orc.ao_admm_run(v, k, w0, h0, min_iter=10**9, max_iter=4, **ao_kw)),
        ("ao_admm", "kl", ao_kw, lambda: orc.ao_admm_run(v, k, w0, h0, distance_type="kl", min_iter=10**9, max_iter=4, **ao_kw)),
    ):
        step, cores = bench.cpu_step_fn(method, dist, m, n, k, extra)
        got = [step() for _ in range(4)]
        np.testing.assert_allclose(got, ref().obj_history[1:], rtol=1e-10)
        assert cores >= 1
