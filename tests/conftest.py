import json
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

GOLDEN_DIR = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


def load_golden(name):
    z = np.load(os.path.join(GOLDEN_DIR, name + ".npz"))
    g = {k: z[k] for k in z.files}
    g["recipe"] = json.loads(str(g["recipe"]))
    kw = g["recipe"]["kw"]
    for key in ("reg_w", "reg_h", "nndsvd_init"):        # json turned tuples into lists
        if key in kw:
            kw[key] = tuple(kw[key])
    return g


def golden_v(spec):
    from oracle.nmf_oracle import planted, planted_with_zeros
    if spec["kind"] == "uniform":
        return np.random.default_rng(spec["seed"]).random((spec["m"], spec["n"]))
    if spec["kind"] == "zeros":
        return planted_with_zeros(spec["m"], spec["n"], spec["k"], spec["seed"]).astype(np.float64)
    return planted(spec["m"], spec["n"], spec["k"], spec["seed"]).astype(np.float64)


def rel_fro(a, b):
    return float(np.linalg.norm(np.asarray(a, np.float64) - b) / np.linalg.norm(b))


def golden_subset(length, stride):
    """Lines of a factor frozen in the mid_* / full_* goldens (tests/golden/make_golden_mid.py)."""
    return np.union1d(np.arange(0, length, stride), np.arange(max(0, length - 256), length))


_ALL = sorted(f[:-4] for f in os.listdir(GOLDEN_DIR) if f.endswith(".npz"))
# SURVEY 8d sizes (make_golden_mid.py: obj_history + a row / column subset of the final factors)
MID_GOLDEN_NAMES = [nm for nm in _ALL if nm.startswith("mid_")]
FULL_GOLDEN_NAMES = [nm for nm in _ALL if nm.startswith("full_")]
GOLDEN_NAMES = [nm for nm in _ALL if nm not in MID_GOLDEN_NAMES and nm not in FULL_GOLDEN_NAMES]
