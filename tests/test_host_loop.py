"""CPU: the outer-iteration driver (`nmf_b200/_loop.py`) against the oracle, with a stub engine whose
"kernels" are the oracle's NumPy update rules.  Covers what the GPU cannot make visible: the one-step
look-ahead, the lagged objective of MUR-KL (objective of iteration i-1 produced inside iteration i),
the factor snapshots that make a stop return the factors OF the stopping iteration, and the
reference's history / index conventions (mur.py:119-146)."""
import os
import sys
from types import SimpleNamespace

import numpy as np
import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

from nmf_b200 import _loop  # noqa: E402
from oracle import nmf_oracle as orc  # noqa: E402


class StubEngine:
    """Holds W^T / H as torch tensors named like the device engine's fields."""

    def __init__(self, v, w0, h0, kind):
        self.device = torch.device("cpu")
        self.v, self.kind = v, kind
        self.scalars = torch.zeros(64, dtype=torch.float64)
        self.w = SimpleNamespace(master=torch.from_numpy(w0.T.copy()))
        self.h = SimpleNamespace(master=torch.from_numpy(h0.copy()))

    def factors(self):
        return self.w.master.numpy().T.copy(), self.h.master.numpy().copy()

    def objective(self):
        w, h = self.factors()
        return orc.objective(self.v, w @ h, self.kind)

    def update(self):
        w, h = self.factors()
        w = orc.mur_w_step(self.kind, self.v, w, h, w @ h, 0.0)
        h = orc.mur_h_step(self.kind, self.v, w, h, w @ h, 0.0)
        self.w.master.copy_(torch.from_numpy(w.T.copy()))
        self.h.master.copy_(torch.from_numpy(h.copy()))


def _case(kind):
    m, n, k = 60, 45, 4
    v = orc.planted(m, n, k, 0).astype(np.float64)
    w0, h0 = orc.seeded_init(m, n, k, 1)
    return v, k, w0, h0


@pytest.mark.parametrize("kind", ["eu", "kl"])
@pytest.mark.parametrize("stop", [False, True])
def test_plain_and_lagged_loops_reproduce_the_reference_loop(kind, stop):
    v, k, w0, h0 = _case(kind)
    kw = dict(min_iter=3, max_iter=400, tol1=1e-3, tol2=0.5) if stop else dict(min_iter=10**9, max_iter=17,
                                                                                tol1=1e-5, tol2=1e-5)
    want = orc.mur_run(v, k, w0, h0, distance_type=kind, **kw)
    if stop:
        assert want.i < kw["max_iter"] - 1, "the case must stop early"

    for lagged in (False, True):
        eng = StubEngine(v, w0, h0, kind)
        obj0 = eng.objective()
        calls = []

        def step_plain(slot):
            eng.update()
            eng.scalars[slot * 8] = eng.objective()

        def step_lagged(slot, prev, mid):
            eng.scalars[prev * 8] = eng.objective()          # objective of the pair the W pass starts from
            calls.append(prev)
            if mid is not None:
                mid()
            eng.update()

        def finish(slot):
            eng.scalars[slot * 8] = eng.objective()

        if lagged:
            lr = _loop.run(eng, step_lagged, obj0=obj0, precision=5, verbose=False, lagged=True, finish=finish, **kw)
        else:
            lr = _loop.run(eng, step_plain, obj0=obj0, precision=5, verbose=False, **kw)
        assert lr.i == want.i
        assert len(lr.obj_history) == len(want.obj_history) == want.i + 2
        np.testing.assert_allclose(lr.obj_history, want.obj_history, rtol=1e-12)
        if lr.masters is not None:                           # stopped: factors of the stopping iteration
            w = lr.masters[0].numpy().T
            h = lr.masters[1].numpy()
        else:
            w, h = eng.factors()
        np.testing.assert_allclose(w, want.w, rtol=1e-12)
        np.testing.assert_allclose(h, want.h, rtol=1e-12)
        if stop:
            assert lr.masters is not None


def test_loop_prints_like_the_reference(capsys):
    v, k, w0, h0 = _case("eu")
    eng = StubEngine(v, w0, h0, "eu")

    def step(slot):
        eng.update()
        eng.scalars[slot * 8] = eng.objective()

    lr = _loop.run(eng, step, obj0=eng.objective(), min_iter=100, max_iter=3, tol1=1e-5, tol2=1e-3, precision=5)
    out = capsys.readouterr().out.splitlines()
    assert out == ['[{}]: {:.5f}'.format(i, lr.obj_history[i + 1]) for i in range(3)]      # mur.py:128


@pytest.mark.parametrize("lagged", [False, True])
def test_checkpoints_hold_the_factors_of_their_iteration(lagged):
    """mur.py:138-140 (the commented `if i % 100 == 0: save_results(...)` hook): every N-th iteration the
    callback sees the factors OF that iteration although the device is already further along."""
    kind = "kl" if lagged else "eu"
    v, k, w0, h0 = _case(kind)
    iters, every = 11, 4
    want = orc.mur_run(v, k, w0, h0, distance_type=kind, min_iter=10**9, max_iter=iters,
                       snap=lambda i: i % every == 0)
    eng = StubEngine(v, w0, h0, kind)
    seen = []

    def on_checkpoint(i, masters, hist):
        seen.append((i, masters[0].numpy().T.copy(), masters[1].numpy().copy(), len(hist)))

    def step_plain(slot):
        eng.update()
        eng.scalars[slot * 8] = eng.objective()

    def step_lagged(slot, prev, mid):
        eng.scalars[prev * 8] = eng.objective()
        if mid is not None:
            mid()
        eng.update()

    def finish(slot):
        eng.scalars[slot * 8] = eng.objective()

    common = dict(obj0=eng.objective(), min_iter=10**9, max_iter=iters, tol1=1e-5, tol2=1e-5, precision=5,
                  verbose=False, checkpoint_every=every, on_checkpoint=on_checkpoint)
    if lagged:
        _loop.run(eng, step_lagged, lagged=True, finish=finish, **common)
    else:
        _loop.run(eng, step_plain, **common)
    assert [s[0] for s in seen] == [0, 4, 8]
    for (i, w, h, nh), (j, w_ref, h_ref) in zip(seen, want.snaps):
        assert i == j and nh == i + 2                       # obj_history up to and including iteration i
        np.testing.assert_allclose(w, w_ref, rtol=1e-12)
        np.testing.assert_allclose(h, h_ref, rtol=1e-12)
