"""CPU: the host side of Engine.upload_data -- pass structure (absmax -> fixed ingest -> floating fallback ->
second plane), row chunking and the chunk cache that lets V cross PCIe once per upload -- with the device
calls stubbed out (no kernel runs here; what the kernels compute is the GPU tests' business)."""
import os
import sys

import numpy as np
import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

from nmf_b200.engine import Engine  # noqa: E402


class HostOnlyEngine(Engine):
    """Engine with every device call replaced by a recorder.  `verdicts` is what finish_ingest reports per
    pass: (vmin, inexact-or-need_lo, non-finite)."""

    ROW_CHUNK_BYTES = 64 * 40 * 8          # 64 rows of 40 doubles per chunk

    def __init__(self, m, n, verdicts, free_bytes):
        self.m, self.n, self.world = m, n, 1
        self.device = torch.device("cpu")
        self.verdicts = list(verdicts)
        self.free_bytes = free_bytes
        self.log = []                      # (pass kind, nv, fixed)
        self.seen = []                     # per pass: list of (r0, chunk)
        self.to_device = 0

    def begin_ingest(self, nv, fixed=False):
        self.fixed = bool(fixed)
        self.log.append(("begin", nv, fixed))

    def absmax_chunk(self, chunk, r0):
        self.log.append(("absmax", r0, chunk.shape[0]))

    def ingest_chunk(self, chunk, r0):
        self.log.append(("ingest", r0, chunk.shape[0]))
        self.seen.append((r0, chunk))

    def finish_ingest(self):
        return self.verdicts.pop(0)


@pytest.fixture
def counting_transfers(monkeypatch):
    """Counts host -> device conversions and fakes the HBM head-room query."""
    state = {"n": 0, "free": 1 << 40}
    real = torch.from_numpy

    def from_numpy(a):
        state["n"] += 1
        return real(a)

    monkeypatch.setattr(torch, "from_numpy", from_numpy)
    monkeypatch.setattr(torch.cuda, "mem_get_info", lambda device=None: (state["free"], 1 << 41))
    return state


def _v(m=200, n=40):
    """Not bf16-exact (values up to 7999): the host sample sends it straight to hi + lo planes."""
    return np.arange(m * n, dtype=np.float64).reshape(m, n)


def _v_exact(m=200, n=40):
    """Every value fits one bf16 plane."""
    return (np.arange(m * n, dtype=np.float64) % 128).reshape(m, n)


def _v_exact_in_the_sample_only(m=200, n=40):
    """bf16-exact in every row the host sample reads, not in one row it skips."""
    sampled = set(np.unique(np.linspace(0, m - 1, num=min(m, 32)).astype(np.int64)).tolist())
    hidden = next(r for r in range(m) if r not in sampled)
    v = _v_exact(m, n)
    v[hidden, 3] = 1000.5
    return v


def test_single_plane_upload_is_one_pass_in_64_row_chunks(counting_transfers):
    v = _v_exact()
    eng = HostOnlyEngine(200, 40, [(0.0, False, False)], 0)
    assert eng.upload_data(v) == 0.0 and eng.vmin == 0.0
    assert [e for e in eng.log if e[0] == "begin"] == [("begin", 1, False)]
    assert [(e[1], e[2]) for e in eng.log if e[0] == "ingest"] == [(0, 64), (64, 64), (128, 64), (192, 8)]
    assert counting_transfers["n"] == 4
    for r0, chunk in eng.seen:
        assert np.array_equal(chunk.numpy(), v[r0:r0 + chunk.shape[0]])
    assert eng._chunk_cache is None and eng._keep_chunks is False      # nothing outlives the call


def test_inexact_data_goes_straight_to_two_planes(counting_transfers):
    """The host sample sees values one bf16 plane cannot hold: one pass with hi + lo planes, V crosses PCIe once."""
    v = _v()
    eng = HostOnlyEngine(200, 40, [(1.0, False, False)], 0)
    assert eng.upload_data(v) == 1.0
    assert [e for e in eng.log if e[0] == "begin"] == [("begin", 2, False)]
    assert counting_transfers["n"] == 4 and len(eng.seen) == 4


def test_a_sample_that_misses_the_inexact_value_costs_a_second_pass(counting_transfers):
    v = _v_exact_in_the_sample_only()
    eng = HostOnlyEngine(200, 40, [(1.0, True, False), (1.0, False, False)], 0)
    assert eng.upload_data(v) == 1.0
    assert [e for e in eng.log if e[0] == "begin"] == [("begin", 1, False), ("begin", 2, False)]
    assert counting_transfers["n"] == 8                                  # floating uploads keep no device chunks
    assert len(eng.seen) == 8 and [s[0] for s in eng.seen] == [0, 64, 128, 192] * 2


def test_explicit_plane_count_is_obeyed(counting_transfers):
    eng = HostOnlyEngine(200, 40, [(1.0, True, False)], 0)
    eng.upload_data(_v(), planes=1)                                      # caller insists on one (lossy) plane
    assert [e for e in eng.log if e[0] == "begin"] == [("begin", 1, False)]


def test_fixed_point_fallback_without_room_transfers_every_pass(counting_transfers):
    counting_transfers["free"] = 1 << 10
    v = _v()
    eng = HostOnlyEngine(200, 40, [(2.0, True, False), (2.0, False, False)], 0)
    eng.upload_data(v, fixed=True)
    # |max| pass + fixed ingest + the floating fallback (straight to two planes): three walks, nothing kept
    assert counting_transfers["n"] == 12


def test_fixed_point_upload_passes_and_fallback(counting_transfers):
    v = _v()
    # exact in two fixed-point planes: absmax pass + ingest pass, one transfer of V
    eng = HostOnlyEngine(200, 40, [(2.0, False, False)], 0)
    assert eng.upload_data(v, fixed=True) == 2.0
    kinds = [e[0] for e in eng.log]
    assert kinds == ["begin"] + ["absmax"] * 4 + ["ingest"] * 4 and eng.log[0] == ("begin", 2, True)
    assert counting_transfers["n"] == 4
    # not exact: falls back to floating planes (the sample says two), still one transfer
    counting_transfers["n"] = 0
    eng = HostOnlyEngine(200, 40, [(2.0, True, False), (2.0, False, False)], 0)
    eng.upload_data(v, fixed=True)
    assert [e for e in eng.log if e[0] == "begin"] == [("begin", 2, True), ("begin", 2, False)]
    assert eng.fixed is False and counting_transfers["n"] == 4
    # same with data the sample finds exact: one floating plane
    counting_transfers["n"] = 0
    eng = HostOnlyEngine(200, 40, [(2.0, True, False), (2.0, False, False)], 0)
    eng.upload_data(_v_exact(), fixed=True)
    assert [e for e in eng.log if e[0] == "begin"] == [("begin", 2, True), ("begin", 1, False)]
    assert eng.fixed is False and counting_transfers["n"] == 4


def test_upload_rejects_bad_input(counting_transfers):
    eng = HostOnlyEngine(200, 40, [(0.0, False, True)], 0)
    with pytest.raises(ValueError, match="non-finite"):
        eng.upload_data(_v())
    assert eng._chunk_cache is None                                      # released on the error path too
    with pytest.raises(ValueError, match="shape"):
        eng.upload_data(_v(100, 40))
    # integer data is promoted (the reference accepts any array NumPy can multiply), Fortran order is copied
    eng = HostOnlyEngine(200, 40, [(0.0, False, False)], 0)
    eng.upload_data(np.asfortranarray(_v().astype(np.int64)))
    assert all(c.dtype == torch.float64 and c.is_contiguous() for _, c in eng.seen)


def test_host_sample_for_the_plane_count():
    """engine._sample_needs_lo_plane: bf16-exact data of either dtype -> one plane; anything one bf16 plane cannot
    hold -> two (non-finite values may fall either way, the ingest pass rejects them); wide matrices are sampled
    with a column stride, tiny ones entirely."""
    from nmf_b200.engine import _sample_needs_lo_plane as needs_lo
    from oracle import nmf_oracle as orc
    exact = orc.planted(300, 200, 8, 0)                                  # float32 on the bf16 grid
    assert not needs_lo(exact) and not needs_lo(exact.astype(np.float64))
    assert needs_lo(np.random.default_rng(0).random((300, 200)))
    assert needs_lo(np.random.default_rng(0).random((300, 200)).astype(np.float32))
    assert needs_lo(np.full((5, 7), np.nan)) and needs_lo(np.full((5, 7), np.inf, dtype=np.float32)) is False
    assert not needs_lo(np.zeros((1, 1)))
    wide = np.zeros((4, 40000), dtype=np.float32)
    wide[:, ::4] = 1.0 + 2.0 ** -12                                      # column stride 4 hits these
    assert needs_lo(wide)
    wide[:] = 0.0
    wide[:, 1::4] = 1.0 + 2.0 ** -12                                     # ... and misses these: the kernel's verdict decides
    assert not needs_lo(wide)
    assert needs_lo(np.array([[1e-45]], dtype=np.float64))               # not representable in fp32 at all
