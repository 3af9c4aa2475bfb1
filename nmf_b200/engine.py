"""Device-side state of one factorisation and the calls that advance it.

PyTorch is used for what it is good at here -- device memory, streams, pinned host buffers, the
process group -- and nothing else: every arithmetic step of every solver is a kernel of libnmf_b200.so
reached through the C ABI (`_lib.py`); no library GEMM runs on the loop.  There is no CPU path; constructing an Engine without a CUDA device raises.
"""
from __future__ import annotations

import ctypes as C
import os
from typing import Optional, Tuple

import numpy as np
import torch

from . import _lib
from ._lib import AdmmState, Data, Factor, PeerCtx, Problem, check
from .sharding import (allgather_flat, allgather_windows, allreduce_max_, allreduce_min_key_, allreduce_sum_,
                       exchange_blocks, live_rows, shard_bounds)


def _ptr(t: Optional[torch.Tensor]) -> Optional[int]:
    return None if t is None else t.data_ptr()


def _key_to_float(key: int) -> float:
    """Inverse of the order-preserving uint32 key the ingest kernel folds min(V) into."""
    if key == 0xFFFFFFFF:
        return float("inf")
    bits = (key & 0x7FFFFFFF) if (key & 0x80000000) else (~key & 0xFFFFFFFF)
    return float(np.array([bits], dtype=np.uint32).view(np.float32)[0])


def _sample_needs_lo_plane(v: np.ndarray) -> bool:
    """True when a strided sample of the host V (<= 32 rows x <= 8192 columns) holds a value that one bf16
    plane cannot represent exactly (non-finite values may fall either way: the full pass over V rejects them)."""
    m, n = v.shape
    rows = np.unique(np.linspace(0, m - 1, num=min(m, 32)).astype(np.int64))
    a = v[rows][:, ::max(1, n // 8192)]
    a32 = a.astype(np.float32, copy=False)
    exact = (np.ascontiguousarray(a32).view(np.uint32) & np.uint32(0xFFFF)) == 0
    if a.dtype != np.float32:
        exact &= a32.astype(a.dtype) == a
    return not bool(exact.all())


class FactorBuffers:
    """master / cplanes / rplanes of one factor (layouts in include/nmf_b200.h)."""

    def __init__(self, L: int, KP: int, device, Lp: Optional[int] = None, win0: int = 0, win_p: int = 0):
        self.L, self.KP = L, KP
        self.Lp = (L + 255) // 256 * 256 if Lp is None else Lp
        self.win0, self.win_p = win0, win_p
        self.master = torch.zeros((KP, self.Lp), dtype=torch.float32, device=device)
        self.cplanes = torch.zeros((3, 256, self.Lp), dtype=torch.bfloat16, device=device)
        self.rplanes = torch.zeros((2, self.Lp, KP), dtype=torch.bfloat16, device=device)
        # sharded runs on GPUs: the peer arena and region the windows of this factor are exchanged through
        # (Engine._attach_exchange), and how many exchanges it has seen (parity = which of the two buffers)
        self.arena, self.xname, self.xcount = None, None, 0

    def struct(self) -> Factor:
        return Factor(self.master.data_ptr(), self.cplanes.data_ptr(), self.rplanes.data_ptr(),
                      self.L, self.Lp, self.win0, self.win_p)


class Engine:
    """Owns the HBM-resident problem: V and V^T planes, both factors, Grams and workspaces."""

    ROW_CHUNK_BYTES = 512 << 20

    def __init__(self, m: int, n: int, k: int, device: Optional[torch.device] = None, rank: int = 0,
                 world: int = 1, group=None):
        self.lib, self.device, self.num_sms = self._open_device(device)
        self.timings = {}                     # wall seconds of the host-side legs (upload_s, download_s): bench e2e
        self.m, self.n, self.k = int(m), int(n), int(k)
        self.rank, self.world = int(rank), int(world)
        self.group = group
        self.KP = check(self.lib.nmfb_padded_components(self.k))
        # 1-D sharding (sharding.py): this rank owns rows [row_range) of V / W and columns [col_range) of V / H
        self.Mp, self.per_m, mb = shard_bounds(self.m, self.world)
        self.Np, self.per_n, nb = shard_bounds(self.n, self.world)
        self.row_win, self.col_win = mb[self.rank], nb[self.rank]
        self.row_range = (self.row_win[0], self.row_win[0] + live_rows(self.m, *self.row_win))
        self.col_range = (self.col_win[0], self.col_win[0] + live_rows(self.n, *self.col_win))
        # judged for ALL ranks on every rank, so that the whole group raises together instead of one rank
        # leaving the others waiting in their first collective
        empty = [p for p in range(self.world)
                 if live_rows(self.m, *mb[p]) <= 0 or live_rows(self.n, *nb[p]) <= 0]
        if empty:
            raise ValueError(f"{self.world} ranks leave rank {empty[0]} without rows or columns of a "
                             f"{self.m} x {self.n} problem (shards are multiples of 256); use fewer ranks")
        with self._device_ctx():
            sharded = self.world > 1
            self.w = FactorBuffers(self.m, self.KP, self.device, self.Mp, self.row_win[0] if sharded else 0,
                                   self.per_m if sharded else 0)
            self.h = FactorBuffers(self.n, self.KP, self.device, self.Np, self.col_win[0] if sharded else 0,
                                   self.per_n if sharded else 0)
            self.gram32 = torch.zeros((2, self.KP, self.KP), dtype=torch.float32, device=self.device)
            self.gram64 = torch.zeros((2, self.KP, self.KP), dtype=torch.float64, device=self.device)
            pbytes = int(self.lib.nmfb_partial_bytes(self.Mp, self.Np, self.k, self.num_sms))
            if self.world > 1:      # a shard has fewer panels -> more K-splits, at the full output pitch
                pbytes = max(pbytes,
                             self.world * int(self.lib.nmfb_partial_bytes(self.per_m, self.Np, self.k, self.num_sms)),
                             self.world * int(self.lib.nmfb_partial_bytes(self.Mp, self.per_n, self.k, self.num_sms)))
            self.partial = torch.empty(pbytes // 4, dtype=torch.float32, device=self.device)
            self.obj_partial_len = int(self.lib.nmfb_obj_partial_len(self.m, self.n, self.num_sms))
            self.obj_partial = torch.zeros(self.obj_partial_len, dtype=torch.float64, device=self.device)
            self.stats = torch.zeros(4, dtype=torch.int32, device=self.device)
            self.scalars = torch.zeros(64, dtype=torch.float64, device=self.device)
            self.ksum = torch.zeros((2, self.KP), dtype=torch.float64, device=self.device)
            self.peer = None
            if self.world > 1 and torch.device(self.device).type == "cuda":
                self._open_peer()
        self.nv = 0
        self.fixed = False                    # plane 0 of V is the fixed-point slice (exact-mode ADMM)
        self.v_rows = None
        self.v_cols = None
        self.vmin = None
        self._problem = None

    # The three places that touch the CUDA runtime directly (tests/softdev.py swaps them for a host stand-in
    # of the device to run the orchestration on the CPU; the product always goes through these).
    def _open_device(self, device):
        """-> (C-ABI library, torch device, SM count); raises without a B200."""
        if not torch.cuda.is_available():
            raise RuntimeError("nmf_b200 needs a CUDA device (B200, sm_100a); there is no CPU path")
        lib = _lib.load()
        dev = torch.device("cuda", torch.cuda.current_device()) if device is None else torch.device(device)
        props = torch.cuda.get_device_properties(dev)
        if props.major != 10:
            raise RuntimeError(f"nmf_b200 is built for sm_100a only; device is sm_{props.major}{props.minor}")
        return lib, dev, props.multi_processor_count

    def _device_ctx(self):
        return torch.cuda.device(self.device)

    def _sync_stream(self):
        torch.cuda.current_stream(self.device).synchronize()

    def _hbm_free(self) -> int:
        return torch.cuda.mem_get_info(self.device)[0]

    @classmethod
    def for_comm(cls, m: int, n: int, k: int, comm) -> "Engine":
        """Engine of this rank in a sharded run: comm = True (the default process group) or a
        torch.distributed group; one process per GPU, the current CUDA device is this rank's."""
        import torch.distributed as dist
        group = None if comm is True else comm
        return cls(m, n, k, rank=dist.get_rank(group), world=dist.get_world_size(group), group=group)

    # ------------------------------------------------------------------ plumbing
    @property
    def stream(self) -> int:
        return torch.cuda.current_stream(self.device).cuda_stream

    def _alloc_data(self, nv: int):
        self.v_rows = self.v_cols = None
        self.nv = nv
        # K-blocked planes: [plane][column block of 64][row][64]  (include/nmf_b200.h)
        self.v_rows = torch.zeros((nv, self.Np // 64, self.per_m, 64), dtype=torch.bfloat16, device=self.device)
        self.v_cols = torch.zeros((nv, self.Mp // 64, self.per_n, 64), dtype=torch.bfloat16, device=self.device)
        self._problem = None

    # ------------------------------------------------------------------ sharded runs
    def _open_peer(self):
        """Peer-memory arena of this rank (peer.py): barrier flags, the ranks' objective / norm partials, their
        terms of the two Gram matrices, and the exchange buffers of W and H -- everything peers write is
        double-buffered ([2] leading)."""
        from .peer import PeerArena
        KP, g, w = self.KP, self.KP * self.KP * 8, self.world
        self._norm_bytes = 4 * (max(self.Mp, self.Np) // 64) * 8
        layout = [("scal", 2 * w * 64), ("gram", 2 * 2 * w * g), ("norm", 2 * self._norm_bytes),
                  ("xw", 2 * KP * self.Mp * 4), ("xh", 2 * KP * self.Np * 4)]
        self.peer = PeerArena(self.lib, layout, self.device, self.rank, self.world, self.group)
        self._attach_exchange(self.w, self.peer, "xw")
        self._attach_exchange(self.h, self.peer, "xh")
        self._scal_count = self._norm_count = self._gram_count = 0
        # MUR-eu: a rank's term of the Gram matrix of the factor it just updated is contracted (tensor cores, a few
        # CTAs) and stored into the peers' tables on a side stream, under the NVLink-bound push of its window
        self.gram_term = torch.zeros((KP, KP), dtype=torch.float64, device=self.device)
        self.gram_term32 = torch.zeros((KP, KP), dtype=torch.float32, device=self.device)
        nbytes = int(self.lib.nmfb_gram_partial_bytes(self.k, max(self.per_m, self.per_n), self.num_sms))
        self.gram_scratch = torch.empty(max(nbytes, 4) // 4, dtype=torch.float32, device=self.device)
        self.gram_stream = torch.cuda.Stream(device=self.device)
        self._gram_events = (torch.cuda.Event(), torch.cuda.Event())
        for ev in self._gram_events:
            ev.record(torch.cuda.current_stream(self.device))            # materialises the cudaEvent_t
        self._peer_ctx = {}
        self.exchange_probe = None            # bench: list that collects the timing events of every exchange

    @staticmethod
    def _attach_exchange(fac: "FactorBuffers", arena, name: str):
        fac.arena, fac.xname, fac.xcount = arena, name, 0

    def gather_factor(self, fac: "FactorBuffers", refresh: bool = True, fixed: bool = False, gram_index=None):
        """Exchange the windows of a factor (or ADMM auxiliary) every rank just updated, then re-derive its bf16
        planes: `refresh` = floating cplanes + rplanes, `fixed` = fixed-point cplanes on top (operand of the
        exact-hi contractions).  On GPUs the window goes straight into the peers' exchange buffers (peer.cu), one
        cross-GPU barrier later every rank adopts all windows; gram_index = 0 (W^T W) / 1 (H H^T) also reduces the
        ranks' window terms of that Gram matrix in the same exchange (k x k all-reduce in rank order).  Without
        peer memory (host stand-in of the CPU tests) the windows are all-gathered by torch.distributed."""
        if self.world == 1:
            return
        fs = fac.struct()
        if self.peer is None:
            allgather_windows(fac.master, fac.win_p, self.group)
            if refresh:
                check(self.lib.nmfb_refresh_planes(self.KP, C.byref(fs), self.stream))
            if fixed:
                check(self.lib.nmfb_fixed_planes(self.KP, C.byref(fs), self.stream))
            return
        # ONE call of the C ABI per exchange (nmfb_exchange_factor): at 8 GPUs an iteration is ~1 ms, and eight
        # ctypes calls per exchange left the GPUs waiting for their hosts
        KP, w = self.KP, self.world
        g = KP * KP * 8
        arena = fac.arena
        par = fac.xcount & 1
        fac.xcount += 1
        xoff = arena.offsets[fac.xname] + par * KP * fac.Lp * 4
        goff, g64, g32 = 0, None, None
        if gram_index is not None:
            gpar = self._gram_count & 1
            self._gram_count += 1
            goff = arena.offsets["gram"] + (gpar * 2 + gram_index) * w * g
            g64 = self.gram64.data_ptr() + gram_index * g
            g32 = self.gram32.data_ptr() + gram_index * (g // 2)
        mode = 0 if not refresh else (2 if gram_index is not None else 1)     # 2: my window is current already
        arena.seq += 1
        check(self.lib.nmfb_exchange_factor(C.byref(self._ctx_of(arena)), KP, C.byref(fs), xoff, mode, int(fixed),
                                            2 if gram_index is not None else 0, goff, g64, g32,
                                            arena.seq & 0xFFFFFFFF, self._probe_events(), self.stream))

    def _ctx_of(self, arena) -> PeerCtx:
        """nmfb_peer_ctx of an arena (built once; the Gram workspaces belong to the main arena's exchanges)."""
        ctx = self._peer_ctx.get(id(arena))
        if ctx is None:
            ctx = PeerCtx(C.cast(arena._ptr_arr, C.c_void_p), arena.mc_ptr or None, self.rank, self.world,
                          self.gram_stream.cuda_stream, int(self._gram_events[0].cuda_event),
                          int(self._gram_events[1].cuda_event), self.gram_scratch.data_ptr(),
                          self.gram_scratch.numel() * 4, self.gram_term.data_ptr(), self.gram_term32.data_ptr(),
                          self.num_sms)
            self._peer_ctx[id(arena)] = ctx
        return ctx

    PROBE_LABELS = ("push", "barrier", "adopt", "gram_sum")

    def _probe_events(self):
        """bench: five timing events per exchange (start | window store + Gram term | barrier | adoption | Gram sum)."""
        if self.exchange_probe is None:
            return None
        evs = [torch.cuda.Event(enable_timing=True) for _ in range(5)]
        for ev in evs:
            ev.record(torch.cuda.current_stream(self.device))            # materialises the cudaEvent_t
        arr = (C.c_void_p * 5)(*[int(ev.cuda_event) for ev in evs])
        self.exchange_probe.append((evs, arr))
        return arr

    def _mark(self, idx: int):
        ev = getattr(self, "_phase_torch", None)
        if ev is not None and idx < len(ev):
            ev[idx].record(torch.cuda.current_stream(self.device))

    def full_grams(self):
        """W^T W and H H^T of the (replicated) factors into gram32 / gram64: start of a sharded MUR-eu run; from
        then on the exchanges keep them current."""
        p = self.problem()
        g = self.KP * self.KP
        for idx, fac in ((0, self.w), (1, self.h)):
            fs = fac.struct()
            check(self.lib.nmfb_gram(C.byref(fs), self.KP, 2, self.partial.data_ptr(), self.partial.numel() * 4,
                                     self.gram32.data_ptr() + 4 * idx * g, self.gram64.data_ptr() + 8 * idx * g,
                                     self.num_sms, self.stream))
        del p

    def mur_eu_window_step(self, side: int, lam: float):
        p = self.problem()
        check(self.lib.nmfb_mur_eu_window_step(C.byref(p), int(side), float(lam), self.stream))

    def gather_norm_partials(self, fac: "FactorBuffers"):
        """AO-ADMM residual test (ao_admm.py:33-43): every rank needs all blocks of the four norm partials."""
        nblk = fac.win_p // 64
        full = self.norm_partial[: 4 * (fac.Lp // 64)]
        if self.peer is None:
            mine = full[4 * nblk * self.rank: 4 * nblk * (self.rank + 1)].clone()
            allgather_flat(full, mine, self.group)
            return
        par = self._norm_count & 1
        self._norm_count += 1
        seg = 4 * nblk * 8
        base = par * self._norm_bytes
        self.peer.push(full.data_ptr() + seg * self.rank, seg, "norm", base + seg * self.rank, seg, 1, seg, self.stream)
        self.peer.barrier(self.stream)
        got = self.peer.view("norm", (2, self._norm_bytes // 8), torch.float64)[par, : full.numel()]
        full.copy_(got, non_blocking=True)

    def mur_kl_half_step(self, side: int, lam: float, obj_start_slot: Optional[int] = None):
        """obj_start_slot: scalar slot that receives the KL objective of the pair this pass starts from."""
        p = self.problem()
        out = None if obj_start_slot is None else self.scalars.data_ptr() + 8 * obj_start_slot
        check(self.lib.nmfb_mur_kl_half_step(C.byref(p), int(side), float(lam), out, self.stream))

    def admm_half_step(self, side: int, rho, lam, prox):
        p = self.problem()
        check(self.lib.nmfb_admm_half_step(C.byref(p), C.byref(self._admm), int(side), float(rho), float(lam),
                                           int(prox), self.stream))

    def admm_publish(self, ring_slot: int):
        check(self.lib.nmfb_admm_publish(C.byref(self._admm), self.scalars.data_ptr() + 64 * ring_slot, self.stream))

    def aoadmm_side(self, side: int, admm_iter: int, lam, prox):
        """One factor update of AO-ADMM in a sharded run (ao_admm.py:46-68): the four norms of the residual
        test are all-gathered per inner iteration so that every rank takes the same early break."""
        p = self.problem()
        fac = self.h if side == 0 else self.w
        slots = check(self.lib.nmfb_aoadmm_side_begin(C.byref(p), C.byref(self._admm), side, self.stream))
        if int(prox) == 2 and float(lam) != 0.0:             # l2n: the Tikhonov system needs this side's rho
            check(self.lib.nmfb_admm_prepare_l2n(C.byref(p), C.byref(self._admm), side, float(lam), self.stream))
        for it in range(admm_iter):
            check(self.lib.nmfb_aoadmm_inner(C.byref(p), C.byref(self._admm), side, 1 if it == 0 else 0, slots,
                                             float(lam), int(prox), self.stream))
            self.gather_norm_partials(fac)
            check(self.lib.nmfb_aoadmm_inner_flag(C.byref(p), C.byref(self._admm), side, self.stream))
        self.gather_factor(fac, refresh=True, fixed=self.fixed)

    def mur_eu_half_step(self, side: int, lam: float):
        p = self.problem()
        check(self.lib.nmfb_mur_eu_half_step(C.byref(p), int(side), float(lam), self.stream))

    def reduce_scalar(self, index: int):
        """Sum a mailbox scalar over ranks (each rank's objective covers its own rows of V)."""
        if self.world == 1:
            return
        if self.peer is None:
            allreduce_sum_(self.scalars[index:index + 1], self.group)
            return
        # every rank stores its partial into slot `rank` of every rank's table; summed in rank order after the
        # barrier, so that all ranks hold the bit-identical objective (and take the same stop decision)
        par = self._scal_count & 1
        self._scal_count += 1
        assert index % 2 == 0                                       # 16-byte granules
        self.peer.seq += 1
        ptr = self.scalars.data_ptr() + 8 * index
        check(self.lib.nmfb_exchange_scalar(C.byref(self._ctx_of(self.peer)), ptr,
                                            self.peer.offsets["scal"] + par * self.world * 64,
                                            self.peer.seq & 0xFFFFFFFF, ptr, self.stream))

    def _data_structs(self) -> Tuple[Data, Data]:
        vr = Data(self.v_rows.data_ptr(), self.row_range[1] - self.row_range[0], self.n, self.per_m, self.Np,
                  self.nv, 1, int(self.fixed), self.row_range[0] if self.world > 1 else 0)
        vc = Data(self.v_cols.data_ptr(), self.col_range[1] - self.col_range[0], self.m, self.per_n, self.Mp,
                  self.nv, 1, int(self.fixed), self.col_range[0] if self.world > 1 else 0)
        return vr, vc

    def problem(self) -> Problem:
        if self._problem is None:
            vr, vc = self._data_structs()
            self._problem = Problem(self.m, self.n, self.k, self.KP, vr, vc, self.w.struct(), self.h.struct(),
                                    self.gram32.data_ptr(), self.gram64.data_ptr(),
                                    self.partial.data_ptr(), self.partial.numel() * 4,
                                    self.obj_partial.data_ptr(), self.obj_partial_len, self.num_sms,
                                    None, 0, self.ksum.data_ptr())
        return self._problem

    def set_phase_events(self, events):
        """Profiling hook: torch.cuda.Event(enable_timing=True) objects (already recorded once, so
        that their cudaEvent_t exists) that the next solver step records at its phase boundaries.
        Pass None to switch the hook off."""
        p = self.problem()
        self._phase_torch = list(events) if events else None
        if not events:
            self._phase_arr = None
            p.phase_events, p.n_phase_events = None, 0
            return
        arr = (C.c_void_p * len(events))(*[int(e.cuda_event) for e in events])
        self._phase_arr = arr                       # keep alive
        p.phase_events = C.cast(arr, C.POINTER(C.c_void_p))
        p.n_phase_events = len(events)

    # ------------------------------------------------------------------ data in
    # chunk-wise ingest (host arrays and on-device synthetic data share it)
    def begin_ingest(self, nv: int, fixed: bool = False):
        self.fixed = bool(fixed)
        self._alloc_data(max(nv, 2) if fixed else nv)
        self.stats.copy_(torch.tensor([-1, 0, 0, 0], dtype=torch.int32))
        if fixed:
            self.rowmax = torch.zeros(self.Mp, dtype=torch.int32, device=self.device)
            self.colmax = torch.zeros(self.Np, dtype=torch.int32, device=self.device)

    def absmax_chunk(self, chunk: torch.Tensor, r0: int):
        """Pass 1 of a fixed-point ingest: fold |max| of rows r0.. of V into the row / column maxima."""
        check(self.lib.nmfb_data_absmax(chunk.data_ptr(), 1 if chunk.dtype == torch.float64 else 0, chunk.stride(0),
                                        r0, chunk.shape[0], self.n, self.rowmax.data_ptr(), self.colmax.data_ptr(),
                                        self.stream))

    def ingest_chunk(self, chunk: torch.Tensor, r0: int):
        vr, vc = self._data_structs()
        is64 = 1 if chunk.dtype == torch.float64 else 0
        if self.fixed:
            check(self.lib.nmfb_ingest_data_fixed(chunk.data_ptr(), is64, chunk.stride(0), r0, chunk.shape[0],
                                                  C.byref(vr), C.byref(vc), self.stats.data_ptr(),
                                                  self.rowmax.data_ptr(), self.colmax.data_ptr(), self.stream))
        else:
            check(self.lib.nmfb_ingest_data(chunk.data_ptr(), is64, chunk.stride(0), r0, chunk.shape[0],
                                            C.byref(vr), C.byref(vc), self.stats.data_ptr(), self.stream))

    def finish_ingest(self):
        if self.world > 1:                    # every rank walked all of V but judged only its shard's planes
            allreduce_max_(self.stats[1:3], self.group)
        st = self.stats.cpu().numpy().view(np.uint32)
        return _key_to_float(int(st[0])), bool(st[1]), bool(st[2])

    def _host_chunks(self, v: np.ndarray):
        """Row chunks of the host V on the device.  A fixed-point upload walks V twice (|max|, then the planes):
        when HBM has room for the whole fp32/fp64 V next to its bf16 planes, the chunks of the first walk are kept
        for the later ones, so V crosses PCIe once.  A floating upload is one walk (the plane count comes from a
        host sample, _upload_passes) and keeps nothing."""
        cache = getattr(self, "_chunk_cache", None)
        if cache is not None and cache[0] is v:
            yield from cache[1]
            return
        free = self._hbm_free()
        planes_bytes = 8 * self.m * self.n // self.world + (64 << 20)         # two copies x two bf16 planes
        keep = getattr(self, "_keep_chunks", False) and v.nbytes + planes_bytes < 0.8 * free
        kept = []
        rows_per = max(64, int(self.ROW_CHUNK_BYTES // max(1, v.shape[1] * v.itemsize)) // 64 * 64)
        for r0 in range(0, self.m, rows_per):
            r1 = min(self.m, r0 + rows_per)
            chunk = torch.from_numpy(v[r0:r1]).to(self.device, non_blocking=False)
            if keep:
                kept.append((r0, chunk))
            yield r0, chunk
        if keep:
            self._chunk_cache = (v, kept)

    # ---- sharded upload: every rank reads ONLY its rows of the host V; the column copy is assembled from the
    # ranks' row shards by an all-to-all over NVLink (1/P of V over each PCIe link instead of all of it)
    def _shard_and_panel(self, v: np.ndarray):
        """(shard, panel): this rank's rows V[I_p, :] and columns V[:, J_p] on the device, in the host dtype.
        The rows come from host memory as one contiguous run: 1/P of V per PCIe link.  On GPUs the columns are then
        assembled over NVLink: every rank stores, for each peer q, the columns J_q of ITS rows straight into q's
        panel (nmfb_peer_scatter into a peer-memory arena, one device barrier) -- V crosses PCIe exactly once in
        total.  Round 1 did that exchange with point-to-point NCCL operations whose channels are built lazily on
        first use (7.4 s at 8 GPUs).  Without peer memory (host stand-in of the CPU tests) the columns are read
        from host memory as a strided 2-D copy."""
        cache = getattr(self, "_chunk_cache", None)
        if cache is not None and cache[0] is v:
            return cache[1][:2]
        (r0, r1), (c0, c1) = self.row_range, self.col_range
        tdt = torch.float64 if v.dtype == np.float64 else torch.float32
        es, n = v.itemsize, self.n
        shard = torch.empty((r1 - r0, n), dtype=tdt, device=self.device)
        host = v.ctypes.data
        check(self.lib.nmfb_upload_2d(shard.data_ptr(), n * es, host + r0 * n * es, n * es, n * es, r1 - r0,
                                      self.stream))
        arena = None
        if self.peer is not None:
            from .peer import PeerArena
            per = self.per_n
            arena = PeerArena(self.lib, [("panel", self.Mp * per * es)], self.device, self.rank, self.world, self.group)
            nb = shard_bounds(self.n, self.world)[2]
            widths = [live_rows(self.n, lo, hi) * es for lo, hi in nb]
            arena.scatter(shard.data_ptr(), n * es, per * es, "panel", r0 * per * es, per * es, r1 - r0, widths,
                          self.stream)
            arena.barrier(self.stream)
            panel = arena.view("panel", (self.Mp, per), tdt)[: self.m, : c1 - c0]
        else:
            panel = torch.empty((self.m, c1 - c0), dtype=tdt, device=self.device)
            check(self.lib.nmfb_upload_2d(panel.data_ptr(), (c1 - c0) * es, host + c0 * es, n * es, (c1 - c0) * es,
                                          self.m, self.stream))
        out = (shard, panel, arena)                      # the arena owns the panel's memory
        self._last_upload = out                          # alive until the ingest kernels have run (upload_data syncs)
        if getattr(self, "_keep_chunks", False):
            self._chunk_cache = (v, out)
        return shard, panel

    def _panel_from_shards(self, shard: torch.Tensor, panel: torch.Tensor):
        """panel (m x my columns) <- every rank's rows of those columns: all-to-all of the row shards."""
        mb = shard_bounds(self.m, self.world)[2]
        nb = shard_bounds(self.n, self.world)[2]
        rows_of = [(lo, lo + live_rows(self.m, lo, hi)) for lo, hi in mb]
        cols_of = [(lo, lo + live_rows(self.n, lo, hi)) for lo, hi in nb]
        send = [shard[:, a:b].contiguous() for a, b in cols_of]                          # my rows of rank q's columns
        recv = [panel[a:b] for a, b in rows_of]                                          # row slices are contiguous
        exchange_blocks(send, recv, self.group)

    def _ingest_shard_and_panel(self, shard, panel, vr, vc, stats, rowmax, colmax, fixed: bool):
        """Row pass: my rows, all columns -> the planes of `vr` only (empty column window).  Column pass: all rows
        of my columns, presented as a matrix with that many columns -> the planes of `vc` only (empty row window).
        fixed: |max| of my rows from the shard, of my columns from the panel; the row maxima are all-reduced
        because the column pass judges exactness with every row's unit."""
        (r0, r1), (c0, c1) = self.row_range, self.col_range
        is64 = 1 if shard.dtype == torch.float64 else 0
        vc_none = Data(vc.planes, 0, vc.cols, vc.rows_p, vc.ld, vc.nv, 1, vc.fixed, 0)
        vr_none = Data(vr.planes, 0, c1 - c0, vr.rows_p, vr.ld, vr.nv, 1, vr.fixed, 0)
        vc_local = Data(vc.planes, c1 - c0, vc.cols, vc.rows_p, vc.ld, vc.nv, 1, vc.fixed, 0)
        st = stats.data_ptr()
        if fixed:
            scratch = torch.zeros(max(self.Mp, self.Np), dtype=torch.int32, device=self.device)
            check(self.lib.nmfb_data_absmax(shard.data_ptr(), is64, shard.stride(0), r0, r1 - r0, self.n,
                                            rowmax.data_ptr(), scratch.data_ptr(), self.stream))
            scratch.zero_()
            check(self.lib.nmfb_data_absmax(panel.data_ptr(), is64, panel.stride(0), 0, self.m, c1 - c0,
                                            scratch.data_ptr(), colmax.data_ptr() + 4 * c0, self.stream))
            allreduce_max_(rowmax, self.group)
            check(self.lib.nmfb_ingest_data_fixed(shard.data_ptr(), is64, shard.stride(0), r0, r1 - r0, C.byref(vr),
                                                  C.byref(vc_none), st, rowmax.data_ptr(), colmax.data_ptr(),
                                                  self.stream))
            check(self.lib.nmfb_ingest_data_fixed(panel.data_ptr(), is64, panel.stride(0), 0, self.m, C.byref(vr_none),
                                                  C.byref(vc_local), st, rowmax.data_ptr(),
                                                  colmax.data_ptr() + 4 * c0, self.stream))
        else:
            check(self.lib.nmfb_ingest_data(shard.data_ptr(), is64, shard.stride(0), r0, r1 - r0, C.byref(vr),
                                            C.byref(vc_none), st, self.stream))
            check(self.lib.nmfb_ingest_data(panel.data_ptr(), is64, panel.stride(0), 0, self.m, C.byref(vr_none),
                                            C.byref(vc_local), st, self.stream))

    def _ingest_pass_sharded(self, v: np.ndarray, nv: int, fixed: bool):
        shard, panel = self._shard_and_panel(v)
        self.begin_ingest(nv, fixed)
        vr, vc = self._data_structs()
        self._ingest_shard_and_panel(shard, panel, vr, vc, self.stats, getattr(self, "rowmax", None),
                                     getattr(self, "colmax", None), fixed)
        allreduce_min_key_(self.stats[0:1], self.group)      # each rank saw its rows and its columns only
        return self.finish_ingest()

    def _ingest_pass(self, v: np.ndarray, nv: int, fixed: bool = False):
        if self.world > 1 and os.environ.get("NMFB_FULL_WALK_UPLOAD", "0") != "1":
            return self._ingest_pass_sharded(v, nv, fixed)
        self.begin_ingest(nv, fixed)
        if fixed:
            for r0, chunk in self._host_chunks(v):
                self.absmax_chunk(chunk, r0)
        for r0, chunk in self._host_chunks(v):
            self.ingest_chunk(chunk, r0)
        return self.finish_ingest()

    def upload_data(self, v: np.ndarray, planes: Optional[int] = None, fixed: bool = False) -> float:
        """Copy the caller's V (host, fp32/fp64) to HBM as bf16 planes of V and V^T.
        Default: the single lossless plane when V is bf16-exact, hi + lo planes otherwise.
        fixed=True (ADMM / AO-ADMM): fixed-point plane + remainder for the exact-hi contractions; falls
        back to the floating planes when two planes cannot hold the data exactly.  Returns min(V)."""
        import time
        t_up = time.perf_counter()
        try:
            return self._upload_data(v, planes, fixed)
        finally:
            if torch.device(self.device).type == "cuda":
                self._sync_stream()
            self._last_upload = None
            t = getattr(self, "timings", None)
            if t is not None:
                t["upload_s"] = t.get("upload_s", 0.0) + time.perf_counter() - t_up

    def _upload_data(self, v: np.ndarray, planes: Optional[int] = None, fixed: bool = False) -> float:
        if v.ndim != 2 or v.shape != (self.m, self.n):
            raise ValueError(f"data must be a 2-D array of shape {(self.m, self.n)}")
        if v.dtype not in (np.float32, np.float64):
            v = v.astype(np.float64)
        if not v.flags.c_contiguous:
            v = np.ascontiguousarray(v)
        # device copies of the chunks are kept only where a second pass over V is certain (fixed-point ingest:
        # |max| pass, then the planes); a floating upload is one pass (see _upload_passes)
        self._chunk_cache, self._keep_chunks = None, bool(fixed)
        try:
            return self._upload_passes(v, planes, fixed)
        finally:
            self._chunk_cache, self._keep_chunks = None, False

    def _upload_passes(self, v: np.ndarray, planes: Optional[int], fixed: bool) -> float:
        if fixed:
            vmin, inexact, bad = self._ingest_pass(v, 2, fixed=True)
            if not inexact:
                if bad:
                    raise ValueError("data contains non-finite values")
                self.vmin = vmin
                return vmin
        # How many bf16 planes V needs is decided by the ingest kernel over ALL of V; a small sample read on the host
        # first lets the common cases take ONE pass over PCIe: a sample that is not bf16-exact goes straight to
        # hi + lo planes, an exact one to the single plane (and only a V that is exact in the sample but not
        # everywhere pays the second pass).
        first = planes or (2 if _sample_needs_lo_plane(v) else 1)
        vmin, need_lo, bad = self._ingest_pass(v, first)
        if need_lo and planes is None and first == 1:
            vmin, need_lo, bad = self._ingest_pass(v, 2)
        if bad:
            raise ValueError("data contains non-finite values")
        self.vmin = vmin
        return vmin

    def upload_data_device(self, v: torch.Tensor, planes: int = 1) -> float:
        """Same as upload_data for a V that already lives on this device (fp32/fp64, row-major)."""
        assert v.is_cuda and v.shape == (self.m, self.n) and v.stride(1) == 1
        self._alloc_data(planes)
        vr, vc = self._data_structs()
        self.stats.copy_(torch.tensor([-1, 0, 0, 0], dtype=torch.int32))
        check(self.lib.nmfb_ingest_data(v.data_ptr(), 1 if v.dtype == torch.float64 else 0, v.stride(0), 0,
                                        self.m, C.byref(vr), C.byref(vc), self.stats.data_ptr(), self.stream))
        st = self.stats.cpu().numpy().view(np.uint32)
        self.vmin = _key_to_float(int(st[0]))
        return self.vmin

    def set_factors(self, w0: np.ndarray, h0: np.ndarray):
        """w0: (m, k), h0: (k, n) host arrays (fp32/fp64)."""
        if w0.shape != (self.m, self.k) or h0.shape != (self.k, self.n):
            raise ValueError("initial factors have the wrong shape")
        for arr, fac, sl, sc in ((w0, self.w, None, None), (h0, self.h, None, None)):
            a = np.ascontiguousarray(arr, dtype=np.float64 if arr.dtype == np.float64 else np.float32)
            d = torch.from_numpy(a).to(self.device)
            if fac is self.w:
                stride_l, stride_c = self.k, 1
            else:
                stride_l, stride_c = 1, self.n
            fs = fac.struct()
            check(self.lib.nmfb_import_factor(d.data_ptr(), 1 if a.dtype == np.float64 else 0, stride_l,
                                              stride_c, self.k, self.KP, C.byref(fs), self.stream))
            self._sync_stream()
            del d
        self._problem = None

    def get_factors(self, dtype=np.float64, masters=None) -> Tuple[np.ndarray, np.ndarray]:
        """Download W (m, k) and H (k, n) as host arrays."""
        import time
        t_dn = time.perf_counter()
        try:
            return self._get_factors(dtype, masters)
        finally:
            self.timings["download_s"] = self.timings.get("download_s", 0.0) + time.perf_counter() - t_dn

    def _get_factors(self, dtype=np.float64, masters=None) -> Tuple[np.ndarray, np.ndarray]:
        tdt = torch.float64 if dtype == np.float64 else torch.float32
        w_d = torch.empty((self.m, self.k), dtype=tdt, device=self.device)
        h_d = torch.empty((self.k, self.n), dtype=tdt, device=self.device)
        wf, hf = self.w.struct(), self.h.struct()
        if masters is not None:
            wf.master, hf.master = masters[0].data_ptr(), masters[1].data_ptr()
        is64 = 1 if tdt == torch.float64 else 0
        check(self.lib.nmfb_export_factor(C.byref(wf), self.k, self.KP, w_d.data_ptr(), is64, self.k, 1, self.stream))
        check(self.lib.nmfb_export_factor(C.byref(hf), self.k, self.KP, h_d.data_ptr(), is64, 1, self.n, self.stream))
        return w_d.cpu().numpy(), h_d.cpu().numpy()

    # ------------------------------------------------------------------ steps
    def objective_eu(self, slot: int = 0, impl: int = 0):
        p = self.problem()
        check(self.lib.nmfb_objective_eu(C.byref(p), impl, self.scalars.data_ptr() + 8 * slot, self.stream))

    def mur_eu_step(self, lambda_w: float, lambda_h: float, obj_slot: Optional[int] = 0):
        p = self.problem()
        out = None if obj_slot is None else self.scalars.data_ptr() + 8 * obj_slot
        check(self.lib.nmfb_mur_eu_step(C.byref(p), float(lambda_w), float(lambda_h), out, self.stream))

    # ------------------------------------------------------------------ ADMM / AO-ADMM state
    def _fixed_planes(self, fac: "FactorBuffers"):
        fs = fac.struct()
        check(self.lib.nmfb_fixed_planes(self.KP, C.byref(fs), self.stream))

    def alloc_admm(self, with_aux: bool, exact: Optional[bool] = None):
        """Duals (zero, admm.py:31-32 / ao_admm.py:26-27), the k x k inverse workspace and, for ADMM,
        the signed auxiliaries initialised to copies of W and H (admm.py:27-28).  exact: contract in the
        exact-hi arithmetic (default: whenever V was ingested as fixed-point planes)."""
        dev, KP = self.device, self.KP
        exact = self.fixed if exact is None else bool(exact)
        Lmax = max(self.Mp, self.Np)
        self.dual_w = torch.zeros((KP, self.Mp), dtype=torch.float32, device=dev)
        self.dual_h = torch.zeros((KP, self.Np), dtype=torch.float32, device=dev)
        self.ainv = torch.zeros((2, KP, KP), dtype=torch.float64, device=dev)
        self.chol_scratch = torch.zeros((KP, KP + 1), dtype=torch.float64, device=dev)
        self.c64 = torch.zeros((KP, Lmax), dtype=torch.float64, device=dev)
        self.admm_scal = torch.zeros(4, dtype=torch.float64, device=dev)
        self.admm_flags = torch.zeros(8, dtype=torch.int32, device=dev)
        self.norm_partial = torch.zeros(4 * (Lmax // 64), dtype=torch.float64, device=dev)
        self.binv = torch.zeros((2, KP, KP), dtype=torch.float64, device=dev)          # prox l2n systems
        self.tik_scratch = torch.zeros(KP * KP + 4, dtype=torch.float64, device=dev)
        if with_aux:
            self.w_aux = FactorBuffers(self.m, KP, dev, self.Mp, self.w.win0, self.w.win_p)
            self.h_aux = FactorBuffers(self.n, KP, dev, self.Np, self.h.win0, self.h.win_p)
            if self.peer is not None:               # exchange buffers of the auxiliaries: a second arena
                from .peer import PeerArena
                self.peer_aux = PeerArena(self.lib, [("xwa", 2 * KP * self.Mp * 4), ("xha", 2 * KP * self.Np * 4)],
                                          self.device, self.rank, self.world, self.group)
                self._attach_exchange(self.w_aux, self.peer_aux, "xwa")
                self._attach_exchange(self.h_aux, self.peer_aux, "xha")
            for src, dst in ((self.w, self.w_aux), (self.h, self.h_aux)):
                dst.master.copy_(src.master)
                fs = dst.struct()
                check(self.lib.nmfb_refresh_planes(KP, C.byref(fs), self.stream))
                if exact:
                    self._fixed_planes(dst)
            wa, ha = self.w_aux.struct(), self.h_aux.struct()
        else:
            self.w_aux = self.h_aux = None
            wa, ha = Factor(), Factor()
            if exact:                           # AO-ADMM contracts against W and H themselves
                self._fixed_planes(self.w)
                self._fixed_planes(self.h)
        side = self._side_stream_handles(Lmax)
        self._admm = AdmmState(wa, ha, self.dual_w.data_ptr(), self.dual_h.data_ptr(), self.ainv.data_ptr(),
                               self.chol_scratch.data_ptr(), self.c64.data_ptr(), self.admm_scal.data_ptr(),
                               self.admm_flags.data_ptr(), self.norm_partial.data_ptr(), int(exact),
                               self.binv.data_ptr(), self.tik_scratch.data_ptr(), *side)

    def _side_stream_handles(self, Lmax: int):
        """(side stream, fork event, join event, Gram scratch, its bytes) of nmfb_admm_state: the k x k work of a
        half step (Gram, single-CTA Cholesky inverse, Tikhonov system) runs next to the numerator GEMM.  The host
        stand-in of the CPU tests has no streams: all NULL, i.e. everything on the one stream."""
        if torch.device(self.device).type != "cuda":
            return (None, None, None, None, 0)
        nbytes = int(self.lib.nmfb_gram_partial_bytes(self.k, Lmax, self.num_sms))
        self.gram_partial = torch.empty(max(nbytes, 4) // 4, dtype=torch.float32, device=self.device)
        self.side_stream = torch.cuda.Stream(device=self.device)
        self._side_events = (torch.cuda.Event(), torch.cuda.Event())
        for ev in self._side_events:
            ev.record(torch.cuda.current_stream(self.device))         # materialises the cudaEvent_t
        return (self.side_stream.cuda_stream, int(self._side_events[0].cuda_event),
                int(self._side_events[1].cuda_event), self.gram_partial.data_ptr(), self.gram_partial.numel() * 4)

    # ------------------------------------------------------------------ ADMM, Kullback-Leibler branch
    def alloc_admm_kl(self, ao: bool = False):
        """admm.py:33,303-315 / ao_admm.py:28-30,71-101: dual_v (zero) and, instead of v_aux itself, the data
        term D = v_aux + dual_v of the right-hand sides (zero at the start), kept as fixed-point bf16 planes
        in the layouts of V and V^T so that the contractions of the Euclidean branch run on it unchanged.
        Sharded runs hold the rank's rows of dual_v / D (the v step is element-wise) and re-build the column
        copy of D by the all-to-all of the sharded upload."""
        dev = self.device
        self.alloc_admm(with_aux=True, exact=True)
        if ao:                                   # AO-ADMM contracts against W and H themselves (ao_admm.py:83)
            self._fixed_planes(self.w)
            self._fixed_planes(self.h)
        # dual_v and D for this rank's rows of V, padded to whole 128 x 128 tiles (per_m rows, pitch Np) so that the
        # fused data step (kl_data.cu) needs no edge cases; the padding stays zero
        # (K-blocked like V: [Np / 64][per_m][64], so that a tile is contiguous runs of HBM)
        self.kl_dual_v = torch.zeros((self.Np // 64, self.per_m, 64), dtype=torch.float32, device=dev)
        self.kl_d = torch.zeros((self.Np // 64, self.per_m, 64), dtype=torch.float32, device=dev)
        # 24-bit row planes of the two operands of P = left . right (nmfb_row_planes3)
        self.kl_pl_a = torch.zeros((3, self.Mp, self.KP), dtype=torch.bfloat16, device=dev)
        self.kl_pl_b = torch.zeros((3, self.Np, self.KP), dtype=torch.bfloat16, device=dev)
        self.kl_scr_r = torch.zeros(self.Mp, dtype=torch.int32, device=dev)
        self.kl_scr_c = torch.zeros(self.Np, dtype=torch.int32, device=dev)
        self.kl_panel = (torch.empty((self.m, self.col_range[1] - self.col_range[0]), dtype=torch.float32, device=dev)
                         if self.world > 1 else None)
        # three planes (fixed-point slice + two bf16 remainders = 23 bits): D is an arbitrary fp32 matrix, and two
        # planes (15 bits) cost 1.2e-4 / 5e-4 of the objective at k = 64 / 128 (profiles/r02c)
        self.d_rows = torch.zeros((3, self.Np // 64, self.per_m, 64), dtype=torch.bfloat16, device=dev)
        self.d_cols = torch.zeros((3, self.Mp // 64, self.per_n, 64), dtype=torch.bfloat16, device=dev)
        self.d_rowmax = torch.zeros(self.Mp, dtype=torch.int32, device=dev)
        self.d_colmax = torch.zeros(self.Np, dtype=torch.int32, device=dev)
        self.d_stats = torch.zeros(4, dtype=torch.int32, device=dev)
        vr, vc = self._data_structs()
        self._d_structs = (Data(self.d_rows.data_ptr(), vr.rows, vr.cols, vr.rows_p, vr.ld, 3, 1, 1, vr.origin),
                           Data(self.d_cols.data_ptr(), vc.rows, vc.cols, vc.rows_p, vc.ld, 3, 1, 1, vc.origin))
        self._problem_d = Problem.from_buffer_copy(self.problem())
        self._problem_d.v_rows, self._problem_d.v_cols = self._d_structs
        self._problem_d.phase_events, self._problem_d.n_phase_events = None, 0

    def kl_data_step(self, left: "FactorBuffers", right: "FactorBuffers", done_flag: Optional[int] = None):
        """admm.py:309-313 / ao_admm.py:88-94 on this rank's rows: ONE tcgen05 kernel forms P = left^T[rows] . right
        tile by tile in tensor memory (never in HBM), does the v_aux / dual_v step and writes D = v_aux + dual_v with
        its row / column |max| (nmfb_kl_data_step); then D -> fixed-point planes of both layouts (sharded: the column
        copy through the all-to-all).  done_flag: device address of the AO-ADMM break flag; once it is set D, dual_v
        and therefore the planes stay what they are."""
        KP, n = self.KP, self.n
        r0, r1 = self.row_range
        ls, rs = left.struct(), right.struct()
        check(self.lib.nmfb_row_planes3(KP, C.byref(ls), self.kl_pl_a.data_ptr(), self.stream))
        check(self.lib.nmfb_row_planes3(KP, C.byref(rs), self.kl_pl_b.data_ptr(), self.stream))
        vr, _vc = self._data_structs()
        d = self.kl_d
        sharded = self.world > 1
        rowmax = self.kl_scr_r if sharded else self.d_rowmax      # sharded: the ingest below takes its own maxima
        colmax = self.kl_scr_c if sharded else self.d_colmax
        check(self.lib.nmfb_kl_data_step(C.byref(vr), self.kl_pl_a.data_ptr(), self.Mp, r0 if sharded else 0,
                                         self.kl_pl_b.data_ptr(), self.Np, KP, self.kl_dual_v.data_ptr(), d.data_ptr(),
                                         self.Np, rowmax.data_ptr(), colmax.data_ptr(), done_flag, self.num_sms,
                                         self.stream))
        dr, dc = self._d_structs
        if sharded:
            self.d_rowmax.zero_()
            self.d_colmax.zero_()
            live = d.permute(1, 0, 2).reshape(self.per_m, self.Np)[: r1 - r0, :n].contiguous()     # row-major rows of D
            self._panel_from_shards(live, self.kl_panel)
            self._ingest_shard_and_panel(live, self.kl_panel, dr, dc, self.d_stats, self.d_rowmax, self.d_colmax, True)
            return
        check(self.lib.nmfb_ingest_data_fixed(d.data_ptr(), 0, -self.per_m, 0, self.m, C.byref(dr), C.byref(dc),
                                              self.d_stats.data_ptr(), self.d_rowmax.data_ptr(),
                                              self.d_colmax.data_ptr(), self.stream))

    def _publish_aux(self, fac: "FactorBuffers"):
        """Exact-mode operand of the next contraction: all-gather the windows (sharded), fixed-point cplanes."""
        if self.world == 1:
            self._fixed_planes(fac)
        else:
            self.gather_factor(fac, refresh=False, fixed=True)

    def admm_kl_step(self, rho, lambda_w, prox_w, lambda_h, prox_h, ring_slot: int):
        """One outer iteration of admm.py:303-324 with distance_type='kl' (sharded: the window half-steps and
        exchanges of the Euclidean branch, solvers.AdmmStepper, on the data term)."""
        pd, s = self._problem_d, self._admm
        check(self.lib.nmfb_admm_half_step(C.byref(pd), C.byref(s), 0, float(rho), float(lambda_h), int(prox_h),
                                           self.stream))                          # admm.py:304,308,320
        self._publish_aux(self.h_aux)
        self.gather_factor(self.h)                                                # operand of the objective
        check(self.lib.nmfb_admm_half_step(C.byref(pd), C.byref(s), 1, float(rho), float(lambda_w), int(prox_w),
                                           self.stream))                          # admm.py:305-306,309,321
        self._publish_aux(self.w_aux)
        self.gather_factor(self.w, refresh=False)                                 # keeps the replicated W current
        self.kl_data_step(self.w_aux, self.h_aux)                                 # admm.py:310-313
        self.objective_kl(slot=ring_slot * 8)                                     # admm.py:324
        self.reduce_scalar(ring_slot * 8)
        self.admm_publish(ring_slot)

    def aoadmm_kl_side(self, side: int, admm_iter: int, lam, prox):
        """ao_admm.py:71-101 (admm_kl_update) for H (side 0: a = W) or W (side 1: a = H^T, everything
        transposed, ao_admm.py:276-284): Gram, rho = tr/k and the Cholesky inverse once; per inner iteration
        the right-hand side a^T (v_aux + dual_v) is contracted anew, then solve / prox / dual / norms, the data
        step with P = a.aux and the break test (whose flag freezes the remaining iterations on the device).
        Sharded: the rank's window of the factor; the norm partials (and, for the H side, the auxiliary that
        P = W.h_aux needs in full) are all-gathered per inner iteration, like Engine.aoadmm_side."""
        pd, s, KP = self._problem_d, self._admm, self.KP
        fac, other = (self.h, self.w) if side == 0 else (self.w, self.h)
        data = self._d_structs[1] if side == 0 else self._d_structs[0]
        slots = check(self.lib.nmfb_aoadmm_side_begin(C.byref(pd), C.byref(s), side, self.stream))
        if int(prox) == 2 and float(lam) != 0.0:
            check(self.lib.nmfb_admm_prepare_l2n(C.byref(pd), C.byref(s), side, float(lam), self.stream))
        flag = self.admm_flags.data_ptr() + 4                                     # flags[1]: inner loop has broken
        for it in range(admm_iter):
            if it > 0:                                                            # ao_admm.py:83: fresh right-hand side
                slots = check(self.lib.nmfb_panel_gemm_exact(C.byref(data), other.cplanes.data_ptr(), other.Lp, KP,
                                                             self.partial.data_ptr() + 4 * fac.win0,
                                                             self.partial.numel() * 4, fac.Lp, self.num_sms,
                                                             self.stream))
            check(self.lib.nmfb_aoadmm_inner(C.byref(pd), C.byref(s), side, 1, slots, float(lam), int(prox),
                                             self.stream))                        # ao_admm.py:83-85,92
            if self.world > 1:
                self.gather_norm_partials(fac)
                if side == 0:
                    self.gather_factor(self.h_aux, refresh=False)
            if side == 0:
                self.kl_data_step(self.w, self.h_aux, flag)                       # ao_admm.py:88-90,94
            else:
                self.kl_data_step(self.w_aux, self.h, flag)
            check(self.lib.nmfb_aoadmm_inner_flag(C.byref(pd), C.byref(s), side, self.stream))   # ao_admm.py:96
        if self.world == 1:
            self._fixed_planes(fac)
        else:
            self.gather_factor(fac, refresh=True, fixed=True)

    def aoadmm_kl_step(self, admm_iter, lambda_w, prox_w, lambda_h, prox_h, ring_slot: int):
        """One outer iteration of ao_admm.py:272-291 with distance_type='kl'."""
        self.aoadmm_kl_side(0, admm_iter, lambda_h, prox_h)
        self.aoadmm_kl_side(1, admm_iter, lambda_w, prox_w)
        self.objective_kl(slot=ring_slot * 8)
        self.reduce_scalar(ring_slot * 8)
        self.admm_publish(ring_slot)

    def admm_step(self, rho, lambda_w, prox_w, lambda_h, prox_h, ring_slot: int):
        p = self.problem()
        check(self.lib.nmfb_admm_step(C.byref(p), C.byref(self._admm), float(rho), float(lambda_w), int(prox_w),
                                      float(lambda_h), int(prox_h), self.scalars.data_ptr() + 64 * ring_slot,
                                      self.stream))

    def aoadmm_step(self, admm_iter, lambda_w, prox_w, lambda_h, prox_h, ring_slot: int):
        p = self.problem()
        check(self.lib.nmfb_aoadmm_step(C.byref(p), C.byref(self._admm), int(admm_iter), float(lambda_w),
                                        int(prox_w), float(lambda_h), int(prox_h),
                                        self.scalars.data_ptr() + 64 * ring_slot, self.stream))

    def objective_kl(self, slot: int = 0):
        p = self.problem()
        check(self.lib.nmfb_objective_kl(C.byref(p), self.scalars.data_ptr() + 8 * slot, self.stream))

    def mur_kl_step(self, lambda_w: float, lambda_h: float, obj_slot: Optional[int] = 0):
        p = self.problem()
        out = None if obj_slot is None else self.scalars.data_ptr() + 8 * obj_slot
        check(self.lib.nmfb_mur_kl_step(C.byref(p), float(lambda_w), float(lambda_h), out, self.stream))

    def read_scalar(self, slot: int = 0) -> float:
        return float(self.scalars[slot].item())

    def launch_count(self) -> int:
        return int(self.lib.nmfb_launch_count())

    def release(self):
        """Drop every device buffer (the object is unusable afterwards)."""
        self.v_rows = self.v_cols = None
        self.w = self.h = None
        self.partial = self.obj_partial = self.gram32 = self.gram64 = None
        self.peer = self.peer_aux = None
        for name in ('dual_w', 'dual_h', 'ainv', 'c64', 'w_aux', 'h_aux', 'norm_partial', 'chol_scratch', 'binv', 'tik_scratch',
                     'kl_dual_v', 'kl_d', 'kl_panel', 'd_rows', 'd_cols', 'kl_pl_a', 'kl_pl_b'):
            setattr(self, name, None)
        self._problem = None
