"""Outer-iteration driver shared by mur / admm / ao_admm.

The GPU runs one iteration ahead of the host: iteration i+1 is enqueued before the objective of
iteration i has been read back, so the per-iteration print + stop test of the reference
(mur.py:127-136, admm.py:324-334, ao_admm.py:291-301) never stalls the device.  When a stop test
can fire, the factor masters of iteration i are snapshotted on the stream (device-to-device) before
iteration i+1 overwrites them, so the returned factors are exactly those of the stopping iteration.
"""
from __future__ import annotations

import logging
from collections import deque

import torch

from .utils import convergence_check

RING = 8          # scalar mailbox ring (slots of 8 doubles each)
LAG = 1


class _NoEvent:
    """Stand-in for torch.cuda.Event when the engine lives on the CPU (host-logic tests only: the
    product engine refuses to exist without a CUDA device)."""

    def record(self, stream=None):
        pass

    def synchronize(self):
        pass


class LoopResult:
    __slots__ = ("i", "obj_history", "masters", "extras")

    def __init__(self):
        self.i = -1
        self.obj_history = []
        self.masters = None
        self.extras = []


def run(engine, enqueue_step, *, obj0, min_iter, max_iter, tol1, tol2, precision, verbose=True,
        on_extra=None, lagged=False, finish=None, checkpoint_every=0, on_checkpoint=None):
    """enqueue_step(slot) must enqueue one outer iteration that leaves the objective in
    engine.scalars[slot*8] (and optional extras in [slot*8+1 : slot*8+8]).

    lagged=True (MUR-KL): the objective of iteration i-1 is a by-product of the first kernel of iteration
    i.  enqueue_step(slot, prev_slot, mid) then writes it to scalars[prev_slot*8] and calls mid() right
    after that kernel; the host reads it while the rest of iteration i runs, and finish(slot) enqueues a
    stand-alone objective for the very last iteration.

    checkpoint_every=N > 0: the reference's commented hook (`if i % 100 == 0: save_results(...)`,
    mur.py:138-140): the factors of every N-th iteration are snapshotted on the stream like those of a
    possible stop, and on_checkpoint(i, (w_master, h_master), obj_history) is called once the objective of
    that iteration has been read -- the device never waits for the file."""
    res = LoopResult()
    res.obj_history = [obj0]
    on_gpu = torch.device(engine.device).type == "cuda"
    host = torch.zeros(RING * 8, dtype=torch.float64)
    if on_gpu:
        host = host.pin_memory()
    snaps = [None, None]
    pending = deque()
    stopped = False

    def process(item):
        nonlocal stopped
        i, slot, ev = item
        ev.synchronize()
        vals = host[slot * 8: slot * 8 + 8].clone()
        obj = float(vals[0])
        res.obj_history.append(obj)
        res.i = i
        if on_extra is not None:
            on_extra(i, vals)
        if verbose:
            print('[{}]: {:.{}f}'.format(i, obj, precision))
        if i > min_iter and convergence_check(res.obj_history[-1], res.obj_history[-2], tol1, tol2):
            logging.warning('Converged.')
            res.masters = snaps[i % 2]
            stopped = True
        if checkpoint_every and on_checkpoint is not None and i % checkpoint_every == 0 and not stopped:
            on_checkpoint(i, snaps[i % 2], list(res.obj_history))

    def publish(i_obj, slot):
        host[slot * 8: slot * 8 + 8].copy_(engine.scalars[slot * 8: slot * 8 + 8], non_blocking=True)
        ev = torch.cuda.Event() if on_gpu else _NoEvent()
        ev.record(torch.cuda.current_stream(engine.device) if on_gpu else None)
        pending.append((i_obj, slot, ev))

    for i in range(max_iter):
        slot = i % RING
        if lagged:
            prev = (i - 1) % RING if i > 0 else RING - 1          # i == 0: recomputes obj0 into a spare slot
            enqueue_step(slot, prev, (lambda: publish(i - 1, prev)) if i > 0 else None)
        else:
            enqueue_step(slot)
            publish(i, slot)
        if i > min_iter or (checkpoint_every and i % checkpoint_every == 0):
            if snaps[i % 2] is None:
                snaps[i % 2] = (torch.empty_like(engine.w.master), torch.empty_like(engine.h.master))
            snaps[i % 2][0].copy_(engine.w.master, non_blocking=True)
            snaps[i % 2][1].copy_(engine.h.master, non_blocking=True)
        while len(pending) > (0 if lagged else LAG) and not stopped:
            process(pending.popleft())
        if stopped:
            break
    if lagged and not stopped and max_iter > 0:
        last = (max_iter - 1) % RING
        finish(last)
        publish(max_iter - 1, last)
    while pending and not stopped:
        process(pending.popleft())
    if not stopped:
        logging.info('Max iteration reached.')
    if on_gpu:
        torch.cuda.current_stream(engine.device).synchronize()
    return res
