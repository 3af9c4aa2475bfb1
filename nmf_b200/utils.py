"""Host-side helpers shared by the solver drivers: stop test, console format, result saving.

Mirrors the role of the reference's nmf/utils.py (convergence_check :4-15, save_results :96-105);
the objective itself (utils.distance :18-33) is computed on the GPU by libnmf_b200.so.
"""
from __future__ import annotations

import numpy as np


def convergence_check(new, old, tol1, tol2):
    """Absolute-tolerance stop test of the reference (utils.py:4-15), same console messages."""
    if new < tol1:
        print('Algorithm converged (1).')
        return True
    if new >= old - tol2:
        print('Algorithm converged (2).')
        return True
    return False


def tol_precision(tol1, tol2):
    """Decimal places shown per iteration (mur.py:93-95).  The reference raises IndexError for a
    zero tolerance; here that case simply prints 2 decimals."""
    tol = min(tol1, tol2)
    if tol < 1:
        parts = format(tol, 'e').split('-')
        return int(parts[1]) if len(parts) > 1 else 2
    return 2


def save_results(save_str, w, h, i, obj_history, experiment):
    """np.savez with the reference's key set (utils.py:103): w, h, i, obj_history, experiment."""
    np.savez(save_str, w=w, h=h, i=i, obj_history=obj_history, experiment=experiment)
    print('Results saved in {}.'.format(save_str))


def make_checkpointer(engine, experiment, checkpoint_every, checkpoint_path):
    """on_checkpoint callback for _loop.run: downloads the snapshotted factors of iteration i and writes
    them with the reference's key set (utils.py:103) to `checkpoint_path` (np.savez adds .npz).  A run is
    resumed by passing the saved w / h as w0 / h0 (exact for `mur`, whose whole state they are; `admm` /
    `ao_admm` restart their duals from zero, like a fresh reference run from that point)."""
    if not checkpoint_every or checkpoint_path is None:
        return None

    def on_checkpoint(i, masters, obj_history):
        if getattr(engine, "rank", 0) != 0:
            return
        w, h = engine.get_factors(masters=masters)
        np.savez(checkpoint_path, w=w, h=h, i=i, obj_history=obj_history, experiment=experiment._asdict())

    return on_checkpoint
