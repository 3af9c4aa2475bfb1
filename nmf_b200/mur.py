"""Multiplicative-update NMF (Lee & Seung) on B200 -- drop-in for the reference's `mur.mur`
(nmf/mur.py:52-146): same keyword arguments, defaults, console output and `Results` layout.

The loop body of the reference (mur.py:119-136) becomes one C-ABI call per outer iteration
(`nmfb_mur_eu_step` / `nmfb_mur_kl_step`); W.H is never formed in HBM.
"""
from __future__ import annotations

import logging
from collections import namedtuple

import numpy as np

from . import _loop
from .engine import Engine
from .solvers import make_stepper
from .utils import make_checkpointer, tol_precision

Experiment = namedtuple('Experiment', 'method components distance_type nndsvd_init max_iter tol1 tol2 lambda_w lambda_h')
Results = namedtuple('Results', 'w h i obj_history experiment')


def mur(x, k, *, distance_type='kl', min_iter=100, max_iter=100000, tol1=1e-5, tol2=1e-5,
        lambda_w=0.0, lambda_h=0.0, nndsvd_init=(False, 'zero'), save_dir='./results/',
        w0=None, h0=None, verbose=True, engine=None, comm=None,
        checkpoint_every=0, checkpoint_path=None):
    """x: (m, n) non-negative data, k: number of components.  Extra (non-reference) keywords:
    w0 / h0 = explicit initial factors, verbose = per-iteration print, engine = reuse device state."""
    experiment = Experiment(method='mur', components=k, distance_type=distance_type,
                            nndsvd_init=nndsvd_init, max_iter=max_iter, tol1=tol1, tol2=tol2,
                            lambda_w=lambda_w, lambda_h=lambda_h)
    if distance_type not in ('eu', 'kl'):
        raise KeyError('Unknown distance type.')
    precision = tol_precision(tol1, tol2)
    x = np.asarray(x) if not isinstance(x, np.ndarray) else x
    m, n = x.shape

    if comm is not None and engine is None:              # sharded over the ranks of a torch.distributed group
        engine = Engine.for_comm(m, n, k, comm)
        verbose = verbose and engine.rank == 0
    eng = engine if engine is not None else Engine(m, n, k)
    if engine is None or eng.v_rows is None:
        vmin = eng.upload_data(x)
        if vmin < 0:                                   # mur.py:99-102: shift the caller's array in place
            amount = abs(np.min(x))
            x += amount
            logging.info('Data elevated by {}.'.format(amount))
            eng.upload_data(x)

    if w0 is None or h0 is None:
        from .init import initial_factors
        w0, h0 = initial_factors(x, k, nndsvd_init, engine=eng)
    eng.set_factors(np.asarray(w0), np.asarray(h0))

    stepper = make_stepper(eng, 'mur', distance_type, lambda_w=lambda_w, lambda_h=lambda_h)
    stepper.prepare()                                  # objective of the initial pair -> scalars[0]
    obj0 = eng.read_scalar(0)
    step = stepper.step

    logging.info('Entering Main Loop.')
    lr = _loop.run(eng, step, obj0=obj0, min_iter=min_iter, max_iter=max_iter, tol1=tol1, tol2=tol2,
                   precision=precision, verbose=verbose, lagged=getattr(stepper, 'lagged', False),
                   finish=getattr(stepper, 'finish', None), checkpoint_every=checkpoint_every,
                   on_checkpoint=make_checkpointer(eng, experiment, checkpoint_every, checkpoint_path))
    w, h = eng.get_factors(masters=lr.masters)
    return Results(w=w, h=h, i=lr.i, obj_history=lr.obj_history, experiment=experiment)
