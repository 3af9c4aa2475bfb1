"""ADMM for NMF on B200 -- drop-in for the reference's `admm.admm` (nmf/admm.py:233-345): same keyword
arguments, defaults, console output and `Results` / `Experiment` layout.

Euclidean branch: one C-ABI call per outer iteration (`nmfb_admm_step`): Gram -> fp64 Cholesky inverse ->
split-bf16 numerator GEMM -> fused solve / prox / dual kernel, for H then W, then the residual objective.
Kullback-Leibler branch (admm.py:303-315): the same half-steps on the data term v_aux + dual_v,
followed by the element-wise v_aux / dual_v kernel (`Engine.admm_kl_step`).
"""
from __future__ import annotations

import logging
from collections import namedtuple

import numpy as np

from . import _loop
from .engine import Engine
from .solvers import make_stepper, prox_code
from .utils import make_checkpointer, tol_precision

Experiment = namedtuple('Experiment', 'method components rho distance_type nndsvd_init min_iter max_iter tol1 tol2 lambda_w prox_w lambda_h prox_h')
Results = namedtuple('Results', 'w h i obj_history experiment')


def _raise_if_not_pd(i, vals):
    if vals[3] != 0:
        raise np.linalg.LinAlgError('Matrix is not positive definite (Gram + rho*I), outer iteration {}.'.format(i))


def admm(v, k, *, rho=1, distance_type='eu', reg_w=(0, 'nn'), reg_h=(0, 'l2n'), min_iter=10,
         max_iter=100000, tol1=1e-3, tol2=1e-3, nndsvd_init=(True, 'zero'), save_dir='./results/',
         w0=None, h0=None, verbose=True, engine=None, comm=None,
      checkpoint_every=0, checkpoint_path=None):
    """v: (m, n) data, k: components.  Extra (non-reference) keywords: w0 / h0, verbose, engine, and
    comm (True or a torch.distributed group: 1-D sharded run, one process per GPU)."""
    experiment = Experiment(method='admm', components=k, rho=rho, distance_type=distance_type,
                            nndsvd_init=nndsvd_init, min_iter=min_iter, max_iter=max_iter, tol1=tol1,
                            tol2=tol2, lambda_w=reg_w[0], prox_w=reg_w[1], lambda_h=reg_h[0],
                            prox_h=reg_h[1])
    if distance_type not in ('eu', 'kl'):
        raise TypeError('Unknown loss type.')
    prox_code(reg_w, k), prox_code(reg_h, k)                 # TypeError('Unknown prox_type.') up front
    precision = tol_precision(tol1, tol2)
    v = np.asarray(v)
    m, n = v.shape

    if comm is not None and engine is None:
        engine = Engine.for_comm(m, n, k, comm)
        verbose = verbose and engine.rank == 0
    eng = engine if engine is not None else Engine(m, n, k)
    if engine is None or eng.v_rows is None:
        eng.upload_data(v, fixed=(distance_type == 'eu'))     # kl: V is read by the objective and the v_aux step only
    if w0 is None or h0 is None:
        from .init import initial_factors
        w0, h0 = initial_factors(v, k, nndsvd_init, engine=eng)
    eng.set_factors(np.asarray(w0), np.asarray(h0))

    stepper = make_stepper(eng, 'admm', distance_type, rho=rho, reg_w=reg_w, reg_h=reg_h)
    stepper.prepare()
    obj0 = eng.read_scalar(0)
    lr = _loop.run(eng, stepper.step, obj0=obj0, min_iter=min_iter, max_iter=max_iter, tol1=tol1, tol2=tol2,
                   precision=precision, verbose=verbose, on_extra=_raise_if_not_pd,
                   checkpoint_every=checkpoint_every,
                   on_checkpoint=make_checkpointer(eng, experiment, checkpoint_every, checkpoint_path))
    w, h = eng.get_factors(masters=lr.masters)
    return Results(w=w, h=h, i=lr.i, obj_history=lr.obj_history, experiment=experiment)
