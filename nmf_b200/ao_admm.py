"""AO-ADMM for NMF on B200 -- drop-in for the reference's `ao_admm.ao_admm` (nmf/ao_admm.py:201-311): same
keyword arguments, defaults, console output and `Results` / `Experiment` layout.

Euclidean branch: one C-ABI call per outer iteration (`nmfb_aoadmm_step`).  The inner ADMM loop of each
factor (ao_admm.py:58-66) runs entirely on the device, early break included; the host only learns how many
inner iterations ran (for the reference's 'ADMM break after N iterations.' line).
Kullback-Leibler branch (ao_admm.py:71-101,272-285): `Engine.aoadmm_kl_side` enqueues the inner
iterations -- each with a fresh right-hand side from v_aux + dual_v and the element-wise data step --, the
early break is still taken on the device (every kernel of a broken loop returns at its first instruction).
"""
from __future__ import annotations

from collections import namedtuple

import numpy as np

from . import _loop
from .admm import _raise_if_not_pd
from .engine import Engine
from .solvers import make_stepper, prox_code
from .utils import make_checkpointer, tol_precision

Experiment = namedtuple('Experiment', 'method components distance_type nndsvd_init min_iter max_iter admm_iter tol1 tol2 lambda_w prox_w lambda_h prox_h')
Results = namedtuple('Results', 'w h i obj_history experiment')


def ao_admm(v, k, *, distance_type='eu', reg_w=(0, 'nn'), reg_h=(0, 'l2n'), min_iter=10, max_iter=100000,
            admm_iter=10, tol1=1e-3, tol2=1e-3, nndsvd_init=(True, 'zero'), save_dir='./results/',
            w0=None, h0=None, verbose=True, engine=None, comm=None,
         checkpoint_every=0, checkpoint_path=None):
    """v: (m, n) data, k: components.  Extra (non-reference) keywords: w0 / h0, verbose, engine, and
    comm (True or a torch.distributed group: 1-D sharded run, one process per GPU)."""
    experiment = Experiment(method='ao_admm', components=k, distance_type=distance_type,
                            nndsvd_init=nndsvd_init, min_iter=min_iter, max_iter=max_iter,
                            admm_iter=admm_iter, tol1=tol1, tol2=tol2, lambda_w=reg_w[0], prox_w=reg_w[1],
                            lambda_h=reg_h[0], prox_h=reg_h[1])
    if distance_type not in ('eu', 'kl'):
        raise TypeError('Unknown loss type.')
    prox_code(reg_w, k), prox_code(reg_h, k)
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

    stepper = make_stepper(eng, 'ao_admm', distance_type, reg_w=reg_w, reg_h=reg_h, admm_iter=admm_iter)
    stepper.prepare()
    obj0 = eng.read_scalar(0)
    inner_used = []

    def on_extra(i, vals):
        _raise_if_not_pd(i, vals)
        used = (int(vals[1]), int(vals[2]))
        inner_used.append(used)
        if verbose:
            for u in used:                                   # ao_admm.py:65, H update first, then W
                if u < admm_iter:
                    print('ADMM break after {} iterations.'.format(u))

    lr = _loop.run(eng, stepper.step, obj0=obj0, min_iter=min_iter, max_iter=max_iter, tol1=tol1, tol2=tol2,
                   precision=precision, verbose=verbose, on_extra=on_extra,
                   checkpoint_every=checkpoint_every,
                   on_checkpoint=make_checkpointer(eng, experiment, checkpoint_every, checkpoint_path))
    w, h = eng.get_factors(masters=lr.masters)
    res = Results(w=w, h=h, i=lr.i, obj_history=lr.obj_history, experiment=experiment)
    ao_admm.last_inner_iterations = inner_used[: lr.i + 1]        # diagnostic for the parity tests
    return res
