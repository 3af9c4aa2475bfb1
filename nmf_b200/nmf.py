"""`NMF` facade -- same constructor, `factorize`, `save_factorization`, `results` and console
output as the reference's nmf/nmf.py:7-135, plus the `w` / `h` attributes its README promises
(README.md:22) but never sets.  Solvers run on the GPU through libnmf_b200.so."""
from __future__ import annotations

import os
from importlib import import_module

from . import utils

METHODS = ('mur', 'anls', 'admm', 'ao_admm')


class NMF:
    """Non-negative matrix factorisation of a 2-D array `data` into `factors` components.

        nmf = NMF(data, factors)
        nmf.factorize(method='mur', **method_params)
        nmf.w, nmf.h, nmf.results
    """

    def __init__(self, data=None, factors=None, saving=True, param_file=None):
        self.data = data
        self.factors = factors
        self.saving = saving
        self.results = None
        if param_file is not None:                    # vestigial in the reference too (nmf.py:38-45)
            try:
                parameters = import_module(param_file)
            except ImportError:
                print('No parameter file found.')
                return
            self.method_params = parameters.method_params

    @property
    def w(self):
        return None if self.results is None else self.results.w

    @property
    def h(self):
        return None if self.results is None else self.results.h

    def factorize(self, method='mur', saving=False, **method_params):
        """Run one solver: 'mur', 'admm' or 'ao_admm' ('anls' is not part of this build)."""
        if method == 'mur':
            from .mur import mur
            self.results = mur(self.data, self.factors, **method_params)
        elif method == 'admm':
            from .admm import admm
            self.results = admm(self.data, self.factors, **method_params)
        elif method == 'ao_admm':
            from .ao_admm import ao_admm
            self.results = ao_admm(self.data, self.factors, **method_params)
        elif method == 'anls':
            raise NotImplementedError(
                "method 'anls' (active-set NNLS, nmf/anls.py) is not provided by the B200 build; "
                "use 'mur', 'admm' or 'ao_admm'")
        else:
            raise Exception('Method not known. Choose one from: mur anls admm ao_admm')

        print('Factorization done.')
        if saving:
            self.save_factorization()

    def default_save_name(self):
        """File stem the reference derives from the experiment (nmf.py:95-126)."""
        e = self.results.experiment
        admm_like = e.method in {'admm', 'ao_admm'}
        parts = [f'nmf_{e.method}', str(self.factors), str(e.distance_type)]
        if e.method == 'admm':
            parts.append(str(e.rho))
        parts.append(f'{e.lambda_w}:{e.prox_w}' if admm_like else f'{e.lambda_w}')
        parts.append(f'{e.lambda_h}:{e.prox_h}' if admm_like else f'{e.lambda_h}')
        parts.append(f'nndsvd{e.nndsvd_init[1][0]}' if e.nndsvd_init[0] else 'random')
        if e.method == 'anls' and getattr(e, 'fcnnls', False):
            parts.append('fcnnls')
        return '_'.join(parts)

    def save_factorization(self, save_dir='./results', save_name=None):
        """np.savez of w, h, i, obj_history, experiment under save_dir (created if missing)."""
        os.makedirs(save_dir, exist_ok=True)
        if save_name is None:
            save_name = self.default_save_name()
        r = self.results
        utils.save_results(os.path.join(save_dir, save_name), w=r.w, h=r.h, i=r.i,
                           obj_history=r.obj_history, experiment=r.experiment._asdict())
