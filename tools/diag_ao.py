import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from oracle import nmf_oracle as orc
from nmf_b200.ao_admm import ao_admm
sys.path.insert(0, 'tests')
from conftest import load_golden, golden_v

g = load_golden('ao_admm_l1n'); r = g['recipe']
v = golden_v(r['v']); k = r['k']
w0, h0 = orc.seeded_init(v.shape[0], v.shape[1], k, r['init_seed'])
kw = dict(r['kw']); kw.pop('nndsvd_init')
tr = orc.ao_admm_run(v, k, w0, h0, **kw)
res = ao_admm(v, k, w0=w0, h0=h0, verbose=False, nndsvd_init=(False, 'zero'), **kw)
used = ao_admm.last_inner_iterations
for i in range(min(len(used), 40)):
    e = abs(res.obj_history[i + 1] - tr.obj_history[i + 1]) / tr.obj_history[i + 1]
    print(i, used[i], tr.inner_breaks[i], f"{e:.2e}", res.obj_history[i + 1], tr.obj_history[i + 1])
# one-iteration check of the first H update only
res1 = ao_admm(v, k, w0=w0, h0=h0, verbose=False, nndsvd_init=(False, 'zero'), **{**kw, 'max_iter': 1})
tr1 = orc.ao_admm_run(v, k, w0, h0, **{**kw, 'max_iter': 1})
print('iter1 relW', np.linalg.norm(res1.w - tr1.w) / np.linalg.norm(tr1.w), 'relH', np.linalg.norm(res1.h - tr1.h) / np.linalg.norm(tr1.h))
for ai in (1, 2, 3):
    kk = {**kw, 'max_iter': 1, 'admm_iter': ai}
    a = ao_admm(v, k, w0=w0, h0=h0, verbose=False, nndsvd_init=(False, 'zero'), **kk)
    b = orc.ao_admm_run(v, k, w0, h0, **kk)
    print('admm_iter', ai, 'relW', np.linalg.norm(a.w - b.w) / np.linalg.norm(b.w), 'relH', np.linalg.norm(a.h - b.h) / np.linalg.norm(b.h), ao_admm.last_inner_iterations, b.inner_breaks)
