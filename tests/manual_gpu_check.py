"""Manual check on a B200 (test infrastructure, not collected by pytest): the product (CUDA) next to the
compiled reference (oracle/_ref) on small problems, printing status / objective / iteration deltas.
   python tests/manual_gpu_check.py"""
import sys, time, os
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from osqp_b200 import Solver, problems
REF = 'oracle/_ref/libosqp_ref.so'

def run(p, **kw):
    out = {}
    for name, lib, ls in (('ref', REF, 0), ('gpu', None, 2)):
        t = time.time()
        s = Solver(p['P'], p['q'], p['A'], p['l'], p['u'], lib_path=lib, linsys_solver=ls, **kw)
        if s.setup_flag:
            print(f"  {name}: setup failed {s.setup_flag}"); out[name] = None; continue
        f = s.solve()
        i = s.info
        extra = ''
        if name == 'gpu':
            extra = str(s.stats())
        print(f"  {name}: flag={f} status={s.status} iter={i.iter} obj={i.obj_val:.8g} pri={i.pri_res:.3e} dua={i.dua_res:.3e} "
              f"setup={i.setup_time:.4f}s solve={i.solve_time:.4f}s rho_upd={i.rho_updates} wall={time.time()-t:.2f}s {extra}", flush=True)
        out[name] = (s.info.status_val, s.info.obj_val, s.x, s.y, s.info.iter)
        s.cleanup()
    if out['ref'] and out['gpu']:
        r, g = out['ref'], out['gpu']
        print(f"  -> status_eq={r[0]==g[0]} obj_rel={abs(r[1]-g[1])/max(1,abs(r[1])):.2e} "
              f"dx={np.max(np.abs(r[2]-g[2])) if len(r[2]) else 0:.2e} dy={np.max(np.abs(r[3]-g[3])) if len(r[3]) else 0:.2e} diter={g[4]-r[4]}")

for gen, kw in ((problems.demo_qp, {}), (problems.basic_qp, dict(polish=1, scaling=0)), (problems.basic_qp2, dict(polish=1))):
    p = gen(); print(p['name'], kw, flush=True); run(p, **kw)
for n in (100, 1000, 3000):
    p = problems.random_qp(n=n)
    for eps in (1e-3, 1e-5):
        kw = dict(eps_abs=eps, eps_rel=eps, adaptive_rho_interval=50)
        print(p['name'], kw, flush=True); run(p, **kw)
p = problems.mpc(horizon=10, nx=4, nu=2)
print(p['name'], flush=True); run(p, eps_abs=1e-5, eps_rel=1e-5, adaptive_rho_interval=50)
p = problems.lasso(n_feat=50, m_samp=500, nnz_per_row=10)
print(p['name'], flush=True); run(p, eps_abs=1e-5, eps_rel=1e-5, adaptive_rho_interval=50)
