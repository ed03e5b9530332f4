"""Sweep the PCG stopping rule (rel tau, residual cap lambda): ADMM iterations, PCG iterations, solve time."""
import sys, os, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from osqp_b200 import Solver, problems
from osqp_b200.capi import Plugin
plug = Plugin()
PROBS = [
    ("random_qp n=30000", problems.random_qp(n=30000), {}),
    ("lasso nf=500", problems.lasso(n_feat=500, m_samp=5000, nnz_per_row=20), {}),
    ("lasso nf=1500", problems.lasso(n_feat=1500, m_samp=20000, nnz_per_row=50), {}),
    ("lasso nf=2000", problems.lasso(n_feat=2000, m_samp=20000, nnz_per_row=20), {}),
    ("lasso nf=10000", problems.lasso(n_feat=10000, m_samp=100000, nnz_per_row=50), {}),
    ("mpc N=50", problems.mpc(), {}),
    ("mpc N=20 nx10", problems.mpc(horizon=20, nx=10, nu=5), {}),
]
COMBOS = [(1e-6, 0.0), (1e-2, 0.0), (0.1, 0.0), (0.3, 0.15), (0.1, 0.15), (0.05, 0.15), (0.1, 0.05), (0.03, 0.5)]
if len(sys.argv) > 1:
    PROBS = [p for p in PROBS if sys.argv[1] in p[0]]
for eps in (1e-3, 1e-5):
    for name, p, kw in PROBS:
        for tol, lam in COMBOS:
            s = Solver(p['P'], p['q'], p['A'], p['l'], p['u'], eps_abs=eps, eps_rel=eps, adaptive_rho_interval=50, max_iter=20000, **kw)
            assert s.setup_flag == 0
            plug.set_pcg(s.engine_handle(), tol, 0.0, 2000)
            plug.set_pcg_lambda(s.engine_handle(), lam)
            s.solve()
            st = s.stats()
            print(f"eps={eps:g} {name:18s} tau={tol:7.0e} lam={lam:4.2f} st={s.info.status_val:2d} iter={s.info.iter:6d} pcg={st['pcg_iters']:8d} "
                  f"pcg/admm={st['pcg_iters']/max(1,st['admm_iters']):6.2f} maxed={st['pcg_maxed']:5d} brk={st['pcg_breakdowns']} launches={st['kernel_launches']} solve={s.info.solve_time*1e3:9.2f}ms obj={s.info.obj_val:.8g} rho_upd={s.info.rho_updates}", flush=True)
            s.cleanup()
