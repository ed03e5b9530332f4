"""Parametric re-solve loop of reference docs/examples/lasso.rst:65-78 on the device: ONE osqp_setup,
then for every gamma osqp_update_lin_cost + a warm-started osqp_solve (the
documented production pattern "setup once, update vectors, warm-start, re-solve").
usage: python tools/lasso_path.py [n_feat=100000] [n_gammas=8] [eps=1e-3]"""
import sys, os, time, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from osqp_b200 import Solver, problems

nf = int(sys.argv[1]) if len(sys.argv) > 1 else 100000
ng = int(sys.argv[2]) if len(sys.argv) > 2 else 8
eps = float(sys.argv[3]) if len(sys.argv) > 3 else 1e-3
p = problems.lasso(n_feat=nf, m_samp=10 * nf, nnz_per_row=50, seed=1)
n = p["A"].shape[1]
gmax = p["gamma"] / 0.2                      # = ||Ad' b||inf: above it the solution is x = 0
gammas = gmax * np.linspace(0.6, 0.05, ng)   # from sparse to dense solutions
kw = dict(eps_abs=eps, eps_rel=eps, adaptive_rho_interval=50, max_iter=20000, warm_start=1)

t = time.time()
s = Solver(p["P"], p["q"], p["A"], p["l"], p["u"], **kw)
assert s.setup_flag == 0
print(json.dumps(dict(workload=p["name"], n=n, m=int(p["A"].shape[0]), nnzA=int(p["A"].nnz),
                      setup_s=round(float(s.info.setup_time), 3), wall_setup_s=round(time.time() - t, 3))), flush=True)
tot_it, tot_s = 0, 0.0
for g in gammas:
    q = np.hstack([np.zeros(n - nf), g * np.ones(nf)])
    assert s.update_lin_cost(q) == 0
    assert s.solve() == 0
    x = s.x
    tot_it += int(s.info.iter); tot_s += float(s.info.solve_time)
    print(json.dumps(dict(gamma_over_max=round(float(g / gmax), 3), status=s.status, iters=int(s.info.iter),
                          solve_s=round(float(s.info.solve_time), 3), obj=float(s.info.obj_val),
                          nnz_x=int(np.sum(np.abs(x[:nf]) > 1e-4 * max(1e-12, np.max(np.abs(x[:nf]))))))), flush=True)
print(json.dumps(dict(total_iters=tot_it, total_solve_s=round(tot_s, 3), iters_per_s=round(tot_it / tot_s, 1))), flush=True)
s.cleanup()
