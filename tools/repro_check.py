"""Debug: is a second solve (after osqp_update_rho back to the initial rho) identical to the first?"""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from osqp_b200 import Solver, problems
base = problems.mpc(horizon=50, nx=20, nu=10, seed=1)
kw = dict(eps_abs=1e-3, eps_rel=1e-3, adaptive_rho_interval=50, warm_start=0)
if len(sys.argv) > 1:      # a different initial state, as tools/config4_batch.py does
    nx = base["nx"]
    x0 = np.random.default_rng(int(sys.argv[1])).standard_normal(nx)
    base["l"][:nx] = -x0
    base["u"][:nx] = -x0
s = Solver(base["P"], base["q"], base["A"], base["l"], base["u"], **kw)
out = []
for k in range(4):
    if k:
        print("update_rho flag", s.update_rho(0.1), "rho setting", s.current_setting("rho"))
    s.solve()
    out.append((s.x.copy(), s.y.copy()))
    print(k, s.status, int(s.info.iter), "rho_updates", int(s.info.rho_updates), "rho", s.current_setting("rho"), s.stats(),
          "dx vs first %.3e" % np.max(np.abs(out[-1][0] - out[0][0])))
s2 = Solver(base["P"], base["q"], base["A"], base["l"], base["u"], **kw)
s2.solve()
print("fresh workspace: dx vs first %.3e" % np.max(np.abs(s2.x - out[0][0])), int(s2.info.iter), s2.stats())
