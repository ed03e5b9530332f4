"""Short run for ncu: one BASELINE config-4 MPC problem (fused single-block kernel)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from osqp_b200 import Solver, problems
p = problems.mpc(horizon=50, nx=20, nu=10, seed=1)
s = Solver(p["P"], p["q"], p["A"], p["l"], p["u"], eps_abs=1e-3, eps_rel=1e-3, adaptive_rho_interval=50, warm_start=0)
assert s.setup_flag == 0
s.solve()
print(s.status, s.info.iter, "solve_ms=%.2f" % (1e3 * s.info.solve_time), s.stats(), flush=True)
