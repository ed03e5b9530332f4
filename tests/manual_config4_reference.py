"""Manual (test infrastructure): the compiled reference (oracle/_ref, QDLDL, 1 core) on one problem of
BASELINE config 4, for the comparison quoted in DESIGN.md section 8.   python tests/manual_config4_reference.py"""
import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from osqp_b200 import Solver, problems
base = problems.mpc(horizon=50, nx=20, nu=10, seed=1)
kw = dict(eps_abs=1e-3, eps_rel=1e-3, adaptive_rho_interval=50, warm_start=0)
s = Solver(base["P"], base["q"], base["A"], base["l"], base["u"], lib_path='oracle/_ref/libosqp_ref.so', linsys_solver=0, **kw)
s.solve()
print("reference QDLDL: setup %.1f ms, solve %.1f ms, %d iterations, %s" % (1e3 * s.info.setup_time, 1e3 * s.info.solve_time, s.info.iter, s.status))
for k in range(3):
    t = time.time(); s.solve(); print("re-solve %.2f ms, %d iterations" % (1e3 * (time.time() - t), s.info.iter))
