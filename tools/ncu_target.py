"""Short run for ncu: config-2 random QP, setup + a few ADMM iterations (plus timing of setup / first solves)."""
import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from osqp_b200 import Solver, problems
n = int(sys.argv[1]) if len(sys.argv) > 1 else 100000
max_iter = int(sys.argv[2]) if len(sys.argv) > 2 else 50
nsolve = int(sys.argv[3]) if len(sys.argv) > 3 else 1
p = problems.random_qp(n=n, seed=1)
t = time.time()
s = Solver(p['P'], p['q'], p['A'], p['l'], p['u'], eps_abs=1e-3, eps_rel=1e-3, adaptive_rho_interval=50, warm_start=0, max_iter=max_iter)
print("setup flag", s.setup_flag, "setup_time(info)=%.3f wall=%.3f" % (s.info.setup_time, time.time() - t), flush=True)
for k in range(nsolve):
    s.solve()
    print(k, s.status, s.info.iter, "solve=%.3f" % s.info.solve_time, "rho_upd", s.info.rho_updates, "rho", s.current_setting("rho"), s.stats(), flush=True)
