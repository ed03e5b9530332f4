"""Where osqp_setup spends its time on config 2 (OSQP_B200_TIMING=1 prints phase times)."""
import sys, os, time
os.environ.setdefault("PYTHONUNBUFFERED", "1")
os.environ["OSQP_B200_TIMING"] = "1"
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from osqp_b200 import Solver, problems
p = problems.random_qp(n=100000, seed=1)
for k in range(3):
    t = time.time()
    s = Solver(p['P'], p['q'], p['A'], p['l'], p['u'], eps_abs=1e-3, eps_rel=1e-3, adaptive_rho_interval=50)
    t1 = time.time()
    print("setup wall %.1f ms (info %.1f ms)" % (1e3 * (t1 - t), 1e3 * s.info.setup_time), flush=True)
    t = time.time(); s.cleanup(); print("cleanup %.1f ms" % (1e3 * (time.time() - t)), flush=True)
