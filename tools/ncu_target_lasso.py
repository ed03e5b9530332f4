"""Short run for ncu: BASELINE config 3 (Lasso 1e5 x 1e6), setup + a few ADMM iterations."""
import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from osqp_b200 import Solver, problems
nf = int(sys.argv[1]) if len(sys.argv) > 1 else 100000
max_iter = int(sys.argv[2]) if len(sys.argv) > 2 else 40
p = problems.lasso(n_feat=nf, m_samp=10 * nf, nnz_per_row=50, seed=1)
s = Solver(p["P"], p["q"], p["A"], p["l"], p["u"], eps_abs=1e-3, eps_rel=1e-3, adaptive_rho_interval=50, warm_start=0, max_iter=max_iter)
print("setup flag", s.setup_flag, flush=True)
s.solve()
print(s.status, s.info.iter, s.stats(), flush=True)
