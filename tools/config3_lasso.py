"""BASELINE config 3 (Lasso, n_f = 1e5 features, 1e6 samples, ~5e7 nnz) on one B200: setup + solve timings."""
import sys, os, time, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from osqp_b200 import Solver, problems
nf = int(sys.argv[1]) if len(sys.argv) > 1 else 100000
t = time.time()
p = problems.lasso(n_feat=nf, m_samp=10 * nf, nnz_per_row=50, seed=1)
gen = time.time() - t
for eps in (1e-3, 1e-5):
    t = time.time()
    s = Solver(p["P"], p["q"], p["A"], p["l"], p["u"], eps_abs=eps, eps_rel=eps, adaptive_rho_interval=50, max_iter=20000)
    setup = time.time() - t
    assert s.setup_flag == 0
    s.solve()
    st = s.stats()
    x = s.x
    # KKT check of the returned point in numpy (size-independent property)
    A, P = p["A"], p["P"]
    Pf = P + P.T - __import__("scipy.sparse").sparse.diags(P.diagonal())
    pri = np.max(np.maximum(A @ x - p["u"], 0) + np.maximum(p["l"] - A @ x, 0))
    dua = np.max(np.abs(Pf @ x + p["q"] + A.T @ s.y))
    print(json.dumps(dict(workload=p["name"], n=int(A.shape[1]), m=int(A.shape[0]), nnzA=int(A.nnz), eps=eps,
                          status=s.status, iters=int(s.info.iter), obj=float(s.info.obj_val),
                          setup_s=round(float(s.info.setup_time), 3), solve_s=round(float(s.info.solve_time), 3),
                          iters_per_s=round(int(s.info.iter) / float(s.info.solve_time), 1),
                          pcg_iters_per_admm=round(st["pcg_iters"] / max(1, st["admm_iters"]), 2),
                          rho_updates=int(s.info.rho_updates), slack_vars=int(st.get("slack_vars", 0)), pri_violation_numpy=float(pri), dua_residual_numpy=float(dua),
                          generate_s=round(gen, 1))), flush=True)
    s.cleanup()
