"""Runs the UNMODIFIED reference (oracle/_ref/libosqp_ref.so, QDLDL backend) on seeded
problems from osqp_b200/problems.py and records status / iterations / objective /
residuals / solution digests in tests/golden/ref_outputs.json.  The oracle restatement
and the CUDA path are both checked against these records on boxes where the reference
library is absent.      python tests/golden/make_ref_outputs.py
"""
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from osqp_b200 import Solver, problems  # noqa: E402

REF = os.path.join(ROOT, "oracle", "_ref", "libosqp_ref.so")
CASES = {
    "demo": (problems.demo_qp, {}, {}),
    "basic_qp": (problems.basic_qp, {}, dict(polish=1, scaling=0, adaptive_rho_interval=100)),
    "basic_qp2": (problems.basic_qp2, {}, dict(polish=1, adaptive_rho_interval=100)),
    "random_qp_n100_e3": (problems.random_qp, dict(n=100), dict(adaptive_rho_interval=50)),
    "random_qp_n100_e5": (problems.random_qp, dict(n=100), dict(adaptive_rho_interval=50, eps_abs=1e-5, eps_rel=1e-5)),
    "random_qp_n500_e3": (problems.random_qp, dict(n=500), dict(adaptive_rho_interval=50)),
    "random_qp_n1000_e5": (problems.random_qp, dict(n=1000), dict(adaptive_rho_interval=50, eps_abs=1e-5, eps_rel=1e-5, max_iter=20000)),
    "random_qp_n2000_e5_polish": (problems.random_qp, dict(n=2000, seed=7), dict(adaptive_rho_interval=50, eps_abs=1e-5, eps_rel=1e-5, max_iter=20000, polish=1)),
    "mpc_small_e5": (problems.mpc, dict(horizon=10, nx=4, nu=2), dict(adaptive_rho_interval=50, eps_abs=1e-5, eps_rel=1e-5)),
    "mpc_N20_e5": (problems.mpc, dict(horizon=20, nx=10, nu=5), dict(adaptive_rho_interval=50, eps_abs=1e-5, eps_rel=1e-5)),
    "lasso_small_e5": (problems.lasso, dict(n_feat=50, m_samp=500, nnz_per_row=10), dict(adaptive_rho_interval=50, eps_abs=1e-5, eps_rel=1e-5)),
    "lasso_n200_e5": (problems.lasso, dict(n_feat=200, m_samp=2000, nnz_per_row=20), dict(adaptive_rho_interval=50, eps_abs=1e-5, eps_rel=1e-5)),
    # SURVEY 8(d) size ladder (QDLDL: n = 1e4 takes ~75 s of setup and minutes of solve on one core)
    "random_qp_n3000_e3": (problems.random_qp, dict(n=3000), dict(adaptive_rho_interval=50)),
    "random_qp_n3000_e5": (problems.random_qp, dict(n=3000), dict(adaptive_rho_interval=50, eps_abs=1e-5, eps_rel=1e-5, max_iter=20000)),
    "random_qp_n10000_e3": (problems.random_qp, dict(n=10000), dict(adaptive_rho_interval=50)),
    "random_qp_n10000_e5": (problems.random_qp, dict(n=10000), dict(adaptive_rho_interval=50, eps_abs=1e-5, eps_rel=1e-5, max_iter=20000)),
    # BASELINE config 4's real shape (horizon 50, 20 states, 10 inputs)
    "mpc_config4_e3": (problems.mpc, dict(horizon=50, nx=20, nu=10), dict(adaptive_rho_interval=50)),
    "mpc_config4_e5": (problems.mpc, dict(horizon=50, nx=20, nu=10), dict(adaptive_rho_interval=50, eps_abs=1e-5, eps_rel=1e-5)),
    # Lasso beyond the single-CTA size: the slack-elimination path (csrc/slack.cuh); n = m = 6000 / 23000
    "lasso_n500_e3": (problems.lasso, dict(n_feat=500, m_samp=5000, nnz_per_row=20), dict(adaptive_rho_interval=50)),
    "lasso_n500_e5": (problems.lasso, dict(n_feat=500, m_samp=5000, nnz_per_row=20), dict(adaptive_rho_interval=50, eps_abs=1e-5, eps_rel=1e-5, max_iter=20000)),
    "lasso_n1500_e3": (problems.lasso, dict(n_feat=1500, m_samp=20000, nnz_per_row=50), dict(adaptive_rho_interval=50)),
}


def main():
    """python tests/golden/make_ref_outputs.py [case,case,...  [out.json]]: the named cases (default: all) are
    (re)recorded and merged into the JSON file."""
    path = sys.argv[2] if len(sys.argv) > 2 else os.path.join(ROOT, "tests", "golden", "ref_outputs.json")
    only = [c for c in sys.argv[1].split(",") if c] if len(sys.argv) > 1 else list(CASES)
    out = json.load(open(path)) if os.path.exists(path) else {}
    for name, (gen, gkw, skw) in CASES.items():
        if name not in only:
            continue
        p = gen(**gkw)
        s = Solver(p["P"], p["q"], p["A"], p["l"], p["u"], lib_path=REF, linsys_solver=0, **skw)
        assert s.setup_flag == 0
        s.solve()
        i = s.info
        out[name] = dict(gen=gen.__name__, gen_kwargs=gkw, settings=skw, status_val=int(i.status_val),
                         iter=int(i.iter), obj_val=float(i.obj_val), pri_res=float(i.pri_res),
                         dua_res=float(i.dua_res), rho_updates=int(i.rho_updates),
                         status_polish=int(i.status_polish), setup_time=float(i.setup_time),
                         solve_time=float(i.solve_time),
                         x_head=s.x[:8].tolist(), y_head=s.y[:8].tolist(),
                         x_norm=float(np.linalg.norm(s.x)), y_norm=float(np.linalg.norm(s.y)))
        print(name, out[name]["status_val"], out[name]["iter"], out[name]["obj_val"])
        s.cleanup()
    with open(path, "w") as f:
        json.dump(out, f, indent=1)


if __name__ == "__main__":
    main()
