"""Slack elimination (osqp_b200/csrc/slack.cuh): singleton columns of A with a
diagonal P entry are removed exactly from the system the CG solves.  The ADMM iteration is the same
mathematics with or without it, so the two must agree on status, objective and solution, and the
elimination must pay in CG iterations."""
import os

import numpy as np
import pytest
import scipy.sparse as sp

from osqp_b200 import Solver, problems

pytestmark = pytest.mark.gpu


def _solve(p, slack, **kw):
    old = os.environ.get("OSQP_B200_SLACK")
    os.environ["OSQP_B200_SLACK"] = slack
    try:
        s = Solver(p["P"], p["q"], p["A"], p["l"], p["u"], **kw)
        assert s.setup_flag == 0 and s.solve() == 0
        out = dict(status=s.info.status_val, iter=int(s.info.iter), obj=s.info.obj_val, x=s.x.copy(), y=s.y.copy(),
                   stats=s.stats(), polish=int(s.info.status_polish))
        s.cleanup()
        return out
    finally:
        if old is None:
            del os.environ["OSQP_B200_SLACK"]
        else:
            os.environ["OSQP_B200_SLACK"] = old


@pytest.mark.parametrize("eps,polish", [(1e-3, 0), (1e-5, 1)])
def test_lasso_with_and_without_elimination(eps, polish):
    p = problems.lasso(n_feat=500, m_samp=5000, nnz_per_row=20, seed=1)     # n = m = 6000: multi-kernel path
    kw = dict(eps_abs=eps, eps_rel=eps, adaptive_rho_interval=50, polish=polish, max_iter=20000)
    on = _solve(p, "1", **kw)
    off = _solve(p, "0", **kw)
    assert on["stats"]["slack_vars"] == 5000 and off["stats"]["slack_vars"] == 0
    assert on["status"] == off["status"] == 1
    assert on["polish"] == off["polish"]
    assert abs(on["obj"] - off["obj"]) <= 10 * eps * max(1.0, abs(off["obj"]))
    assert np.max(np.abs(on["x"] - off["x"])) <= 30 * eps * max(1.0, np.max(np.abs(off["x"])))
    per_on = on["stats"]["pcg_iters"] / on["iter"]
    per_off = off["stats"]["pcg_iters"] / off["iter"]
    print(f"eps {eps}: ADMM iterations {on['iter']} / {off['iter']}, CG per ADMM iteration {per_on:.2f} / {per_off:.2f}")
    assert per_on * 2.0 <= per_off, (per_on, per_off)
    # the returned point satisfies the KKT conditions to the tolerance, recomputed on the host
    A, P = sp.csc_matrix(p["A"]), sp.csc_matrix(p["P"])
    x, y = on["x"], on["y"]
    Ax = A @ x
    pri = np.max(np.abs(Ax - np.clip(Ax, p["l"], p["u"])))
    dua = np.max(np.abs(P @ x + p["q"] + A.T @ y))
    assert pri <= eps * (1 + np.max(np.abs(Ax)))
    assert dua <= eps * (1 + max(np.max(np.abs(P @ x)), np.max(np.abs(A.T @ y)), np.max(np.abs(p["q"]))))


def test_mixed_structure_partial_slack_set():
    """A random QP with slack columns appended: half with a diagonal P entry, half with P_jj = 0
    (D_j = sigma + a_j^2 rho_r), and 100 that share a partner row with an earlier one (only the first
    of a row is eliminated)."""
    rng = np.random.default_rng(5)
    base = problems.random_qp(n=1500, seed=3)
    P0, A0 = sp.csc_matrix(base["P"]), sp.csc_matrix(base["A"])
    n0, m0 = A0.shape[1], A0.shape[0]
    ns = 1200
    rows = np.concatenate([rng.permutation(m0)[:ns - 100], rng.integers(0, 50, 100)])
    S = sp.csc_matrix((rng.choice([-1.0, 1.0, 0.5], ns), (rows, np.arange(ns))), shape=(m0, ns))
    has_p = rng.random(ns) < 0.5
    P = sp.block_diag([P0, sp.diags(np.where(has_p, rng.uniform(0.5, 2.0, ns), 0.0))], format="csc")
    P.eliminate_zeros()
    A = sp.hstack([A0, S], format="csc")
    q = np.concatenate([base["q"], np.where(has_p, 0.1 * rng.standard_normal(ns), 0.0)])
    p = dict(P=P, q=q, A=A, l=base["l"], u=base["u"])
    kw = dict(eps_abs=1e-4, eps_rel=1e-4, adaptive_rho_interval=50, max_iter=20000)
    on = _solve(p, "1", **kw)
    off = _solve(p, "0", **kw)
    print("slack_vars", on["stats"]["slack_vars"], "iters", on["iter"], off["iter"])
    assert on["stats"]["slack_vars"] == len(np.unique(rows))
    assert on["status"] == off["status"] == 1
    assert abs(on["obj"] - off["obj"]) <= 1e-3 * max(1.0, abs(off["obj"]))


def test_singleton_columns_in_inequality_rows():
    """Slack columns (one entry, diagonal P) whose partner rows are INEQUALITIES and loose rows."""
    rng = np.random.default_rng(9)
    base = problems.random_qp(n=2500, seed=4)
    P0, A0 = sp.csc_matrix(base["P"]), sp.csc_matrix(base["A"])
    n0, m0 = A0.shape[1], A0.shape[0]
    ns = 2000
    rows = rng.permutation(m0)[:ns]
    S = sp.csc_matrix((rng.choice([-1.0, 1.0, 2.0], ns), (rows, np.arange(ns))), shape=(m0, ns))
    P = sp.block_diag([P0, sp.diags(rng.uniform(0.5, 2.0, ns))], format="csc")
    A = sp.hstack([A0, S], format="csc")
    q = np.concatenate([base["q"], 0.1 * rng.standard_normal(ns)])
    p = dict(P=P, q=q, A=A, l=base["l"], u=base["u"])
    kw = dict(eps_abs=1e-5, eps_rel=1e-5, adaptive_rho_interval=50, max_iter=20000)
    on = _solve(p, "1", **kw)
    off = _solve(p, "0", **kw)
    assert on["stats"]["slack_vars"] == ns
    assert on["status"] == off["status"] == 1
    assert abs(on["iter"] - off["iter"]) <= 100, (on["iter"], off["iter"])
    assert abs(on["obj"] - off["obj"]) <= 1e-4 * max(1.0, abs(off["obj"]))
    assert np.max(np.abs(on["x"] - off["x"])) <= 1e-3 * max(1.0, np.max(np.abs(off["x"])))
