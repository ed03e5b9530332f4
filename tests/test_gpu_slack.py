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


def test_matrix_and_bound_updates_refresh_the_elimination():
    """osqp_update_P_A / osqp_update_bounds / osqp_update_rho on a problem with eliminated slacks: the
    entries a_j, P_jj, the row types and rho all enter D_j and rho~, which engine_refresh_linsys
    recomputes; the re-solve must agree with a fresh setup on the updated data."""
    rng = np.random.default_rng(21)
    p = problems.lasso(n_feat=400, m_samp=4000, nnz_per_row=20, seed=6)      # n = m = 4800: multi-kernel path
    P, A = sp.csc_matrix(sp.triu(p["P"])), sp.csc_matrix(p["A"])
    P.sort_indices(); A.sort_indices()
    kw = dict(eps_abs=1e-5, eps_rel=1e-5, adaptive_rho_interval=50, max_iter=20000)
    old = os.environ.get("OSQP_B200_SLACK")
    os.environ["OSQP_B200_SLACK"] = "1"
    try:
        s = Solver(P, p["q"], A, p["l"], p["u"], **kw)
        assert s.setup_flag == 0 and s.solve() == 0
        assert s.stats()["slack_vars"] == 4000
        # new values on the same patterns: every entry of A scaled (the slacks' -1 included), P's diagonal block too
        Ax2 = A.data * (1.0 + 0.2 * rng.random(A.nnz))
        Px2 = P.data * 1.7
        assert s.update_P_A(Px2, None, Ax2, None) == 0
        # some equality rows become inequalities (rho and rho~ change on those partner rows)
        l2, u2 = p["l"].copy(), p["u"].copy()
        u2[:500] += 0.05
        assert s.update_bounds(l2, u2) == 0
        assert s.update_rho(0.3) == 0
        assert s.solve() == 0
        P2 = sp.csc_matrix((Px2, P.indices, P.indptr), shape=P.shape)
        A2 = sp.csc_matrix((Ax2, A.indices, A.indptr), shape=A.shape)
        f = Solver(P2, p["q"], A2, l2, u2, rho=0.3, **kw)
        assert f.setup_flag == 0 and f.solve() == 0
        assert s.info.status_val == f.info.status_val == 1
        assert abs(s.info.obj_val - f.info.obj_val) <= 1e-4 * max(1.0, abs(f.info.obj_val))
        assert np.max(np.abs(s.x - f.x)) <= 1e-3 * max(1.0, np.max(np.abs(f.x)))
        s.cleanup(); f.cleanup()
    finally:
        if old is None:
            del os.environ["OSQP_B200_SLACK"]
        else:
            os.environ["OSQP_B200_SLACK"] = old
