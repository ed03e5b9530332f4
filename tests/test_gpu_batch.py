"""osqp_solve_batch (BASELINE config 4, additive API): a batch of MPC QPs that differ in their
initial state must give exactly what osqp_solve gives one problem at a time."""
import numpy as np
import pytest
import scipy.sparse as sp

from osqp_b200 import Solver, problems
from osqp_b200.capi import solve_batch


@pytest.mark.gpu
@pytest.mark.timeout(600)
def test_batch_matches_individual_solves():
    base = problems.mpc(horizon=10, nx=4, nu=2, seed=2)
    nx = base["nx"]
    rng = np.random.default_rng(5)
    kw = dict(eps_abs=1e-4, eps_rel=1e-4, adaptive_rho_interval=25, warm_start=0)
    solvers = []
    for _ in range(12):
        x0 = 0.5 * rng.standard_normal(nx)
        l, u = base["l"].copy(), base["u"].copy()
        l[:nx] = -x0
        u[:nx] = -x0
        s = Solver(base["P"], base["q"], base["A"], l, u, **kw)
        assert s.setup_flag == 0
        solvers.append(s)
    ref = []
    for s in solvers:
        assert s.solve() == 0
        ref.append((s.x.copy(), s.y.copy(), int(s.info.iter), int(s.info.status_val)))
    assert len({r[2] for r in ref}) >= 1
    for threads in (1, 4):
        for s in solvers:
            s.update_rho(0.1)      # adaptive rho persists across solves (as in the reference): same start
        assert solve_batch(solvers, threads) == 0
        for s, (x, y, it, st) in zip(solvers, ref):
            assert int(s.info.status_val) == st and int(s.info.iter) == it
            assert np.array_equal(s.x, x) and np.array_equal(s.y, y)
    for s in solvers:
        s.cleanup()


@pytest.mark.gpu
def test_fused_and_multikernel_paths_agree(monkeypatch):
    """Small problems run whole ADMM iterations in one thread block (admm_fused.cuh); with
    OSQP_B200_FUSED=0 the same problem goes through the multi-kernel engine.  Same algorithm,
    different summation order: same status and iteration count, solutions equal to rounding."""
    cases = [problems.basic_qp(), problems.mpc(horizon=10, nx=4, nu=2, seed=2), problems.random_qp(n=60, seed=4)]
    for p in cases:
        out = []
        for fused in ("1", "0"):
            monkeypatch.setenv("OSQP_B200_FUSED", fused)
            s = Solver(p["P"], p["q"], p["A"], p["l"], p["u"], eps_abs=1e-5, eps_rel=1e-5,
                       adaptive_rho_interval=25, polish=0)
            assert s.setup_flag == 0 and s.solve() == 0
            out.append((int(s.info.status_val), int(s.info.iter), s.x.copy(), s.y.copy(), s.stats()["kernel_launches"]))
            s.cleanup()
        (st1, it1, x1, y1, l1), (st0, it0, x0, y0, l0) = out
        assert st1 == st0 and it1 == it0, (p["name"], it1, it0)
        assert np.max(np.abs(x1 - x0)) <= 1e-9 * max(1.0, np.max(np.abs(x0)))
        assert np.max(np.abs(y1 - y0)) <= 1e-9 * max(1.0, np.max(np.abs(y0)))
        assert l1 < l0          # the fused path really is the one that ran


@pytest.mark.gpu
@pytest.mark.timeout(600)
def test_lockstep_batch_mixed_problems_and_fallback(monkeypatch):
    """The lockstep driver on a batch whose members differ in size, finish at different check
    intervals, adapt rho at different times and include a primal-infeasible problem; then the same
    batch with one deviating schedule (falls back to the threaded path) and with the lockstep
    driver switched off.  Every run must reproduce osqp_solve problem by problem."""
    kw = dict(eps_abs=1e-4, eps_rel=1e-4, adaptive_rho_interval=25, warm_start=0, polish=1)
    probs = [problems.random_qp(n=30 + 7 * k, seed=10 + k) for k in range(6)]
    probs += [problems.mpc(horizon=8, nx=4, nu=2, seed=k) for k in range(3)]
    inf = problems.random_qp(n=40, seed=3)
    A2 = sp.vstack([inf["A"], inf["A"][:1]], format="csc")       # row 0 twice with disjoint bounds
    inf = dict(P=inf["P"], q=inf["q"], A=A2, l=np.hstack([inf["l"], inf["u"][0] + 1.0]),
               u=np.hstack([inf["u"], inf["u"][0] + 2.0]))
    probs.append(inf)
    solvers = [Solver(p["P"], p["q"], p["A"], p["l"], p["u"], **kw) for p in probs]
    assert all(s.setup_flag == 0 for s in solvers)
    ref = []
    for s in solvers:
        assert s.solve() == 0
        ref.append((int(s.info.status_val), int(s.info.iter), int(s.info.status_polish), s.x.copy(), s.y.copy()))
    assert len({r[1] for r in ref}) > 1               # they really finish at different intervals
    assert any(r[0] == -3 for r in ref)               # the infeasible one is detected

    def run_and_check(threads):
        for s in solvers:
            s.update_rho(0.1)
        assert solve_batch(solvers, threads) == 0
        for s, (st, it, pol, x, y) in zip(solvers, ref):
            assert int(s.info.status_val) == st and int(s.info.iter) == it and int(s.info.status_polish) == pol
            assert np.array_equal(s.x, x, equal_nan=True) and np.array_equal(s.y, y, equal_nan=True)

    run_and_check(1)                                   # lockstep
    solvers[2].update_setting("max_iter", 3999)        # one schedule differs: threaded fallback
    run_and_check(3)
    solvers[2].update_setting("max_iter", 4000)
    monkeypatch.setenv("OSQP_B200_LOCKSTEP", "0")
    run_and_check(2)
    for s in solvers:
        s.cleanup()
