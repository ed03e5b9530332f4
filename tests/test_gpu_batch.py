"""osqp_solve_batch (BASELINE config 4, additive API): a batch of MPC QPs that differ in their
initial state must give exactly what osqp_solve gives one problem at a time."""
import numpy as np
import pytest

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
