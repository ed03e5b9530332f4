"""The strict drop-in: the reference's OWN osqp_solve loop
(src/osqp.c:247-625, its own vector kernels, residual checks, adaptive rho and polish) with
libosqp_cuda_pcg.so as its KKT solver.  oracle/_ref/libosqp_ref_cuda.so is the unmodified reference
compiled with the dispatcher of oracle/ref_config/lin_sys_cuda_pcg.c (`make -C oracle ref_cuda`); the
backend is chosen the way a user would: settings->linsys_solver = 2."""
import os

import numpy as np
import pytest

from conftest import ROOT, TESTS_TOL, full_sym
from osqp_b200 import Solver, capi, problems

REF_CUDA = os.path.join(ROOT, "oracle", "_ref", "libosqp_ref_cuda.so")
needs_lib = pytest.mark.skipif(not os.path.exists(REF_CUDA), reason="needs oracle/_ref/libosqp_ref_cuda.so")


@needs_lib
def test_patched_reference_still_solves_with_qdldl_and_rejects_unknown_backends(fixtures):
    d = fixtures["basic_qp"]
    args = (full_sym(d["P"]), d["q"], d["A"], d["l"], d["u"])
    s = Solver(*args, lib_path=REF_CUDA, linsys_solver=0, polish=1, eps_abs=1e-5, eps_rel=1e-5)
    assert s.setup_flag == 0 and s.solve() == 0 and s.info.status_val == 1
    assert np.max(np.abs(s.x - d["sol_x_test"])) < TESTS_TOL
    s.cleanup()
    s = Solver(*args, lib_path=REF_CUDA, linsys_solver=7)
    assert s.setup_flag == 2          # OSQP_SETTINGS_VALIDATION_ERROR


@needs_lib
def test_backend_fails_loudly_without_a_gpu(monkeypatch):
    """no CPU fallback behind the reference's loader either: error 3 = OSQP_LINSYS_SOLVER_LOAD_ERROR"""
    import torch
    if torch.cuda.is_available():
        pytest.skip("this box has a GPU")
    monkeypatch.setenv("OSQP_CUDA_PCG_LIB", capi.PLUGIN_LIB)
    p = problems.random_qp(n=20, seed=1)
    s = Solver(p["P"], p["q"], p["A"], p["l"], p["u"], lib_path=REF_CUDA, linsys_solver=2)
    assert s.setup_flag == 3


@needs_lib
@pytest.mark.gpu
def test_reference_loop_with_cuda_pcg_backend(fixtures, monkeypatch):
    monkeypatch.setenv("OSQP_CUDA_PCG_LIB", capi.PLUGIN_LIB)
    d = fixtures["basic_qp"]
    cases = [dict(P=full_sym(d["P"]), q=d["q"], A=d["A"], l=d["l"], u=d["u"], name="basic_qp"),
             problems.random_qp(n=150, seed=21), problems.mpc(horizon=6, nx=4, nu=2, seed=3)]
    kw = dict(eps_abs=1e-5, eps_rel=1e-5, adaptive_rho_interval=50, polish=1)
    for p in cases:
        out = []
        for backend in (0, 2):
            s = Solver(p["P"], p["q"], p["A"], p["l"], p["u"], lib_path=REF_CUDA, linsys_solver=backend, **kw)
            assert s.setup_flag == 0, (p["name"], backend, s.setup_flag)
            assert s.solve() == 0
            out.append((int(s.info.status_val), int(s.info.iter), float(s.info.obj_val), int(s.info.status_polish),
                        s.x.copy(), s.y.copy()))
            s.cleanup()
        (st0, it0, ob0, pol0, x0, y0), (st2, it2, ob2, pol2, x2, y2) = out
        print(p["name"], "qdldl iters", it0, "cuda pcg iters", it2, "polish", pol0, pol2)
        assert st0 == st2 == 1
        assert it0 == it2                     # CG to 1e-10: the reference's trajectory, iteration for iteration
        assert abs(ob0 - ob2) <= 1e-6 * max(1.0, abs(ob0))
        assert np.max(np.abs(x0 - x2)) <= 1e-5 * max(1.0, np.max(np.abs(x0)))
        assert np.max(np.abs(y0 - y2)) <= 1e-4 * max(1.0, np.max(np.abs(y0)))
    d = fixtures["basic_qp"]
    assert np.max(np.abs(out[0][4] - out[1][4])) < 1.0      # (last case sanity; golden check below)
    s = Solver(full_sym(d["P"]), d["q"], d["A"], d["l"], d["u"], lib_path=REF_CUDA, linsys_solver=2, **kw)
    assert s.setup_flag == 0 and s.solve() == 0
    assert np.max(np.abs(s.x - d["sol_x_test"])) < TESTS_TOL and np.max(np.abs(s.y - d["sol_y_test"])) < TESTS_TOL
    s.cleanup()
