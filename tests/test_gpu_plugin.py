"""The lin_sys plugin boundary on the GPU, in the reference's strict host-pointer form:
load_linsys_solver -> init_linsys_solver -> s->solve(s, b) / update_rho_vec / update_matrices /
free, exactly the calls of reference tests/solve_linsys/test_solve_linsys.h:12-46."""
import ctypes as C

import numpy as np
import pytest
import scipy.sparse as sp

import osqp_oracle as O
from conftest import TESTS_TOL, full_sym
from osqp_b200 import capi, problems

pytestmark = pytest.mark.gpu


def _plugin(P, A, sigma, rho_vec, polish=0):
    lib = capi.load_library()
    lib.load_linsys_solver.argtypes = [C.c_int]
    lib.load_linsys_solver.restype = capi.c_int
    lib.init_linsys_solver.argtypes = [C.POINTER(C.POINTER(capi.LinSysSolver)), C.POINTER(capi.csc),
                                       C.POINTER(capi.csc), capi.c_float, capi.c_float_p, C.c_int,
                                       capi.c_int]
    lib.init_linsys_solver.restype = capi.c_int
    assert lib.load_linsys_solver(capi.CUDA_PCG_SOLVER) == 0
    Ph, Ah = capi.CscHolder(P, upper=True), capi.CscHolder(A)
    s = C.POINTER(capi.LinSysSolver)()
    rho = np.ascontiguousarray(rho_vec, dtype=np.float64) if rho_vec is not None else None
    flag = lib.init_linsys_solver(C.byref(s), C.byref(Ph.struct), C.byref(Ah.struct), sigma,
                                  rho.ctypes.data_as(capi.c_float_p) if rho is not None else None,
                                  capi.CUDA_PCG_SOLVER, polish)
    return flag, s, (Ph, Ah, rho)


def test_solve_kkt_fixture(fixtures):
    """reference tests/solve_linsys/test_solve_linsys.h:12-46 with LINSYS_SOLVER = cuda_pcg"""
    d = fixtures["solve_linsys"]
    m = int(d["test_solve_KKT_m"])
    rho_vec = np.full(m, float(d["test_solve_KKT_rho"]))
    flag, s, keep = _plugin(full_sym(d["test_solve_KKT_Pu"]), d["test_solve_KKT_A"],
                            float(d["test_solve_KKT_sigma"]), rho_vec)
    assert flag == 0
    assert s.contents.type == capi.CUDA_PCG_SOLVER and s.contents.nthreads >= 1
    b = np.array(d["test_solve_KKT_rhs"], dtype=np.float64)
    assert s.contents.solve(s, b.ctypes.data_as(capi.c_float_p)) == 0
    assert np.max(np.abs(b - d["test_solve_KKT_x"])) < TESTS_TOL
    s.contents.free(s)


def test_plugin_matches_oracle_and_updates():
    """strict-mode solve on a larger system, then update_rho_vec and update_matrices"""
    p = problems.random_qp(n=300, seed=4)
    P, A = p["P"], p["A"]
    n, m = P.shape[0], A.shape[0]
    rng = np.random.default_rng(0)
    sigma, rho = 1e-3, 0.5
    flag, s, keep = _plugin(P, A, sigma, np.full(m, rho))
    assert flag == 0
    for trial in range(2):
        rhs = rng.standard_normal(n + m)
        want = O.solve_kkt(P, A, sigma, rho, rhs)
        b = rhs.copy()
        assert s.contents.solve(s, b.ctypes.data_as(capi.c_float_p)) == 0
        assert np.max(np.abs(b - want)) <= 1e-6 * max(1.0, np.max(np.abs(want)))
    rho2 = np.full(m, 2.0)
    assert s.contents.update_rho_vec(s, rho2.ctypes.data_as(capi.c_float_p)) == 0
    rhs = rng.standard_normal(n + m)
    want = O.solve_kkt(P, A, sigma, 2.0, rhs)
    b = rhs.copy()
    assert s.contents.solve(s, b.ctypes.data_as(capi.c_float_p)) == 0
    assert np.max(np.abs(b - want)) <= 1e-6 * max(1.0, np.max(np.abs(want)))
    # same pattern, new values
    P2 = sp.csc_matrix(P * 1.5)
    A2 = sp.csc_matrix(A * 0.5)
    Ph, Ah = capi.CscHolder(P2, upper=True), capi.CscHolder(A2)
    assert s.contents.update_matrices(s, C.byref(Ph.struct), C.byref(Ah.struct)) == 0
    want = O.solve_kkt(P2, A2, sigma, 2.0, rhs)
    b = rhs.copy()
    assert s.contents.solve(s, b.ctypes.data_as(capi.c_float_p)) == 0
    assert np.max(np.abs(b - want)) <= 1e-6 * max(1.0, np.max(np.abs(want)))
    s.contents.free(s)


def test_plugin_polish_mode():
    """polish = 1: rho_vec NULL, sigma carries delta, solve returns [x; nu]
    (reference qdldl_interface.c:261-272, 354-356; src/polish.c:232-234)"""
    rng = np.random.default_rng(1)
    n, m = 40, 15
    M = rng.standard_normal((n, n))
    P = sp.csc_matrix(M @ M.T + np.eye(n))
    A = sp.csc_matrix(rng.standard_normal((m, n)))
    delta = 1e-4
    flag, s, keep = _plugin(P, A, delta, None, polish=1)
    assert flag == 0
    rhs = rng.standard_normal(n + m)
    K = sp.bmat([[P + delta * sp.identity(n), A.T], [A, -delta * sp.identity(m)]], format="csc")
    want = sp.linalg.spsolve(K, rhs)
    b = rhs.copy()
    assert s.contents.solve(s, b.ctypes.data_as(capi.c_float_p)) == 0
    assert np.max(np.abs(b[:n] - want[:n])) <= 1e-5 * max(1.0, np.max(np.abs(want[:n])))
    assert np.max(np.abs(b[n:] - want[n:])) <= 1e-3 * max(1.0, np.max(np.abs(want[n:])))
    s.contents.free(s)


def test_plugin_detects_nonconvex(fixtures):
    """reference qdldl_interface.c:93-98,305-310: init returns OSQP_NONCVX_ERROR"""
    d = fixtures["non_cvx"]
    m = d["A"].shape[0]
    flag, s, keep = _plugin(full_sym(d["P"]), d["A"], 1e-6, np.full(m, 0.1))
    assert flag == capi.OSQP_NONCVX_ERROR
    assert not s
