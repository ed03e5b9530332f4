"""Kernel-level parity (SURVEY 8c row "SpMV, norms, quad form"): the SpMV launches of
the solve path through the C-ABI check point osqp_cuda_engine_spmv, against the reference's own lin_alg
vectors (reference tests/lin_alg/generate_problem.py:42-78, asserted at tests/lin_alg/test_lin_alg.h:112-291)
and against scipy on matrices large enough to take the column-blocked and sliced-ELL kernels; the Ruiz
scaling D, E, c of osqp_setup (reference src/scaling.c:44-156) against the oracle's restatement, which
tests/test_oracle.py pins to the compiled reference."""
import ctypes as C

import os

import numpy as np
import pytest
import scipy.sparse as sp

import osqp_oracle as O
from conftest import full_sym
from osqp_b200 import Solver, capi, problems

TOL = 1e-12          # fp64 products, summation order differs from numpy's: relative to the largest entry


def _lib():
    L = capi.Plugin().lib
    L.osqp_cuda_engine_spmv.argtypes = [C.c_void_p, capi.c_int, capi.c_float_p, capi.c_float_p]
    L.osqp_cuda_engine_spmv.restype = capi.c_int
    L.osqp_cuda_engine_create.argtypes = [C.POINTER(C.c_void_p), capi.c_int, capi.c_int, C.POINTER(capi.csc),
                                          C.POINTER(capi.csc), capi.c_float_p, capi.c_float_p, capi.c_float_p]
    L.osqp_cuda_engine_create.restype = capi.c_int
    L.osqp_cuda_engine_free.argtypes = [C.c_void_p]
    L.osqp_cuda_engine_free.restype = None
    return L


class _RawEngine:
    """A bare device engine (osqp_cuda_engine_create): the matrices on the device, no linear system --
    the reference's lin_alg matrices are not positive semidefinite, osqp_setup rightly refuses them."""

    def __init__(self, Pfull, A):
        self.L = _lib()
        m, n = A.shape
        self.keep = (capi.CscHolder(Pfull, upper=True), capi.CscHolder(sp.csc_matrix(A)))
        z = np.zeros(max(n, m, 1))
        self.e = C.c_void_p()
        fp = lambda a: a.ctypes.data_as(capi.c_float_p)
        assert self.L.osqp_cuda_engine_create(C.byref(self.e), n, m, C.byref(self.keep[0].struct),
                                              C.byref(self.keep[1].struct), fp(z), fp(z), fp(z)) == 0

    def engine_handle(self):
        return self.e

    def free(self):
        self.L.osqp_cuda_engine_free(self.e)


def _spmv(s, which, x, nout):
    L = _lib()
    x = np.ascontiguousarray(x, dtype=np.float64)
    y = np.full(nout, np.nan)
    assert L.osqp_cuda_engine_spmv(s.engine_handle(), which, x.ctypes.data_as(capi.c_float_p),
                                   y.ctypes.data_as(capi.c_float_p)) == 0
    return y


def _close(got, want):
    return np.max(np.abs(got - want)) <= TOL * max(1.0, np.max(np.abs(want)))


def _check_all(s, Pf, A, rng):
    m, n = A.shape
    x, y, x2 = rng.standard_normal(n), rng.standard_normal(m), rng.standard_normal(n)
    assert _close(_spmv(s, 0, x, m), A @ x)
    assert _close(_spmv(s, 1, np.concatenate([x, y]), n), Pf @ x + A.T @ y)
    got = _spmv(s, 2, np.concatenate([x, x2]), 2 * m)
    assert _close(got[:m], A @ x) and _close(got[m:], A @ x2)


@pytest.mark.gpu
def test_mat_vec_reference_vectors(fixtures):
    """reference tests/lin_alg/test_lin_alg.h:195-262 (test_mat_vec_multiplication): A x, A' y and the
    symmetric P x completed from its upper triangle, with the reference's stored answers."""
    d = fixtures["lin_alg"]
    A, Pu = sp.csc_matrix(d["test_mat_vec_A"]), sp.csc_matrix(d["test_mat_vec_Pu"])
    x, y = np.asarray(d["test_mat_vec_x"], float), np.asarray(d["test_mat_vec_y"], float)
    m, n = A.shape
    s = _RawEngine(full_sym(Pu), A)
    assert np.max(np.abs(_spmv(s, 0, x, m) - np.asarray(d["test_mat_vec_Ax"]))) <= 1e-12
    # [P | A'] [x ; 0] = P x ;  [P | A'] [0 ; y] = A' y ;  both at once = their sum
    assert np.max(np.abs(_spmv(s, 1, np.concatenate([x, 0 * y]), n) - np.asarray(d["test_mat_vec_Px"]))) <= 1e-12
    assert np.max(np.abs(_spmv(s, 1, np.concatenate([0 * x, y]), n) - np.asarray(d["test_mat_vec_ATy"]))) <= 1e-12
    both = _spmv(s, 1, np.concatenate([x, y]), n)
    assert np.max(np.abs(both - (np.asarray(d["test_mat_vec_Px"]) + np.asarray(d["test_mat_vec_ATy"])))) <= 1e-12
    # the "cumulative" variants of the reference (y += A x) follow from the same products
    assert np.max(np.abs(y + _spmv(s, 0, x, m) - np.asarray(d["test_mat_vec_Ax_cum"]))) <= 1e-12
    # quad_form (reference test_lin_alg.h:278-291): x' P x / 2 from the same P x pass
    Pq, xq = sp.csc_matrix(d["test_qpform_Pu"]), np.asarray(d["test_qpform_x"], float)
    nq = Pq.shape[0]
    s2 = _RawEngine(full_sym(Pq), sp.csc_matrix(np.eye(nq)))
    Px = _spmv(s2, 1, np.concatenate([xq, np.zeros(nq)]), nq)
    assert abs(0.5 * xq @ Px - float(d["test_qpform_value"])) <= 1e-12
    s.free(); s2.free()


@pytest.mark.gpu
@pytest.mark.parametrize("case", ["small_direct", "sell", "tile", "tile_ragged", "tile_lasso"])
def test_spmv_kernels_against_scipy(case, monkeypatch):
    monkeypatch.setenv("OSQP_B200_TILE_ROWS_MAXPCT", "400")     # keep the ROWS kernel on random sparsity (default bar: 60 %)
    rng = np.random.default_rng(17)
    if case == "small_direct":          # fused-size problem: CSR kernels
        p = problems.random_qp(n=60, seed=2)
    elif case == "sell":                # sliced-ELL mirrors, too small / too sparse per block for the tiles
        p = problems.random_qp(n=6000, seed=3)
    elif case == "tile":                # BASELINE config 2's shape at 0.4 scale: A in ROWS mode, A' in PARTIAL mode
        p = problems.random_qp(n=40000, seed=11)
    elif case == "tile_lasso":          # A = [A_d -I 0; I 0 -I; I 0 I]: ROWS mode with most (CTA, block) phases empty,
        p = problems.lasso(n_feat=5000, m_samp=50000, nnz_per_row=30, seed=4)   # row-capped ranges over the 2-entry rows
    else:                               # empty rows / columns, 2-entry rows next to 900-entry rows
        n, m = 60000, 300000
        rows = np.repeat(np.arange(m), 16)
        cols = rng.integers(0, n, size=16 * m)
        keep = ((rows % 7) != 0) & ~((cols >= 100) & (cols < 140))
        long_rows = np.array([40, 41, 5000, m - 1])
        lr = np.repeat(long_rows, 900)
        lc = np.concatenate([rng.choice(n, 900, replace=False) for _ in long_rows])
        A = sp.csc_matrix((rng.standard_normal(int(keep.sum()) + lr.size),
                           (np.concatenate([rows[keep], lr]), np.concatenate([cols[keep], lc]))), shape=(m, n))
        A.sum_duplicates(); A.sort_indices()
        Mx = sp.random(n, n, density=2.0 / n, random_state=8, format="csc")
        P = sp.csc_matrix(Mx @ Mx.T + 1e-2 * sp.eye(n))
        p = dict(P=P, q=rng.standard_normal(n), A=A, l=-np.ones(m), u=np.ones(m))
    s = Solver(p["P"], p["q"], p["A"], p["l"], p["u"], scaling=0)
    assert s.setup_flag == 0
    A = sp.csc_matrix(p["A"])
    Pf = full_sym(sp.triu(sp.csc_matrix(p["P"])))
    st0 = s.stats()
    _check_all(s, Pf, A, rng)
    st1 = s.stats()
    if case == "tile_lasso":
        assert st1["tile_launches"] - st0["tile_launches"] in (1, 2), "A should take the column-blocked ROWS kernel"
    elif case.startswith("tile"):
        assert st1["tile_launches"] - st0["tile_launches"] == 2, "A and [P | A'] should both take a column-blocked kernel"
    else:
        assert st1["tile_launches"] == st0["tile_launches"]
    s.cleanup()


@pytest.mark.gpu
@pytest.mark.parametrize("case", ["basic_qp2", "random", "lasso"])
def test_ruiz_scaling_matches_reference_restatement(case):
    """work->scaling->{D, E, c} after osqp_setup (reference scale_data, src/scaling.c:44-156, 10 iterations)."""
    if case == "basic_qp2":
        p = problems.basic_qp2()
    elif case == "random":
        p = problems.random_qp(n=2000, seed=7)
    else:
        p = problems.lasso(n_feat=300, m_samp=3000, nnz_per_row=20, seed=2)
    s = Solver(p["P"], p["q"], p["A"], p["l"], p["u"])
    assert s.setup_flag == 0
    n, m = sp.csc_matrix(p["P"]).shape[0], sp.csc_matrix(p["A"]).shape[0]
    sc = s.work.contents.scaling.contents
    D = np.ctypeslib.as_array(sc.D, shape=(n,)).copy()
    E = np.ctypeslib.as_array(sc.E, shape=(m,)).copy()
    Pf = full_sym(sp.triu(sp.csc_matrix(p["P"])))
    _, _, _, _, _, D0, E0, c0 = O.scale_data(Pf, np.asarray(p["q"], float), sp.csc_matrix(p["A"]),
                                              np.asarray(p["l"], float), np.asarray(p["u"], float), 10)
    assert np.max(np.abs(D / D0 - 1.0)) <= 1e-12
    assert np.max(np.abs(E / E0 - 1.0)) <= 1e-12
    assert abs(sc.c / c0 - 1.0) <= 1e-12
    assert abs(sc.cinv * sc.c - 1.0) <= 1e-14
    s.cleanup()
