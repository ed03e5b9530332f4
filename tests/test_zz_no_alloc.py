"""osqp_cuda_engine_get_solution_noalloc (include/osqp_cuda_pcg.h): the allocation-free variant of
osqp_cuda_engine_get_solution, whose staging area is a file-scope static device array.  It has no
counterpart in the reference.  Called through the
C ABI, two calls in a row, against the existing entry point; sizes beyond the static maximum are
rejected.  Collected last on purpose."""
import ctypes as C

import numpy as np
import pytest

from osqp_b200 import Solver, capi, problems

TOL = 1e-12          # the two entry points run the same un-scaling kernels: equal to rounding


def _bind(plug):
    L = plug.lib
    for name in ("osqp_cuda_engine_get_solution", "osqp_cuda_engine_get_solution_noalloc"):
        f = getattr(L, name)
        f.argtypes = [C.c_void_p, capi.c_float_p, capi.c_float_p]
        f.restype = capi.c_int
    L.osqp_cuda_noalloc_max.argtypes = []
    L.osqp_cuda_noalloc_max.restype = capi.c_int
    return L


def _read(fn, eng, n, m):
    x, y = np.full(n, np.nan), np.full(m, np.nan)
    rc = fn(eng, x.ctypes.data_as(capi.c_float_p), y.ctypes.data_as(capi.c_float_p))
    return int(rc), x, y


@pytest.mark.gpu
def test_noalloc_solution_matches_existing_entry_point_twice():
    L = _bind(capi.Plugin())
    p = problems.random_qp(n=300, seed=4)            # n + m = 3300 doubles of the static scratch
    s = Solver(p["P"], p["q"], p["A"], p["l"], p["u"], eps_abs=1e-5, eps_rel=1e-5, adaptive_rho_interval=50)
    assert s.setup_flag == 0 and s.solve() == 0 and s.info.status_val == 1
    n, m = p["P"].shape[0], p["A"].shape[0]
    assert n + m <= int(L.osqp_cuda_noalloc_max())
    eng = s.engine_handle()
    rc, x_ref, y_ref = _read(L.osqp_cuda_engine_get_solution, eng, n, m)
    assert rc == 0
    assert np.max(np.abs(x_ref - s.x)) == 0.0 and np.max(np.abs(y_ref - s.y)) == 0.0
    # a second, smaller problem in between leaves its own data in the shared static array
    p2 = problems.random_qp(n=100, seed=9)
    s2 = Solver(p2["P"], p2["q"], p2["A"], p2["l"], p2["u"], adaptive_rho_interval=50)
    assert s2.setup_flag == 0 and s2.solve() == 0
    for call in range(2):
        rc, x, y = _read(L.osqp_cuda_engine_get_solution_noalloc, eng, n, m)
        assert rc == 0, call
        assert np.max(np.abs(x - x_ref)) <= TOL * max(1.0, np.max(np.abs(x_ref))), call
        assert np.max(np.abs(y - y_ref)) <= TOL * max(1.0, np.max(np.abs(y_ref))), call
        n2, m2 = p2["P"].shape[0], p2["A"].shape[0]
        rc2, x2, y2 = _read(L.osqp_cuda_engine_get_solution_noalloc, s2.engine_handle(), n2, m2)
        assert rc2 == 0
        assert np.max(np.abs(x2 - s2.x)) <= TOL * max(1.0, np.max(np.abs(s2.x)))
        assert np.max(np.abs(y2 - s2.y)) <= TOL * max(1.0, np.max(np.abs(s2.y)))
    # x only / y only
    xo = np.full(n, np.nan)
    assert L.osqp_cuda_engine_get_solution_noalloc(eng, xo.ctypes.data_as(capi.c_float_p), None) == 0
    assert np.max(np.abs(xo - x_ref)) <= TOL * max(1.0, np.max(np.abs(x_ref)))
    s.cleanup(); s2.cleanup()


@pytest.mark.gpu
def test_noalloc_rejects_sizes_beyond_the_static_scratch():
    L = _bind(capi.Plugin())
    p = problems.random_qp(n=25000, seed=2)          # n + m = 275000 > 262144
    s = Solver(p["P"], p["q"], p["A"], p["l"], p["u"], adaptive_rho_interval=50, max_iter=50)
    assert s.setup_flag == 0
    s.solve()
    n, m = p["P"].shape[0], p["A"].shape[0]
    assert n + m > int(L.osqp_cuda_noalloc_max())
    rc, x, y = _read(L.osqp_cuda_engine_get_solution_noalloc, s.engine_handle(), n, m)
    assert rc == 2
    assert np.all(np.isnan(x)) and np.all(np.isnan(y))        # nothing was written
    rc, x, y = _read(L.osqp_cuda_engine_get_solution, s.engine_handle(), n, m)
    assert rc == 0 and np.all(np.isfinite(x))
    s.cleanup()
