"""The column-blocked shared-memory-gather mirrors (osqp_b200/csrc/spmv_tile.cuh) against the
sliced-ELL kernel (OSQP_B200_TILE=0): same algorithm, the layout changes the summation order
inside a row (per column block, then over the blocks), so status and iteration count must match
and the solutions agree to rounding.  The mirrors are only built for large random-sparsity
problems (>= 1M nonzeros, enough entries per row and column block), hence the sizes here.  The
choice is read once per process, hence the subprocesses.  Kernel-level checks against scipy:
tests/test_gpu_linalg.py."""
import json
import os
import subprocess
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

SCRIPT = r"""
import sys, json, numpy as np, scipy.sparse as sp
sys.path.insert(0, %r)
from osqp_b200 import Solver, problems
case, out = sys.argv[1], sys.argv[2]
if case == "random":        # config-2 shape at 0.4 scale
    p = problems.random_qp(n=40000, seed=11)
    kw = dict(eps_abs=1e-4, eps_rel=1e-4)
elif case == "random_polish":
    p = problems.random_qp(n=33000, seed=5)
    kw = dict(eps_abs=1e-5, eps_rel=1e-5, polish=1, max_iter=20000)
elif case == "lasso":       # only A gets a mirror (ROWS mode, empty phases skipped); slack elimination on
    p = problems.lasso(n_feat=5000, m_samp=50000, nnz_per_row=30, seed=4)
    kw = dict(eps_abs=1e-4, eps_rel=1e-4, max_iter=20000)
else:
    raise SystemExit("unknown case")
s = Solver(p["P"], p["q"], p["A"], p["l"], p["u"], adaptive_rho_interval=50, **kw)
assert s.setup_flag == 0
assert s.solve() == 0
st = s.stats()
np.savez(out, x=s.x, y=s.y)
print(json.dumps(dict(status=int(s.info.status_val), iters=int(s.info.iter), obj=float(s.info.obj_val),
                      polish=int(s.info.status_polish), tile=int(st["tile_launches"]), sell=int(st["sell_launches"]),
                      launches=int(st["kernel_launches"]), solve_time=float(s.info.solve_time))))
"""


def _run(case, tile, tmp_path):
    env = dict(os.environ)
    env["OSQP_B200_TILE"] = "1" if tile else "0"
    env["OSQP_B200_TILE_ROWS_MAXPCT"] = "400"        # the ROWS kernel on random sparsity too (default bar: 60 %)
    out = str(tmp_path / f"{case}_{tile}.npz")
    r = subprocess.run([sys.executable, "-c", SCRIPT % ROOT, case, out], cwd=ROOT, env=env,
                       stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True, timeout=800)
    assert r.returncode == 0, r.stderr[-2000:]
    info = json.loads(r.stdout.strip().splitlines()[-1])
    z = np.load(out)
    return info, z["x"], z["y"]


@pytest.mark.gpu
@pytest.mark.timeout(1800)
@pytest.mark.parametrize("case", ["random", "random_polish", "lasso"])
def test_tile_and_sell_kernels_agree(case, tmp_path):
    a, xa, ya = _run(case, True, tmp_path)
    b, xb, yb = _run(case, False, tmp_path)
    print(a, b)
    assert a["tile"] > 0, "the column-blocked kernels did not run"
    assert b["tile"] == 0 and b["sell"] > 0
    assert a["status"] == b["status"] == 1, (a, b)
    assert a["polish"] == b["polish"]
    # inexact CG + a different summation order: the trajectories may part by a check interval
    assert abs(a["iters"] - b["iters"]) <= 50, (a, b)
    assert abs(a["obj"] - b["obj"]) <= 1e-6 * max(1.0, abs(b["obj"])), (a, b)
    assert np.max(np.abs(xa - xb)) <= 1e-4 * max(1.0, np.max(np.abs(xb)))
    assert np.max(np.abs(ya - yb)) <= 1e-3 * max(1.0, np.max(np.abs(yb)))


@pytest.mark.gpu
@pytest.mark.timeout(900)
def test_tile_ragged_kkt_solve_satisfies_the_system(monkeypatch):
    """A linear solve has one answer whatever the kernel: the strict plugin call (reference
    lin_sys/direct/qdldl/qdldl_interface.c:350-376, host b in/out) on a tile-sized matrix with
    empty rows, empty columns, 2-entry rows next to 900-entry rows and a different rho on every
    row.  The KKT system is too large for a sparse LU here, so the check is the system itself,
    evaluated in numpy: (P + sigma I + A' rho A) x~ = b1 + A' (rho .* b2) and z~ = A x~."""
    import ctypes as C
    import scipy.sparse as sp
    from osqp_b200 import capi
    from test_gpu_plugin import _plugin
    monkeypatch.setenv("OSQP_B200_TILE_ROWS_MAXPCT", "400")    # both mirrors wanted here (default bar for ROWS: 60 %)
    rng = np.random.default_rng(5)
    n, m = 60000, 300000
    rows = np.repeat(np.arange(m), 16)                   # enough stream per row range for the ROWS mirror to pay
    cols = rng.integers(0, n, size=16 * m)               # for its table loads (tile_build's traffic rule)
    vals = rng.standard_normal(16 * m)
    keep = (rows % 7) != 0                               # every 7th row empty
    keep &= ~((cols >= 100) & (cols < 140))              # empty columns
    long_rows = np.array([40, 41, 5000, m - 1])          # 900-entry rows inside blocks of short ones
    lr = np.repeat(long_rows, 900)
    lc = np.concatenate([rng.choice(n, 900, replace=False) for _ in long_rows])
    A = sp.csc_matrix((np.concatenate([vals[keep], rng.standard_normal(lr.size)]),
                       (np.concatenate([rows[keep], lr]), np.concatenate([cols[keep], lc]))), shape=(m, n))
    A.sum_duplicates(); A.sort_indices()
    Mx = sp.random(n, n, density=2.0 / n, random_state=8, format="csc")
    P = sp.csc_matrix(sp.triu(Mx @ Mx.T + 1e-2 * sp.eye(n)))
    P.sort_indices()
    sigma = 1e-3
    rho_vec = 10.0 ** rng.uniform(-2, 0, m)
    flag, s, keep_alive = _plugin(P, A, sigma, rho_vec)
    assert flag == 0
    plug = capi.Plugin()
    plug.lib.cuda_pcg_set_tolerance.argtypes = [C.c_void_p, capi.c_float, capi.c_int]
    plug.lib.cuda_pcg_set_tolerance.restype = None
    plug.lib.cuda_pcg_set_tolerance(C.cast(s, C.c_void_p), 1e-12, 5000)
    plug.lib.cuda_pcg_uses_tile.argtypes = [C.c_void_p]
    plug.lib.cuda_pcg_uses_tile.restype = capi.c_int
    assert plug.lib.cuda_pcg_uses_tile(C.cast(s, C.c_void_p)) == 3, "the column-blocked mirrors were not built"
    Pf = sp.csc_matrix(P + sp.triu(P, 1).T)
    for trial in range(2):
        rhs = rng.standard_normal(n + m)
        b = rhs.copy()
        assert s.contents.solve(s, b.ctypes.data_as(capi.c_float_p)) == 0
        xt, zt = b[:n], b[n:]
        want = rhs[:n] + A.T @ (rho_vec * rhs[n:])
        got = Pf @ xt + sigma * xt + A.T @ (rho_vec * (A @ xt))
        assert np.max(np.abs(got - want)) <= 1e-8 * max(1.0, np.max(np.abs(want)))
        assert np.max(np.abs(zt - A @ xt)) <= 1e-10 * max(1.0, np.max(np.abs(zt)))
    s.contents.free(s)
