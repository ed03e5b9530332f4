"""The sliced-ELL mirror (osqp_b200/csrc/spmv_sell.cuh) against the direct CSR kernel
(OSQP_B200_SPMV=direct): same algorithm, the layout only changes the summation order inside a
row, so status and iteration count must match and the solutions agree to rounding.  The kernel
choice is read once per process, hence the subprocesses."""
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
if case == "random":        # short uniform rows in A (1 lane per row), ~100-entry rows in [P | A']
    p = problems.random_qp(n=3000, seed=11)
elif case == "lasso":       # ragged: 50-entry sample rows next to 2-entry bound rows, equality rows
    p = problems.lasso(n_feat=600, m_samp=6000, nnz_per_row=50, seed=3)
else:
    raise SystemExit("unknown case")
s = Solver(p["P"], p["q"], p["A"], p["l"], p["u"], eps_abs=1e-5, eps_rel=1e-5, adaptive_rho_interval=50,
           polish=1, max_iter=20000)
assert s.setup_flag == 0
assert s.solve() == 0
st = s.stats()
np.savez(out, x=s.x, y=s.y)
print(json.dumps(dict(status=int(s.info.status_val), iters=int(s.info.iter), obj=float(s.info.obj_val),
                      polish=int(s.info.status_polish), sell=int(st["sell_launches"]), launches=int(st["kernel_launches"]))))
"""


def _run(case, kind, tmp_path):
    env = dict(os.environ)
    env.pop("OSQP_B200_SPMV", None)
    if kind:
        env["OSQP_B200_SPMV"] = kind
    out = str(tmp_path / f"{case}_{kind or 'sell'}.npz")
    r = subprocess.run([sys.executable, "-c", SCRIPT % ROOT, case, out], cwd=ROOT, env=env,
                       stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True, timeout=500)
    assert r.returncode == 0, r.stderr[-2000:]
    info = json.loads(r.stdout.strip().splitlines()[-1])
    z = np.load(out)
    return info, z["x"], z["y"]


@pytest.mark.gpu
@pytest.mark.timeout(900)
@pytest.mark.parametrize("case", ["random", "lasso"])
def test_sell_and_direct_kernels_agree(case, tmp_path):
    a, xa, ya = _run(case, None, tmp_path)
    b, xb, yb = _run(case, "direct", tmp_path)
    assert a["sell"] > 0, "the sliced-ELL kernel did not run"
    assert b["sell"] == 0
    assert a["status"] == b["status"] == 1, (a, b)
    assert a["polish"] == b["polish"]
    # inexact CG + a different summation order: the trajectories may part by a check interval
    assert abs(a["iters"] - b["iters"]) <= 50, (a, b)
    assert abs(a["obj"] - b["obj"]) <= 1e-6 * max(1.0, abs(b["obj"])), (a, b)
    assert np.max(np.abs(xa - xb)) <= 1e-4 * max(1.0, np.max(np.abs(xb)))
    assert np.max(np.abs(ya - yb)) <= 1e-3 * max(1.0, np.max(np.abs(yb)))


@pytest.mark.gpu
def test_sell_ragged_kkt_solve_matches_exact_solve():
    """A linear solve has one answer whatever the trajectory: the strict plugin call on a matrix
    with empty rows, empty columns, 4-entry rows next to 700-entry rows (several lanes-per-row
    values in one mirror) and a different rho on every row, against the oracle's sparse LU."""
    import ctypes as C
    import scipy.sparse as sp
    import osqp_oracle as O
    from osqp_b200 import capi
    from test_gpu_plugin import _plugin
    rng = np.random.default_rng(5)
    n, m = 2500, 9000
    A = sp.random(m, n, density=4.0 / n, random_state=7, format="lil")
    for r in range(0, m, 7):
        A.rows[r] = []; A.data[r] = []
    for r in (40, 41, 5000, 8999):                      # long rows inside blocks of short ones
        cols = np.sort(rng.choice(n, 700, replace=False))
        A.rows[r] = list(cols); A.data[r] = list(rng.standard_normal(700))
    A = sp.csc_matrix(A.tocsc())
    A[:, 100:140] = 0.0                                 # empty columns
    A.eliminate_zeros()
    M = sp.random(n, n, density=2.0 / n, random_state=8, format="csc")
    P = sp.csc_matrix(M @ M.T + 1e-2 * sp.eye(n))
    sigma = 1e-3
    rho_vec = 10.0 ** rng.uniform(-2, 1, m)
    flag, s, keep = _plugin(P, A, sigma, rho_vec)
    assert flag == 0
    plug = capi.Plugin()
    plug.lib.cuda_pcg_set_tolerance.argtypes = [C.c_void_p, capi.c_float, capi.c_int]
    plug.lib.cuda_pcg_set_tolerance.restype = None
    plug.lib.cuda_pcg_set_tolerance(C.cast(s, C.c_void_p), 1e-13, 5000)
    for trial in range(2):
        rhs = rng.standard_normal(n + m)
        want = O.solve_kkt(P, A, sigma, rho_vec, rhs)
        b = rhs.copy()
        assert s.contents.solve(s, b.ctypes.data_as(capi.c_float_p)) == 0
        assert np.max(np.abs(b - want)) <= 1e-6 * max(1.0, np.max(np.abs(want)))
    s.contents.free(s)
