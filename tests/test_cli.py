"""osqp_run (osqp_b200/csrc/osqp_run.c): on-disk problems through the C API without Python (SURVEY 8f
rank 4).  The runner only uses the reference's own entry points and dlopen()s the library it is given,
so the CPU tests drive the compiled reference with it (QDLDL) and the GPU test the product library --
same binary, same file, same settings: the ABI is the drop-in boundary."""
import json
import os
import subprocess

import numpy as np
import pytest

from conftest import REF_LIB, ROOT
from osqp_b200 import problems

RUN = os.path.join(ROOT, "osqp_b200", "lib", "osqp_run")


def _run(args):
    r = subprocess.run([RUN] + args, stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True, timeout=300)
    return r.returncode, [json.loads(l) for l in r.stdout.splitlines() if l.startswith("{")], r.stderr


def test_problem_file_round_trip_and_header(tmp_path):
    p = problems.random_qp(n=40, seed=5)
    f = str(tmp_path / "p.bin")
    problems.save_problem(f, p["P"], p["q"], p["A"], p["l"], p["u"])
    back = problems.load_problem(f)
    import scipy.sparse as sp
    assert (sp.triu(p["P"]) != back["P"]).nnz == 0 and (sp.csc_matrix(p["A"]) != back["A"]).nnz == 0
    for k in ("q", "l", "u"):
        assert np.array_equal(np.asarray(p[k], float), back[k])
    rc, out, err = _run([f, "--info"])
    assert rc == 0, err
    assert out[0]["n"] == 40 and out[0]["m"] == 400 and out[0]["nnzA"] == sp.csc_matrix(p["A"]).nnz
    bad = str(tmp_path / "bad.bin")
    open(bad, "wb").write(b"not a problem")
    assert _run([bad, "--info"])[0] == 2
    open(bad, "wb").write(open(f, "rb").read()[:200])
    assert _run([bad, "--lib", "/nonexistent.so"])[0] == 2          # truncated file: caught before the library
    assert _run([f, "--lib", "/nonexistent.so"])[0] == 3


@pytest.mark.skipif(not os.path.exists(REF_LIB), reason="needs oracle/_ref (built where /root/reference exists)")
def test_runner_drives_the_reference_library(tmp_path, fixtures):
    """basic_qp of reference tests/basic_qp/generate_problem.py through the runner + the compiled reference"""
    d = fixtures["basic_qp"]
    f, o = str(tmp_path / "basic.bin"), str(tmp_path / "basic.sol")
    from conftest import full_sym
    problems.save_problem(f, full_sym(d["P"]), d["q"], d["A"], d["l"], d["u"])
    rc, out, err = _run([f, "--lib", REF_LIB, "--solver", "0", "--eps", "1e-5", "--polish", "1", "--out", o])
    assert rc == 0, err
    assert out[1]["status_val"] == 1
    sol = problems.load_solution(o)
    assert np.max(np.abs(sol["x"] - d["sol_x_test"])) < 1e-4 and np.max(np.abs(sol["y"] - d["sol_y_test"])) < 1e-4
    assert abs(sol["obj_val"] - d["sol_obj_value_test"]) < 1e-4 and sol["iter"] == out[1]["iter"]


@pytest.mark.gpu
def test_runner_drives_the_product_library(tmp_path):
    import osqp_oracle as O
    p = problems.random_qp(n=300, seed=8)
    f, o = str(tmp_path / "p.bin"), str(tmp_path / "p.sol")
    problems.save_problem(f, p["P"], p["q"], p["A"], p["l"], p["u"])
    rc, out, err = _run([f, "--eps", "1e-5", "--out", o])            # default library: libosqp_b200.so, cuda pcg
    assert rc == 0, err
    ref = O.solve(p["P"], p["q"], p["A"], p["l"], p["u"], eps_abs=1e-5, eps_rel=1e-5, adaptive_rho_interval=50)
    sol = problems.load_solution(o)
    assert sol["status_val"] == ref.status_val == 1
    assert abs(sol["obj_val"] - ref.obj_val) <= 1e-4 * max(1.0, abs(ref.obj_val))
    assert np.max(np.abs(sol["x"] - ref.x)) <= 1e-3 * max(1.0, np.max(np.abs(ref.x)))
