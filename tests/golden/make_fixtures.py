"""Regenerates tests/golden/reference_fixtures.json from the reference's OWN fixture
generators (reference tests/*/generate_problem.py), run here in the build container with
`utils.codegen_utils` replaced by a recorder: instead of printing C headers
(tests/utils/codegen_utils.py:172-461) the two entry points capture the problem data and
the golden answers.  /root/reference cannot travel to the GPU box; the JSON does.

    python tests/golden/make_fixtures.py        # needs /root/reference
"""
import json
import os
import runpy
import sys
import types

import numpy as np
import scipy.sparse as sp

REF_TESTS = "/root/reference/tests"
OUT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "reference_fixtures.json")
CASES = ["basic_qp", "basic_qp2", "lin_alg", "non_cvx", "primal_dual_infeasibility",
         "primal_infeasibility", "solve_linsys", "unconstrained", "update_matrices"]

captured = {}


def enc(v):
    if sp.issparse(v):
        M = sp.csc_matrix(v)
        M.sort_indices()
        return {"__csc__": True, "shape": list(M.shape), "p": M.indptr.tolist(),
                "i": M.indices.tolist(), "x": [float(t) for t in M.data]}
    if isinstance(v, np.ndarray):
        a = np.where(np.isposinf(v), 1e30, np.where(np.isneginf(v), -1e30, v)) if v.dtype.kind == "f" else v
        return {"__nd__": True, "shape": list(v.shape), "dtype": str(v.dtype), "data": a.ravel().tolist()}
    if isinstance(v, (np.floating, float)):
        return float(v)
    if isinstance(v, (np.integer, int)):
        return int(v)
    if isinstance(v, str):
        return v
    raise TypeError(type(v))


def generate_problem_data(P, q, A, l, u, problem_name, sols_data={}):
    d = {"P": P, "q": np.asarray(q, float), "A": A, "l": np.asarray(l, float), "u": np.asarray(u, float)}
    d.update({"sol_" + k: v for k, v in sols_data.items()})
    captured[problem_name] = {k: enc(v) for k, v in d.items()}


def generate_data(problem_name, sols_data):
    captured[problem_name] = {k: enc(v) for k, v in sols_data.items()}


def main():
    stub = types.ModuleType("utils.codegen_utils")
    stub.generate_problem_data = generate_problem_data
    stub.generate_data = generate_data
    pkg = types.ModuleType("utils")
    pkg.codegen_utils = stub
    sys.modules["utils"] = pkg
    sys.modules["utils.codegen_utils"] = stub
    # scipy >= 1.12 dropped the numpy re-exports this generator used (sp.randn)
    import scipy
    if not hasattr(scipy, "randn"):
        scipy.randn = np.random.randn
    cwd = os.getcwd()
    os.chdir(REF_TESTS)
    try:
        for c in CASES:
            runpy.run_path(os.path.join(REF_TESTS, c, "generate_problem.py"), run_name="__main__")
    finally:
        os.chdir(cwd)
    with open(OUT, "w") as f:
        json.dump(captured, f)
    print("wrote", OUT, {k: len(v) for k, v in captured.items()})


if __name__ == "__main__":
    main()
