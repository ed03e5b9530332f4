import json
import os
import sys

import numpy as np
import pytest
import scipy.sparse as sp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "oracle")):
    if p not in sys.path:
        sys.path.insert(0, p)

REF_LIB = os.path.join(ROOT, "oracle", "_ref", "libosqp_ref.so")
GOLDEN = os.path.join(ROOT, "tests", "golden")
TESTS_TOL = 1e-4          # reference tests/osqp_tester.h:9


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a B200 (run with -m gpu on the GPU box)")


def _dec(v):
    if isinstance(v, dict) and v.get("__csc__"):
        return sp.csc_matrix((np.array(v["x"], float), np.array(v["i"], np.int64),
                              np.array(v["p"], np.int64)), shape=tuple(v["shape"]))
    if isinstance(v, dict) and v.get("__nd__"):
        return np.array(v["data"], dtype=v["dtype"]).reshape(v["shape"])
    return v


@pytest.fixture(scope="session")
def fixtures():
    """The reference's own test fixtures (tests/golden/make_fixtures.py)."""
    with open(os.path.join(GOLDEN, "reference_fixtures.json")) as f:
        raw = json.load(f)
    return {k: {kk: _dec(vv) for kk, vv in d.items()} for k, d in raw.items()}


@pytest.fixture(scope="session")
def ref_outputs():
    with open(os.path.join(GOLDEN, "ref_outputs.json")) as f:
        return json.load(f)


@pytest.fixture(scope="session")
def ref_lib():
    if not os.path.exists(REF_LIB):
        pytest.skip("oracle/_ref/libosqp_ref.so not built (make -C oracle ref needs /root/reference)")
    return REF_LIB


def full_sym(Pu):
    Pu = sp.csc_matrix(Pu)
    return sp.csc_matrix(sp.triu(Pu) + sp.triu(Pu, 1).T)
