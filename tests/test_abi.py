"""CPU-only checks of the drop-in boundary: both shared libraries load, export every symbol
that include/*.h declares, agree on struct layouts with the reference ABI, validate inputs
like the reference does, and FAIL LOUDLY (no CPU fallback) when no B200 is present."""
import ctypes as C
import os
import re
import shutil
import subprocess

import numpy as np
import pytest

from conftest import ROOT
from osqp_b200 import Solver, capi, problems


def _declared(header):
    src = open(os.path.join(ROOT, "include", header)).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    names = re.findall(r"^\s*(?:c_int|c_float|void|csc|void\s*\*|csc\s*\*)\s*\*?\s*(\w+)\s*\(", src, flags=re.M)
    return sorted(set(names))


def test_host_library_exports_osqp_h():
    lib = C.CDLL(capi.PRODUCT_LIB)
    names = _declared("osqp.h")
    assert "osqp_setup" in names and "osqp_update_P_A" in names and "init_linsys_solver" in names
    assert len(names) >= 35
    for n in names:
        assert hasattr(lib, n), n
    assert C.c_char_p.in_dll(lib, "LINSYS_SOLVER_NAME") is not None


def test_plugin_exports_osqp_cuda_pcg_h():
    lib = C.CDLL(capi.PLUGIN_LIB)
    names = _declared("osqp_cuda_pcg.h")
    for must in ("init_linsys_solver_cuda_pcg", "solve_linsys_cuda_pcg",
                 "update_linsys_solver_matrices_cuda_pcg", "update_linsys_solver_rho_vec_cuda_pcg",
                 "free_linsys_solver_cuda_pcg", "osqp_cuda_engine_iterate", "osqp_cuda_engine_residuals"):
        assert must in names
    for n in names:
        assert hasattr(lib, n), n


def test_struct_layouts_match_c(tmp_path):
    """sizeof/offsetof as seen by gcc for include/osqp.h == the ctypes mirror (== reference ABI)."""
    if not shutil.which("gcc"):
        pytest.skip("gcc missing")
    src = tmp_path / "lay.c"
    src.write_text('#include <stdio.h>\n#include <stddef.h>\n#include "osqp.h"\n'
                   'int main(){printf("%zu %zu %zu %zu %zu %zu %zu %zu %zu\\n", sizeof(csc), sizeof(OSQPSettings),'
                   'sizeof(OSQPInfo), sizeof(OSQPData), offsetof(OSQPWorkspace, settings),'
                   'offsetof(OSQPWorkspace, info), offsetof(OSQPWorkspace, summary_printed),'
                   'offsetof(OSQPSettings, linsys_solver), offsetof(OSQPInfo, rho_estimate));return 0;}\n')
    exe = tmp_path / "lay"
    subprocess.check_call(["gcc", "-I", os.path.join(ROOT, "include"), str(src), "-o", str(exe)])
    got = [int(t) for t in subprocess.check_output([str(exe)]).split()]
    W, S, I = capi.OSQPWorkspace, capi.OSQPSettings, capi.OSQPInfo
    want = [C.sizeof(capi.csc), C.sizeof(S), C.sizeof(I), C.sizeof(capi.OSQPData), W.settings.offset,
            W.info.offset, W.summary_printed.offset, S.linsys_solver.offset, I.rho_estimate.offset]
    assert got == want


def test_default_settings_match_reference_constants():
    """reference include/constants.h:58-118"""
    lib = capi.load_library()
    s = capi.OSQPSettings()
    lib.osqp_set_default_settings(C.byref(s))
    assert (s.rho, s.sigma, s.scaling, s.max_iter, s.eps_abs, s.eps_rel) == (0.1, 1e-6, 10, 4000, 1e-3, 1e-3)
    assert (s.eps_prim_inf, s.eps_dual_inf, s.alpha, s.delta) == (1e-4, 1e-4, 1.6, 1e-6)
    assert (s.polish, s.polish_refine_iter, s.check_termination, s.warm_start) == (0, 3, 25, 1)
    assert (s.adaptive_rho, s.adaptive_rho_interval, s.adaptive_rho_tolerance, s.adaptive_rho_fraction) == (1, 0, 5.0, 0.4)
    assert s.linsys_solver == capi.CUDA_PCG_SOLVER and s.time_limit == 0.0


BAD_SETTINGS = [dict(scaling=-1), dict(adaptive_rho=2), dict(adaptive_rho_interval=-1),
                dict(adaptive_rho_fraction=-1.5), dict(adaptive_rho_tolerance=0.5),
                dict(polish_refine_iter=-1), dict(rho=-1.0), dict(sigma=-0.1), dict(delta=-1.0),
                dict(max_iter=-1), dict(eps_abs=-1.0), dict(eps_rel=-1.0), dict(eps_abs=0.0, eps_rel=0.0),
                dict(eps_prim_inf=-0.1), dict(eps_dual_inf=-0.1), dict(alpha=2.0), dict(linsys_solver=5),
                dict(verbose=2), dict(scaled_termination=2), dict(check_termination=-1),
                dict(warm_start=5), dict(time_limit=-0.2)]


@pytest.mark.parametrize("bad", BAD_SETTINGS)
def test_settings_validation_errors(bad):
    """reference tests/basic_qp/test_basic_qp.h:113-290 -> OSQP_SETTINGS_VALIDATION_ERROR,
    raised before any device work (so it is checkable without a GPU)."""
    p = problems.basic_qp()
    s = Solver(p["P"], p["q"], p["A"], p["l"], p["u"], **bad)
    assert s.setup_flag == capi.OSQP_SETTINGS_VALIDATION_ERROR


def test_data_validation_errors():
    """reference tests/basic_qp/test_basic_qp.h:292-376 -> OSQP_DATA_VALIDATION_ERROR"""
    import scipy.sparse as sp
    p = problems.basic_qp()
    # l > u
    s = Solver(p["P"], p["q"], p["A"], p["u"] + 1.0, p["u"])
    assert s.setup_flag == capi.OSQP_DATA_VALIDATION_ERROR
    # A with the wrong number of columns
    s = Solver(p["P"], p["q"], sp.csc_matrix(np.ones((4, 3))), p["l"], p["u"])
    assert s.setup_flag == capi.OSQP_DATA_VALIDATION_ERROR
    # P not upper triangular: hand the full matrix through the raw holder
    Pfull = capi.CscHolder(sp.csc_matrix(np.array([[4.0, 1.0], [1.0, 2.0]])), upper=False)
    s = Solver(Pfull, p["q"], p["A"], p["l"], p["u"])
    assert s.setup_flag == capi.OSQP_DATA_VALIDATION_ERROR
    # non-square P
    Pbad = capi.CscHolder(sp.csc_matrix(np.ones((2, 3))), upper=False)
    s = Solver(Pbad, p["q"], p["A"], p["l"], p["u"])
    assert s.setup_flag == capi.OSQP_DATA_VALIDATION_ERROR


def _have_gpu():
    try:
        import torch
        return torch.cuda.is_available()
    except Exception:
        return False


@pytest.mark.skipif(_have_gpu(), reason="checks the no-GPU failure mode")
def test_no_cpu_fallback():
    """Without a usable B200 osqp_setup must fail with OSQP_LINSYS_SOLVER_LOAD_ERROR."""
    p = problems.demo_qp()
    s = Solver(p["P"], p["q"], p["A"], p["l"], p["u"])
    assert s.setup_flag == capi.OSQP_LINSYS_SOLVER_LOAD_ERROR


def test_product_does_not_link_oracle():
    """Neither product library references oracle/ or carries the reference's factorisation."""
    for lib in (capi.PRODUCT_LIB, capi.PLUGIN_LIB):
        blob = open(lib, "rb").read()
        assert b"libosqp_ref" not in blob and b"osqp_oracle" not in blob
        for sym in (b"QDLDL_factor", b"QDLDL_solve", b"amd_l_order", b"form_KKT"):
            assert sym not in blob, sym
