"""Row-partitioned mode on real GPUs (needs >= 2): the partitioned solve must reproduce the
single-GPU solve of the same QP (status, objective, solution) -- tools/dist_check.py."""
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.gpu
@pytest.mark.timeout(900)
@pytest.mark.parametrize("polish", [0, 1])
def test_partitioned_matches_single_gpu(polish):
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2",
           "--master-addr", "127.0.0.1", "--master-port", str(29533 + polish),
           os.path.join(ROOT, "tools", "dist_check.py"), "3000", "1e-4", str(polish)]
    r = subprocess.run(cmd, cwd=ROOT, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=800)
    assert r.returncode == 0, r.stdout[-3000:]
    assert r.stdout.count("ok=True") == 2, r.stdout[-3000:]


@pytest.mark.gpu
@pytest.mark.parametrize("polish", [0, 1])
def test_partitioned_code_path_on_one_gpu(polish):
    """The row-partitioned code path in a world of one (osqp_cuda_set_partition_loopback): every
    [P | A'] product stores its partial row sums, joins them (identity here) and runs its fused
    epilogue in the separate kernel (launch_reduced / k_apply_epi); P arrives as a share.  Same
    arithmetic per row as the ordinary solve, so status, iterations and solution must agree
    (dot products are grouped differently: rounding-level differences only)."""
    import ctypes as C
    import numpy as np
    from osqp_b200 import Solver, capi, problems
    from osqp_b200 import dist as qd
    plug = capi.Plugin()
    L = plug.lib
    L.osqp_cuda_set_partition_loopback.argtypes = [capi.c_int]
    L.osqp_cuda_set_partition_loopback.restype = capi.c_int
    p = problems.random_qp(n=3000, seed=7)
    kw = dict(eps_abs=1e-4, eps_rel=1e-4, adaptive_rho_interval=50, polish=polish)
    s1 = Solver(p["P"], p["q"], p["A"], p["l"], p["u"], **kw)
    assert s1.setup_flag == 0 and s1.solve() == 0
    assert L.osqp_cuda_set_partition_loopback(1) == 0
    try:
        sl = qd.slab(p, 0, 1)
        s = Solver(sl["P"], sl["q"], sl["A"], sl["l"], sl["u"], **kw)
        assert s.setup_flag == 0 and s.solve() == 0
        part = plug.get_partition()
        assert part["allreduces"] > 0, "the partitioned code path did not run"
        assert s.info.status_val == s1.info.status_val == 1
        assert s.info.status_polish == s1.info.status_polish
        assert abs(int(s.info.iter) - int(s1.info.iter)) <= 25
        assert abs(s.info.obj_val - s1.info.obj_val) <= 1e-6 * max(1.0, abs(s1.info.obj_val))
        assert np.max(np.abs(s.x - s1.x)) <= 1e-5 * max(1.0, np.max(np.abs(s1.x)))
        assert np.max(np.abs(s.y - s1.y)) <= 1e-4 * max(1.0, np.max(np.abs(s1.y)))
        s.cleanup()
    finally:
        L.osqp_cuda_set_partition_loopback(0)
    s1.cleanup()
