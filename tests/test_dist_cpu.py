"""Host-side logic of the row-partitioned mode on CPU: two gloo ranks cut a QP
into row slabs and must reproduce the un-partitioned [P | A'] product and Ruiz column norms
after the same collectives the CUDA library issues (sum / max allreduce of n doubles)."""
import os
import socket

import numpy as np
import pytest
import scipy.sparse as sp

from osqp_b200 import dist as qd
from osqp_b200 import problems


def test_row_bounds_tile_the_rows():
    for m in (1, 2, 7, 1000, 1000003):
        for world in (1, 2, 3, 8):
            b = qd.row_bounds(m, world)
            assert b[0] == 0 and b[-1] == m and len(b) == world + 1
            sizes = np.diff(b)
            assert (sizes >= 0).all() and sizes.max() - sizes.min() <= 1


def test_slabs_partition_the_problem():
    p = problems.random_qp(n=40, seed=3)
    world = 3
    slabs = [qd.slab(p, g, world) for g in range(world)]
    A = sp.vstack([s["A"] for s in slabs]).toarray()
    assert np.array_equal(A, p["A"].toarray())
    assert np.array_equal(np.concatenate([s["l"] for s in slabs]), p["l"])
    assert np.array_equal(np.concatenate([s["u"] for s in slabs]), p["u"])
    # P: upper-triangular shares (column slabs of triu(P)) whose symmetric completions sum to P
    assert all(s["P"].shape == (40, 40) and sp.tril(s["P"], -1).nnz == 0 for s in slabs)
    assert sum(s["P"].nnz for s in slabs) == sp.triu(sp.csc_matrix(p["P"])).nnz
    assert max(s["P"].nnz for s in slabs) < sp.triu(sp.csc_matrix(p["P"])).nnz      # nobody carries all of it
    assert abs(sum(qd.full_sym(s["P"]) for s in slabs) - sp.csc_matrix(p["P"])).max() == 0.0
    assert all(np.array_equal(s["q"], p["q"]) for s in slabs)


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, out):
    import torch
    import torch.distributed as dist
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        p = problems.random_qp(n=60, seed=5)
        m, n = p["A"].shape
        rng = np.random.default_rng(11)          # same stream on both ranks = replicated vectors
        x, v = rng.standard_normal(n), rng.standard_normal(m)
        s = qd.slab(p, rank, world)
        lo, hi = s["rows"]
        # the exchange inside every CG iteration: partial [P | A'] [x; v] -> allreduce(sum)
        t = torch.from_numpy(qd.local_operator_partial(s, x, v[lo:hi]).copy())
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
        full = sp.csc_matrix(p["P"]) @ x + p["A"].T @ v
        err_op = float(np.max(np.abs(t.numpy() - full)))
        # the exchange of a Ruiz pass: column inf-norms of [P; A] -> allreduce(max)
        c = torch.from_numpy(qd.local_col_absmax(s).copy())
        dist.all_reduce(c, op=dist.ReduceOp.MAX)
        ref = np.maximum(np.asarray(abs(sp.csc_matrix(p["P"])).max(axis=0).todense()).ravel(),
                         np.asarray(abs(p["A"]).max(axis=0).todense()).ravel())
        err_norm = float(np.max(np.abs(c.numpy() - ref)))
        # the dual slabs come back in row order
        y = qd.gather_rows(v[lo:hi], world)
        ok_gather = bool(np.array_equal(y, v))
        # every rank ends up with identical bits (replicated state stays replicated)
        bits = [None] * world
        dist.all_gather_object(bits, t.numpy().tobytes())
        out.put((rank, err_op, err_norm, ok_gather, all(b == bits[0] for b in bits)))
    finally:
        dist.destroy_process_group()


@pytest.mark.timeout(240)
def test_two_rank_gloo_exchange_matches_unpartitioned():
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    out = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, out)) for r in range(2)]
    for pr in procs:
        pr.start()
    res = [out.get(timeout=200) for _ in procs]
    for pr in procs:
        pr.join(timeout=60)
        assert pr.exitcode == 0
    for rank, err_op, err_norm, ok_gather, same_bits in res:
        assert err_op < 1e-12, (rank, err_op)
        assert err_norm == 0.0, (rank, err_norm)
        assert ok_gather and same_bits
