"""Host-side helpers of the row-partitioned mode (include/osqp_cuda_pcg.h).

One process per GPU.  Rank g owns rows [bounds[g], bounds[g+1]) of A, l, u (and of z, y);
x, q and all scalars are replicated; P's nonzeros are spread over the ranks (rank g passes the g-th
slab of columns of triu(P): the symmetric completions of the shares sum to P, and every use of P on
the path -- operator, Jacobi diagonal, column norms, objective -- goes through the same sum / max
allreduce as the A' partial sums, so no rank carries all of P's nonzeros).  The data-path collective
(ncclAllReduce of the [P | A'] partial sums inside every CG iteration) lives in the CUDA
library; this module only cuts the problem, bootstraps NCCL and gathers the dual slabs.
The cutting logic is pure numpy/scipy so that it is testable on CPU (gloo, world_size 2).
"""
import numpy as np
import scipy.sparse as sp


def row_bounds(m, world):
    """Contiguous, balanced row slabs: bounds[g] = g * m // world."""
    return [(g * m) // world for g in range(world + 1)]


def p_share(P, rank, world):
    """Rank's share of P as an upper-triangular n x n CSC matrix: columns [n g / G, n (g + 1) / G) of
    triu(P), the others empty.  sum_g full_sym(p_share(P, g, G)) == full symmetric P."""
    Pu = sp.triu(sp.csc_matrix(P), format="csc")
    n = Pu.shape[0]
    if world <= 1:
        return Pu
    c = row_bounds(n, world)
    lo, hi = c[rank], c[rank + 1]
    indptr = np.zeros(n + 1, dtype=Pu.indptr.dtype)
    a, e = Pu.indptr[lo], Pu.indptr[hi]
    indptr[lo:hi + 1] = Pu.indptr[lo:hi + 1] - a
    indptr[hi + 1:] = e - a
    return sp.csc_matrix((Pu.data[a:e], Pu.indices[a:e], indptr), shape=(n, n))


def full_sym(Pu):
    """Symmetric completion of an upper-triangular share (what the device builds from it)."""
    Pu = sp.csc_matrix(Pu)
    return sp.csc_matrix(Pu + sp.triu(Pu, 1).T)


def slab(prob, rank, world):
    """The rank's share of a QP dict(P, q, A, l, u): its rows of A, l, u and its share of P."""
    A = sp.csr_matrix(prob["A"])
    m, n = A.shape
    b = row_bounds(m, world)
    lo, hi = b[rank], b[rank + 1]
    P = p_share(prob["P"], rank, world)
    return dict(P=P, q=np.asarray(prob["q"], dtype=np.float64),
                A=sp.csc_matrix(A[lo:hi, :]), l=np.asarray(prob["l"][lo:hi], dtype=np.float64),
                u=np.asarray(prob["u"][lo:hi], dtype=np.float64), rows=(lo, hi))


def local_operator_partial(s, x, v_local):
    """What one rank contributes to [P | A'] [x; v]: P_g x + A_g' v_g  (numpy; the check the
    CPU tests run against the un-partitioned product after a sum-allreduce)."""
    return full_sym(s["P"]) @ x + s["A"].T @ v_local


def local_col_absmax(s):
    """The rank's contribution to the Ruiz column norms of [P; A] (joined by a max-allreduce)."""
    n = s["A"].shape[1]
    out = np.zeros(n)
    for M in (full_sym(s["P"]), s["A"]):
        M = sp.csc_matrix(M)
        if M.nnz:
            out = np.maximum(out, np.asarray(abs(M).max(axis=0).todense()).ravel())
    return out


def init_partition(plugin, rank, world):
    """Create the NCCL communicator of the CUDA library: rank 0 makes the id, torch.distributed
    (any backend) ships it.  Call once per process before setup; world <= 1 leaves the mode."""
    if world <= 1:
        plugin.set_partition(0, 1, None)
        return
    import torch.distributed as dist
    box = [plugin.nccl_unique_id() if rank == 0 else None]
    dist.broadcast_object_list(box, src=0)
    plugin.set_partition(rank, world, box[0])


def gather_rows(y_local, world):
    """Concatenate the per-rank slabs of an m-vector on every rank (host side, any backend)."""
    if world <= 1:
        return np.asarray(y_local)
    import torch.distributed as dist
    parts = [None] * world
    dist.all_gather_object(parts, np.asarray(y_local))
    return np.concatenate(parts)
