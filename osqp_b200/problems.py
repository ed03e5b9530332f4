"""Synthetic QP generators for the configs of BASELINE.json.

    minimize 0.5 x'Px + q'x   subject to  l <= Ax <= u

Every generator returns a dict {P, q, A, l, u, name} with scipy CSC matrices (P full
symmetric; the C API takes its upper triangle) and numpy fp64 vectors.  Seeds are
fixed so the CPU reference and the CUDA backend see identical inputs.  The tiny
literal problems restate the reference's own fixtures (file:line cited).
"""
import numpy as np
import scipy.sparse as sp

INF = 1e30   # OSQP_INFTY, reference include/constants.h:99


def demo_qp():
    """reference examples/osqp_demo.c:6-18 (BASELINE.json config 1)."""
    P = sp.csc_matrix(np.array([[4.0, 1.0], [1.0, 2.0]]))
    A = sp.csc_matrix(np.array([[1.0, 1.0], [1.0, 0.0], [0.0, 1.0]]))
    return dict(name="demo", P=P, q=np.array([1.0, 1.0]), A=A,
                l=np.array([1.0, 0.0, 0.0]), u=np.array([1.0, 0.7, 0.7]))


def basic_qp():
    """reference tests/basic_qp/generate_problem.py:5-24."""
    P = sp.csc_matrix(np.array([[4.0, 1.0], [1.0, 2.0]]))
    A = sp.csc_matrix(np.array([[1.0, 1.0], [1.0, 0.0], [0.0, 1.0], [0.0, 1.0]]))
    return dict(name="basic_qp", P=P, q=np.ones(2), A=A, l=np.array([1.0, 0.0, 0.0, -INF]),
                u=np.array([1.0, 0.7, 0.7, INF]),
                x_test=np.array([0.3, 0.7]), y_test=np.array([-2.9, 0.0, 0.2, 0.0]), obj_test=1.88)


def basic_qp2():
    """reference tests/basic_qp2/generate_problem.py:5-30."""
    P = sp.csc_matrix(np.array([[11.0, 0.0], [0.0, 0.0]]))
    A = sp.csc_matrix(np.array([[-1.0, 0.0], [0.0, -1.0], [-1.0, 3.0], [2.0, 5.0], [3.0, 4.0]]))
    return dict(name="basic_qp2", P=P, q=np.array([3.0, 4.0]), A=A, l=-INF * np.ones(5),
                u=np.array([0.0, 0.0, -15.0, 100.0, 80.0]),
                x_test=np.array([15.0, -0.0]), y_test=np.array([0.0, 508.0, 168.0, 0.0, 0.0]),
                obj_test=1282.5,
                q_new=np.array([1.0, 1.0]), u_new=np.array([-2.0, 0.0, -20.0, 100.0, 80.0]),
                x_test_new=np.array([20.0, -0.0]), y_test_new=np.array([0.0, 664.0, 221.0, 0.0, 0.0]),
                obj_test_new=2220.0)


def random_qp(n=100000, m_factor=10, nnz_per_row=10, seed=1, p_cols=3):
    """BASELINE.json configs 2 and 5: A has `nnz_per_row` N(0,1) entries per
    row at uniform random columns; P = M M' + 1e-2 I with M n x n, `p_cols` N(0,1) entries
    per column; q ~ N(0,1); u = A v + U(0,1), l = -inf."""
    rng = np.random.default_rng(seed)
    m = int(m_factor * n)
    cols = rng.integers(0, n, size=(m, nnz_per_row), dtype=np.int64)
    vals = rng.standard_normal((m, nnz_per_row))
    indptr = np.arange(0, m * nnz_per_row + 1, nnz_per_row, dtype=np.int64)
    A = sp.csr_matrix((vals.ravel(), cols.ravel(), indptr), shape=(m, n))
    A.sum_duplicates()
    A = A.tocsc()
    A.sort_indices()
    mrows = rng.integers(0, n, size=(n, p_cols), dtype=np.int64)
    mvals = rng.standard_normal((n, p_cols))
    M = sp.csc_matrix((mvals.ravel(), mrows.ravel(),
                       np.arange(0, n * p_cols + 1, p_cols, dtype=np.int64)), shape=(n, n))
    M.sum_duplicates()
    P = (M @ M.T + 1e-2 * sp.identity(n, format="csc")).tocsc()
    P.sort_indices()
    q = rng.standard_normal(n)
    v = rng.standard_normal(n)
    u = A @ v + rng.random(m)
    l = -INF * np.ones(m)
    return dict(name=f"random_qp_n{n}_m{m}", P=P, q=q, A=A, l=l, u=u)


def random_qp_rows(n=1000000, m_factor=10, nnz_per_row=20, seed=1, rows=None, nblocks=64, p_cols=3):
    """The same family as random_qp, generated in `nblocks` independent row blocks so that a rank
    of the row-partitioned mode can build ONLY its slab [rows[0], rows[1]) of A, l, u, and every
    world size sees the same global problem (BASELINE.json config 5: n=1e6, m=1e7, 20 nnz/row).
    The slab boundaries must fall on block boundaries (m / nblocks rows per block).
    Returns dict(P, q, A (slab, CSC), l, u (slab), rows, m)."""
    m = int(m_factor * n)
    assert m % nblocks == 0, "m must be a multiple of nblocks"
    rb = m // nblocks
    lo, hi = rows if rows is not None else (0, m)
    assert lo % rb == 0 and hi % rb == 0, "slab boundaries must be block boundaries"
    rng0 = np.random.default_rng([seed, 0])
    mrows = rng0.integers(0, n, size=(n, p_cols), dtype=np.int64)
    mvals = rng0.standard_normal((n, p_cols))
    M = sp.csc_matrix((mvals.ravel(), mrows.ravel(),
                       np.arange(0, n * p_cols + 1, p_cols, dtype=np.int64)), shape=(n, n))
    M.sum_duplicates()
    P = (M @ M.T + 1e-2 * sp.identity(n, format="csc")).tocsc()
    P.sort_indices()
    q = rng0.standard_normal(n)
    v = rng0.standard_normal(n)
    blocks, us = [], []
    for b in range(lo // rb, hi // rb):
        rng = np.random.default_rng([seed, 1 + b])
        cols = rng.integers(0, n, size=(rb, nnz_per_row), dtype=np.int32)
        vals = rng.standard_normal((rb, nnz_per_row))
        Ab = sp.csr_matrix((vals.ravel(), cols.ravel(),
                            np.arange(0, rb * nnz_per_row + 1, nnz_per_row, dtype=np.int64)), shape=(rb, n))
        Ab.sum_duplicates()
        us.append(Ab @ v + rng.random(rb))
        blocks.append(Ab)
    A = sp.vstack(blocks, format="csr") if blocks else sp.csr_matrix((0, n))
    A = A.tocsc()
    A.sort_indices()
    u = np.concatenate(us) if us else np.zeros(0)
    return dict(name=f"random_qp_n{n}_m{m}_rows{lo}_{hi}", P=P, q=q, A=A, l=-INF * np.ones(hi - lo), u=u,
                rows=(lo, hi), m=m)


def lasso(n_feat=100000, m_samp=1000000, nnz_per_row=50, seed=1, gamma_frac=0.2):
    """BASELINE.json config 3; formulation of reference docs/examples/lasso.rst:40-63:
    variables (x, y, t), minimize y'y + gamma 1't  s.t.  y = Ad x - b, -t <= x <= t."""
    rng = np.random.default_rng(seed)
    nf, ms = int(n_feat), int(m_samp)
    cols = rng.integers(0, nf, size=(ms, nnz_per_row), dtype=np.int64)
    vals = rng.standard_normal((ms, nnz_per_row))
    Ad = sp.csr_matrix((vals.ravel(), cols.ravel(),
                        np.arange(0, ms * nnz_per_row + 1, nnz_per_row, dtype=np.int64)),
                       shape=(ms, nf))
    Ad.sum_duplicates()
    x_true = (rng.random(nf) > 0.8) * rng.standard_normal(nf) / np.sqrt(nf)
    b = Ad @ x_true + 0.5 * rng.standard_normal(ms)
    gamma = gamma_frac * np.max(np.abs(Ad.T @ b))
    In, Im = sp.identity(nf, format="csc"), sp.identity(ms, format="csc")
    P = sp.block_diag([sp.csc_matrix((nf, nf)), 2.0 * Im, sp.csc_matrix((nf, nf))], format="csc")
    q = np.hstack([np.zeros(nf + ms), gamma * np.ones(nf)])
    A = sp.vstack([sp.hstack([Ad, -Im, sp.csc_matrix((ms, nf))]),
                   sp.hstack([In, sp.csc_matrix((nf, ms)), -In]),
                   sp.hstack([In, sp.csc_matrix((nf, ms)), In])], format="csc")
    A.sort_indices()
    l = np.hstack([b, -INF * np.ones(nf), np.zeros(nf)])
    u = np.hstack([b, np.zeros(nf), INF * np.ones(nf)])
    return dict(name=f"lasso_n{nf}_m{ms}", P=P, q=q, A=A, l=l, u=u, gamma=gamma)


def mpc(horizon=50, nx=20, nu=10, seed=1, x0=None):
    """BASELINE.json config 4; structure of reference docs/examples/mpc.rst:81-101:
    x = (x(0..N), u(0..N-1)); dynamics as equalities, boxes on states and inputs."""
    rng = np.random.default_rng(seed)
    N = int(horizon)
    Ad = rng.standard_normal((nx, nx))
    Ad *= 0.95 / np.max(np.abs(np.linalg.eigvals(Ad)))
    Bd = rng.standard_normal((nx, nu))
    Q, QN, R = sp.identity(nx), sp.identity(nx), 0.1 * sp.identity(nu)
    if x0 is None:
        x0 = rng.standard_normal(nx)
    P = sp.block_diag([sp.kron(sp.identity(N), Q), QN, sp.kron(sp.identity(N), R)], format="csc")
    q = np.zeros((N + 1) * nx + N * nu)
    Ax = sp.kron(sp.identity(N + 1), -sp.identity(nx)) + sp.kron(sp.eye(N + 1, k=-1), sp.csc_matrix(Ad))
    Bu = sp.kron(sp.vstack([sp.csc_matrix((1, N)), sp.identity(N)]), sp.csc_matrix(Bd))
    Aeq = sp.hstack([Ax, Bu])
    leq = np.hstack([-x0, np.zeros(N * nx)])
    Aineq = sp.identity((N + 1) * nx + N * nu)
    lineq = np.hstack([-5.0 * np.ones((N + 1) * nx), -1.0 * np.ones(N * nu)])
    A = sp.vstack([Aeq, Aineq], format="csc")
    A.sort_indices()
    l = np.hstack([leq, lineq])
    u = np.hstack([leq, -lineq])
    return dict(name=f"mpc_N{N}_nx{nx}_nu{nu}", P=P, q=q, A=A, l=l, u=u, nx=nx, nu=nu, N=N,
                Ad=Ad, Bd=Bd)


def algorithmic_bytes(prob):
    """Algorithmic HBM bytes of the device kernels for this problem:
    fp64 values + int32 indices = 12 B per stored nonzero, 4 B per row pointer, every
    vector element counted once per kernel that must touch it."""
    P, A = sp.csc_matrix(prob["P"]), sp.csc_matrix(prob["A"])
    n, m = P.shape[0], A.shape[0]
    nnzPf, nnzA = P.nnz, A.nnz

    def spmv(nnz, rows, cols):
        return 12.0 * nnz + 4.0 * (rows + 1) + 8.0 * cols + 8.0 * rows

    bA = spmv(nnzA, m, n)
    bM = spmv(nnzPf + nnzA, n, n + m)           # [Pf | A'] in one pass
    pcg_iter = (bA + 8.0 * m) + (bM + 8.0 * n) + 8.0 * 7 * n + 8.0 * 4 * n
    # rhs pass over [Pf | A'] + the ADMM update, which is a pure vector kernel: z~ = A x~ comes from the
    # recurrence inside the CG (no pass over A), 14 m-vector and 4 n-vector reads / writes
    admm_fixed = (bM + 8.0 * 5 * n) + (8.0 * 4 * n + 8.0 * 14 * m)
    check = (bA + 8.0 * n + 8.0 * 11 * m) + (bM + 8.0 * (n + m) + 8.0 * 5 * n)
    return dict(n=n, m=m, nnzA=nnzA, nnzPf=nnzPf, spmv_A=bA, spmv_M=bM, pcg_iter=pcg_iter,
                admm_fixed=admm_fixed, check=check)


# ------------------------------------------------------------------ on-disk problems
# One binary file per QP for the C runner osqp_b200/csrc/osqp_run.c (format in its header);
# the reference only has text dumps of single objects (src/util.c:393-472).
def save_problem(path, P, q, A, l, u):
    Pu = sp.triu(sp.csc_matrix(P), format="csc"); Pu.sort_indices()
    Ac = sp.csc_matrix(A); Ac.sort_indices()
    n, m = Pu.shape[0], Ac.shape[0]
    with open(path, "wb") as f:
        f.write(b"OSQPB200")
        np.array([n, m, Pu.nnz, Ac.nnz], dtype="<i8").tofile(f)
        for M in (Pu, Ac):
            M.indptr.astype("<i8").tofile(f)
            M.indices.astype("<i8").tofile(f)
            M.data.astype("<f8").tofile(f)
        for v in (q, l, u):
            np.asarray(v, dtype="<f8").tofile(f)


def load_problem(path):
    with open(path, "rb") as f:
        if f.read(8) != b"OSQPB200":
            raise ValueError(f"{path} is not an OSQPB200 problem file")
        n, m, nnzP, nnzA = (int(v) for v in np.fromfile(f, "<i8", 4))
        mats = []
        for nnz, rows in ((nnzP, n), (nnzA, m)):
            p = np.fromfile(f, "<i8", n + 1); i = np.fromfile(f, "<i8", nnz); x = np.fromfile(f, "<f8", nnz)
            mats.append(sp.csc_matrix((x, i, p), shape=(rows, n)))
        q, l, u = np.fromfile(f, "<f8", n), np.fromfile(f, "<f8", m), np.fromfile(f, "<f8", m)
    return dict(P=mats[0], q=q, A=mats[1], l=l, u=u)


def load_solution(path):
    with open(path, "rb") as f:
        if f.read(8) != b"OSQPSOL0":
            raise ValueError(f"{path} is not an OSQPSOL0 solution file")
        n, m, status_val, it = (int(v) for v in np.fromfile(f, "<i8", 4))
        obj, pri, dua = np.fromfile(f, "<f8", 3)
        return dict(status_val=status_val, iter=it, obj_val=float(obj), pri_res=float(pri), dua_res=float(dua),
                    x=np.fromfile(f, "<f8", n), y=np.fromfile(f, "<f8", m))
