"""The two exact transformations of the reduced KKT system that round 2 added, restated in numpy and
checked against dense linear algebra (CPU; the CUDA code follows these formulas line by line:
osqp_b200/csrc/slack.cuh, osqp_b200/csrc/band.cuh)."""
import numpy as np
import scipy.sparse as sp

from osqp_b200 import problems


def _K(P, A, rho, sigma):
    Pf = sp.csc_matrix(P)
    Pf = Pf + sp.triu(Pf, 1).T if (abs(Pf - Pf.T)).nnz else Pf
    return (Pf + sigma * sp.eye(Pf.shape[0]) + A.T @ sp.diags(rho) @ A).toarray()


def test_slack_elimination_is_the_schur_complement_with_a_modified_rho():
    """slack.cuh: S = K_XX - K_XY D^-1 K_YX = P_XX + sigma I + A_X' diag(rho~) A_X and the three
    element-wise steps around the CG (k_slack_pre, the masked CG on S, the close in k_admm_update)
    solve K d = r0 exactly."""
    rng = np.random.default_rng(0)
    p = problems.lasso(n_feat=12, m_samp=40, nnz_per_row=5, seed=3)
    P, A = sp.csc_matrix(p["P"]), sp.csc_matrix(p["A"])
    n, m = A.shape[1], A.shape[0]
    sigma = 1e-6
    rho = np.where(p["l"] == p["u"], 1e3 * 0.1, 0.1) * (1 + rng.random(m))      # a different rho on every row
    K = _K(P, A, rho, sigma)
    # the slack set as slack_setup finds it: singleton column of A, diagonal-only P, first come per row
    Pd = P.toarray(); Pd = Pd + np.triu(Pd, 1).T
    Y, rowof = [], {}
    for j in range(n):
        col = A[:, j].nonzero()[0]
        if len(col) == 1 and np.count_nonzero(Pd[j]) <= 1 and np.count_nonzero(Pd[:, j]) <= 1 and col[0] not in rowof:
            rowof[col[0]] = j
            Y.append(j)
    Y = np.array(Y); X = np.setdiff1d(np.arange(n), Y)
    assert len(Y) == 40                                   # the y block of [x; y; t]
    r = np.array([A[:, j].nonzero()[0][0] for j in Y])
    a = np.array([A[r[k], Y[k]] for k in range(len(Y))])
    pjj = Pd[Y, Y]
    D = pjj + sigma + a * a * rho[r]
    assert np.allclose(np.diag(K[np.ix_(Y, Y)]), D) and np.allclose(K[np.ix_(Y, Y)], np.diag(D))
    rho_t = rho.copy(); rho_t[r] = rho[r] * (pjj + sigma) / D
    S = K[np.ix_(X, X)] - K[np.ix_(X, Y)] @ np.diag(1.0 / D) @ K[np.ix_(Y, X)]
    AX = A.toarray()[:, X]
    S2 = Pd[np.ix_(X, X)] + sigma * np.eye(len(X)) + AX.T @ np.diag(rho_t) @ AX
    assert np.allclose(S, S2, rtol=1e-12, atol=1e-12)
    # one ADMM linear solve: r0 = -(q + P x + A' v), K d = r0
    x, v, q = rng.standard_normal(n), rng.standard_normal(m), rng.standard_normal(n)
    r0 = -(q + Pd @ x + A.T @ v)
    d_ref = np.linalg.solve(K, r0)
    r0Y = -(q[Y] + pjj * x[Y] + a * v[r])                # k_slack_pre: row Y of r0 through the partner row
    assert np.allclose(r0Y, r0[Y])
    vt = v.copy(); vt[r] += rho[r] * a * r0Y / D          # ... folded into v
    r0S = -(q + Pd @ x + A.T @ vt)[X]                     # the rhs pass, rows X
    dX = np.linalg.solve(S2, r0S)                         # the CG on S (exact here)
    AdX = AX @ dX                                         # the z~ accumulator
    dY = (r0Y - a * rho[r] * AdX[r]) / D                  # close: per slack, through its partner row
    d = np.zeros(n); d[X] = dX; d[Y] = dY
    assert np.allclose(d, d_ref, rtol=1e-9, atol=1e-10)
    Ad = AdX.copy(); Ad[r] += a * dY                      # A x~ on the partner rows
    assert np.allclose(Ad, A @ d_ref, rtol=1e-9, atol=1e-10)


def test_block_tridiagonal_factor_and_its_application_by_matvecs():
    """band.cuh: with G_k = L_kk^-1, H_k = G_k L_{k,k-1}, JT_{k-1} = (L_{k,k-1} G_{k-1})', the solve is
    y_k = G_k r_k - H_k y_{k-1} (forward), x_k = G_k' y_k - JT_k x_{k+1} (backward): mat-vecs only."""
    rng = np.random.default_rng(1)
    B, nl = 6, 7
    n = B * nl
    Kd = np.zeros((n, n))
    for k in range(nl):
        M = rng.standard_normal((B, B))
        Kd[k * B:(k + 1) * B, k * B:(k + 1) * B] = M @ M.T + 4 * B * np.eye(B)
        if k:
            E = rng.standard_normal((B, B))
            Kd[k * B:(k + 1) * B, (k - 1) * B:k * B] = E
            Kd[(k - 1) * B:k * B, k * B:(k + 1) * B] = E.T
    # block Cholesky (k_band_factor)
    G, H, JT = [None] * nl, [None] * nl, [None] * nl
    Lprev = None
    for k in range(nl):
        Dk = Kd[k * B:(k + 1) * B, k * B:(k + 1) * B].copy()
        if k:
            E = Kd[k * B:(k + 1) * B, (k - 1) * B:k * B]
            Lsub = E @ G[k - 1].T                          # L_{k,k-1} = E_k L_{k-1,k-1}^-T
            Dk -= Lsub @ Lsub.T
        Lkk = np.linalg.cholesky(Dk)
        G[k] = np.linalg.inv(Lkk)
        if k:
            H[k] = G[k] @ Lsub
            JT[k - 1] = (Lsub @ G[k - 1]).T
    r = rng.standard_normal(n)
    y = np.zeros(n)
    for k in range(nl):
        y[k * B:(k + 1) * B] = G[k] @ r[k * B:(k + 1) * B] - (H[k] @ y[(k - 1) * B:k * B] if k else 0.0)
    x = np.zeros(n)
    for k in range(nl - 1, -1, -1):
        x[k * B:(k + 1) * B] = G[k].T @ y[k * B:(k + 1) * B] - (JT[k] @ x[(k + 1) * B:(k + 2) * B] if k < nl - 1 else 0.0)
    assert np.allclose(x, np.linalg.solve(Kd, r), rtol=1e-10, atol=1e-12)


def test_mpc_reduced_matrix_is_block_tridiagonal_in_bfs_level_order():
    """band_plan_compute: breadth-first level sets of the graph of K = P + sigma I + A' rho A from a
    peripheral vertex give a block-tridiagonal K (an edge joins equal or adjacent levels)."""
    p = problems.mpc(horizon=8, nx=4, nu=2, seed=3)
    P, A = sp.csc_matrix(p["P"]), sp.csc_matrix(p["A"])
    n = A.shape[1]
    K = _K(P, A, np.ones(A.shape[0]), 1e-6)
    adj = [np.nonzero(K[i])[0] for i in range(n)]

    def bfs(s):
        lvl = -np.ones(n, int); lvl[s] = 0
        cur, k = [s], 0
        while cur:
            nxt = []
            for v in cur:
                for u in adj[v]:
                    if lvl[u] < 0:
                        lvl[u] = k + 1; nxt.append(u)
            cur, k = nxt, k + 1
        return lvl
    lvl = bfs(0)
    lvl = bfs(int(np.argmax(lvl)))                         # one George-Liu sweep: restart from a farthest vertex
    assert lvl.min() == 0
    i, j = np.nonzero(K)
    assert np.max(np.abs(lvl[i] - lvl[j])) <= 1
    assert np.bincount(lvl).max() <= 3 * (4 + 2)           # levels of about one stage (nx + nu)
