"""Manual prototype behind osqp_b200/csrc/slack.cuh (not collected by pytest): Jacobi-preconditioned CG
iterations on the reduced KKT matrix K of a Lasso problem against the Schur complement S obtained by
eliminating the slack block (K_YY diagonal).  Uses the oracle's scaling / rho vectors.
    python tests/manual_lasso_schur_proto.py
"""
import sys, numpy as np, scipy.sparse as sp, scipy.sparse.linalg as spla
import os
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, 'oracle'))
from osqp_b200 import problems
import osqp_oracle as O
from scipy.sparse.csgraph import reverse_cuthill_mckee

def pcg(K, b, Minv, tol=1e-2, maxit=2000):
    x = np.zeros_like(b); r = b.copy(); z = Minv(r); p = z.copy(); rz = r @ z; r0 = np.max(np.abs(r))
    for it in range(1, maxit + 1):
        Kp = K @ p; a = rz / (p @ Kp); x += a * p; r -= a * Kp
        if np.max(np.abs(r)) <= tol * r0: return it
        z = Minv(r); rz2 = r @ z; p = z + (rz2 / rz) * p; rz = rz2
    return maxit

def build_K(p, rho=0.1, sigma=1e-6):
    Pf = sp.csc_matrix(p["P"]); A = sp.csc_matrix(p["A"])
    Ps, qs, As, ls, us, D, E, c = O.scale_data(Pf, p["q"], A, p["l"], p["u"], 10)
    rho_vec = np.asarray(O.rho_vectors(ls, us, rho)[1], dtype=float)
    assert rho_vec.max() > rho_vec.min() or True
    K = (Ps + sigma * sp.eye(Ps.shape[0]) + As.T @ sp.diags(rho_vec) @ As).tocsc()
    return K, rho_vec

def block_jacobi(K, bs):
    n = K.shape[0]; Kd = K.toarray(); inv = []
    for a in range(0, n, bs):
        e = min(n, a + bs); inv.append((a, e, np.linalg.inv(Kd[a:e, a:e])))
    def f(r):
        z = np.empty_like(r)
        for a, e, Bi in inv: z[a:e] = Bi @ r[a:e]
        return z
    return f

rng = np.random.default_rng(0)
for kw in (dict(n_feat=200, m_samp=2000, nnz_per_row=20), dict(n_feat=2000, m_samp=20000, nnz_per_row=50)):
    p = problems.lasso(seed=1, **kw)
    for rho in (0.1, 1.0):
        K, rho_vec = build_K(p, rho=rho)
        n = K.shape[0]
        A = sp.csc_matrix(p["A"]); Pf = sp.csc_matrix(p["P"])
        # Y: variables whose column of A has exactly one entry, in an equality row, and whose P column is diagonal only
        ctype = O.rho_vectors(p["l"], p["u"], rho)[0]
        cnt = np.diff(A.indptr)
        Pcnt = np.diff(Pf.indptr)
        isY = np.zeros(n, bool)
        for j in np.where(cnt == 1)[0]:
            r = A.indices[A.indptr[j]]
            if ctype[r] == 1 and (Pcnt[j] == 0 or (Pcnt[j] == 1 and Pf.indices[Pf.indptr[j]] == j)): isY[j] = True
        # at most one Y per row
        seen = set(); 
        for j in np.where(isY)[0]:
            r = A.indices[A.indptr[j]]
            if r in seen: isY[j] = False
            seen.add(r)
        Y = np.where(isY)[0]; X = np.where(~isY)[0]
        Kxx = K[X][:, X].tocsc(); Kxy = K[X][:, Y].tocsc(); Kyy = K[Y][:, Y].tocsc()
        offd = abs(Kyy - sp.diags(Kyy.diagonal())).max()
        Dy = Kyy.diagonal()
        S = (Kxx - Kxy @ sp.diags(1.0 / Dy) @ Kxy.T).tocsc()
        b = rng.standard_normal(n)
        d = K.diagonal()
        it_K = pcg(K, b, lambda r: r / d)
        bs = b[X] - Kxy @ (b[Y] / Dy)
        ds = S.diagonal()
        it_S = pcg(S, bs, lambda r: r / ds)
        print(kw, "rho", rho, "n", n, "|Y|", len(Y), "Kyy offdiag", offd, "PCG(K, Jacobi):", it_K, " PCG(S, Jacobi):", it_S, flush=True)
