"""CPU oracle: a numpy/scipy restatement of the reference's ADMM path.

TEST INFRASTRUCTURE ONLY.  Nothing under osqp_b200/ imports this file; only tests/,
__graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may.

Each function cites the reference file:line it follows.  The KKT system
[P + sigma I, A'; A, -diag(1/rho)] is solved exactly with scipy's sparse LU (the same
call the reference's own fixture generator uses for its golden answers,
tests/solve_linsys/generate_problem.py:30-33) in place of QDLDL's LDL'.

Pinned (tests/test_oracle.py): against the reference's golden vectors restated in
tests/golden/reference_fixtures.json (basic_qp, basic_qp2, unconstrained,
primal/dual infeasibility, update_matrices, solve_linsys) and against outputs of the
reference itself compiled by oracle/Makefile (tests/golden/ref_outputs.json, generated
by tests/golden/make_ref_outputs.py): same status, iteration count and objective.
"""
from types import SimpleNamespace

import numpy as np
import scipy.sparse as sp
import scipy.sparse.linalg as spla

# reference include/constants.h:17-29,58-121
OSQP_SOLVED, OSQP_SOLVED_INACCURATE = 1, 2
OSQP_PRIMAL_INFEASIBLE, OSQP_PRIMAL_INFEASIBLE_INACCURATE = -3, 3
OSQP_DUAL_INFEASIBLE, OSQP_DUAL_INFEASIBLE_INACCURATE = -4, 4
OSQP_MAX_ITER_REACHED, OSQP_NON_CVX, OSQP_UNSOLVED = -2, -7, -10
OSQP_INFTY = 1e30
OSQP_NAN = float(0x7FC00000)
OSQP_DIVISION_TOL = 1.0 / OSQP_INFTY
RHO_MIN, RHO_MAX, RHO_TOL, RHO_EQ_OVER_RHO_INEQ = 1e-6, 1e6, 1e-4, 1e3
MIN_SCALING, MAX_SCALING = 1e-4, 1e4

DEFAULTS = dict(rho=0.1, sigma=1e-6, scaling=10, adaptive_rho=1, adaptive_rho_interval=0,
                adaptive_rho_tolerance=5.0, adaptive_rho_fraction=0.4, max_iter=4000,
                eps_abs=1e-3, eps_rel=1e-3, eps_prim_inf=1e-4, eps_dual_inf=1e-4, alpha=1.6,
                delta=1e-6, polish=0, polish_refine_iter=3, verbose=0, scaled_termination=0,
                check_termination=25, warm_start=1, time_limit=0.0, linsys_solver=0)


def _norm_inf(v):
    return float(np.max(np.abs(v))) if v.size else 0.0


def _limit_scaling(d):
    """reference src/scaling.c:7-14"""
    d = np.where(d < MIN_SCALING, 1.0, d)
    return np.where(d > MAX_SCALING, MAX_SCALING, d)


def _col_norms(M):
    """column inf-norms of a sparse matrix (reference src/lin_alg.c:325-343)"""
    M = sp.csc_matrix(M)
    out = np.zeros(M.shape[1])
    if M.nnz:
        a = abs(M).tocsc()
        out = np.asarray(a.max(axis=0).todense()).ravel()
    return out


def scale_data(P, q, A, l, u, iters):
    """Ruiz equilibration + cost normalisation, reference src/scaling.c:44-156.
    P full symmetric.  Returns scaled copies and D, E, c."""
    n, m = P.shape[0], A.shape[0]
    P, A, q = sp.csc_matrix(P, dtype=float).copy(), sp.csc_matrix(A, dtype=float).copy(), q.astype(float).copy()
    D, E, c = np.ones(n), np.ones(m), 1.0
    for _ in range(int(iters)):
        # column norms of [P A'; A 0]   (scaling.c:28-42)
        Dt = np.maximum(_col_norms(P), _col_norms(A))
        Et = _col_norms(A.T) if m else np.zeros(0)
        Dt = 1.0 / np.sqrt(_limit_scaling(Dt))
        Et = 1.0 / np.sqrt(_limit_scaling(Et))
        P = sp.diags(Dt) @ P @ sp.diags(Dt)
        A = sp.diags(Et) @ A @ sp.diags(Dt) if m else A
        q = Dt * q
        D, E = D * Dt, E * Et
        # cost normalisation (scaling.c:107-141)
        c_temp = float(np.mean(_col_norms(P)))
        nq = float(_limit_scaling(np.array([_norm_inf(q)]))[0])
        c_temp = float(_limit_scaling(np.array([max(c_temp, nq)]))[0])
        c_temp = 1.0 / c_temp
        P, q, c = P * c_temp, q * c_temp, c * c_temp
    return sp.csc_matrix(P), q, sp.csc_matrix(A), E * l, E * u, D, E, c


def rho_vectors(l, u, rho):
    """reference src/auxil.c:76-98 (set_rho_vec)"""
    big = OSQP_INFTY * MIN_SCALING
    loose = (l < -big) & (u > big)
    eq = (~loose) & (u - l < RHO_TOL)
    ctype = np.where(loose, -1, np.where(eq, 1, 0))
    rho_vec = np.where(loose, RHO_MIN, np.where(eq, RHO_EQ_OVER_RHO_INEQ * rho, rho))
    return ctype, rho_vec.astype(float)


class KKT:
    """Exact solve of the quasi-definite KKT system, the role of
    reference lin_sys/direct/qdldl/qdldl_interface.c:177-376."""

    def __init__(self, P, A, sigma, rho_vec, polish=False):
        self.P, self.A, self.sigma, self.polish = P, A, sigma, polish
        self.n, self.m = P.shape[0], A.shape[0]
        self.update_rho_vec(rho_vec)

    def update_rho_vec(self, rho_vec):
        n, m = self.n, self.m
        if self.polish:                       # (2,2) block -delta I, qdldl_interface.c:261-272
            self.rho_inv = np.full(m, self.sigma)
        else:
            self.rho_inv = 1.0 / rho_vec
        K = sp.bmat([[self.P + self.sigma * sp.identity(n), self.A.T if m else None],
                     [self.A if m else None, -sp.diags(self.rho_inv) if m else None]], format="csc")
        self.lu = spla.splu(K)
        # inertia check of the reference (qdldl_interface.c:93-98): exactly n positive pivots
        if n + m <= 2000:
            ev = np.linalg.eigvalsh(K.toarray())
            self.positive = int(np.sum(ev > 0))
        else:
            self.positive = n

    def solve(self, b):
        """ADMM mode: in [sigma x_prev - q ; z_prev - y/rho], out [x~ ; z~] (qdldl_interface.c:350-376)"""
        sol = self.lu.solve(b)
        if self.polish:
            return sol
        out = b.copy()
        out[:self.n] = sol[:self.n]
        out[self.n:] = b[self.n:] + self.rho_inv * sol[self.n:]
        return out


class Workspace:
    """State of one solver instance (reference OSQPWorkspace, include/types.h:182-288)."""

    def __init__(self, P, q, A, l, u, **settings):
        s = dict(DEFAULTS)
        for k, v in settings.items():
            if k not in s:
                raise KeyError(k)
            s[k] = v
        self.s = SimpleNamespace(**s)
        P = sp.csc_matrix(P, dtype=float)
        P = sp.triu(P) + sp.triu(P, 1).T          # full symmetric from (possibly) upper
        self.n, self.m = P.shape[0], sp.csc_matrix(A).shape[0]
        self.P0, self.A0 = sp.csc_matrix(P), sp.csc_matrix(A, dtype=float).reshape(self.m, self.n)
        self.q0, self.l0, self.u0 = np.asarray(q, float), np.asarray(l, float), np.asarray(u, float)
        self._scale()
        self.s.rho = min(max(self.s.rho, RHO_MIN), RHO_MAX)
        self.ctype, self.rho_vec = rho_vectors(self.l, self.u, self.s.rho)
        self.kkt = KKT(self.P, self.A, self.s.sigma, self.rho_vec)
        self.setup_flag = 0 if self.kkt.positive == self.n else 5   # OSQP_NONCVX_ERROR
        self.x, self.z, self.y = np.zeros(self.n), np.zeros(self.m), np.zeros(self.m)
        self.rho_updates = 0

    def _scale(self):
        if self.s.scaling:
            (self.P, self.q, self.A, self.l, self.u, self.D, self.E, self.c) = scale_data(
                self.P0, self.q0, self.A0, self.l0, self.u0, self.s.scaling)
        else:
            self.P, self.q, self.A = self.P0, self.q0.copy(), self.A0
            self.l, self.u = self.l0.copy(), self.u0.copy()
            self.D, self.E, self.c = np.ones(self.n), np.ones(self.m), 1.0
        self.Dinv, self.Einv, self.cinv = 1.0 / self.D, 1.0 / self.E, 1.0 / self.c

    # ---- residuals, reference src/auxil.c:227-359
    def _unscaled(self):
        return bool(self.s.scaling) and not self.s.scaled_termination

    def residuals(self, x, z, y):
        r = SimpleNamespace()
        r.Ax, r.Px = self.A @ x, self.P @ x
        r.Aty = self.A.T @ y if self.m else np.zeros(self.n)
        r.pri_vec = r.Ax - z
        r.dua_vec = self.q + r.Px + r.Aty
        if self._unscaled():
            r.pri_res = _norm_inf(self.Einv * r.pri_vec)
            r.dua_res = self.cinv * _norm_inf(self.Dinv * r.dua_vec)
        else:
            r.pri_res, r.dua_res = _norm_inf(r.pri_vec), _norm_inf(r.dua_vec)
        if self.m == 0:
            r.pri_res = 0.0
        r.obj = (0.5 * float(x @ r.Px) + float(self.q @ x)) * (self.cinv if self.s.scaling else 1.0)
        return r

    def _tolerances(self, r, z, eps_abs, eps_rel):
        if self._unscaled():
            pri = max(_norm_inf(self.Einv * z), _norm_inf(self.Einv * r.Ax))
            dua = self.cinv * max(_norm_inf(self.Dinv * self.q), _norm_inf(self.Dinv * r.Aty),
                                  _norm_inf(self.Dinv * r.Px))
        else:
            pri = max(_norm_inf(z), _norm_inf(r.Ax))
            dua = max(_norm_inf(self.q), _norm_inf(r.Aty), _norm_inf(r.Px))
        return eps_abs + eps_rel * pri, eps_abs + eps_rel * dua

    def _is_primal_infeasible(self, eps):
        """reference src/auxil.c:361-424 (projects delta_y in place)"""
        big = OSQP_INFTY * MIN_SCALING
        dy = self.delta_y
        up_inf, lo_inf = self.u > big, self.l < -big
        dy[up_inf & lo_inf] = 0.0
        only_up = up_inf & ~lo_inf
        dy[only_up] = np.minimum(dy[only_up], 0.0)
        only_lo = lo_inf & ~up_inf
        dy[only_lo] = np.maximum(dy[only_lo], 0.0)
        nrm = _norm_inf(self.E * dy) if self._unscaled() else _norm_inf(dy)
        if nrm > OSQP_DIVISION_TOL:
            lhs = float(self.u @ np.maximum(dy, 0.0) + self.l @ np.minimum(dy, 0.0))
            if lhs < eps * nrm:
                Atdy = self.A.T @ dy
                if self._unscaled():
                    Atdy = self.Dinv * Atdy
                return _norm_inf(Atdy) < eps * nrm
        return False

    def _is_dual_infeasible(self, eps):
        """reference src/auxil.c:426-512"""
        big = OSQP_INFTY * MIN_SCALING
        dx = self.delta_x
        if self._unscaled():
            nrm, cs = _norm_inf(self.D * dx), self.c
        else:
            nrm, cs = _norm_inf(dx), 1.0
        if nrm > OSQP_DIVISION_TOL and float(self.q @ dx) < cs * eps * nrm:
            Pdx = self.P @ dx
            if self._unscaled():
                Pdx = self.Dinv * Pdx
            if _norm_inf(Pdx) < cs * eps * nrm:
                Adx = self.A @ dx
                if self._unscaled():
                    Adx = self.Einv * Adx
                bad = ((self.u < big) & (Adx > eps * nrm)) | ((self.l > -big) & (Adx < -eps * nrm))
                return not bool(np.any(bad))
        return False

    def _check_termination(self, info, r, approximate):
        """reference src/auxil.c:681-786; returns True when a status was set"""
        s = self.s
        eps_abs, eps_rel, epi, edi = s.eps_abs, s.eps_rel, s.eps_prim_inf, s.eps_dual_inf
        if info.pri_res > OSQP_INFTY or info.dua_res > OSQP_INFTY:
            info.status_val, info.obj_val = OSQP_NON_CVX, OSQP_NAN
            return True
        if approximate:
            eps_abs, eps_rel, epi, edi = 10 * eps_abs, 10 * eps_rel, 10 * epi, 10 * edi
        eps_prim, eps_dual = self._tolerances(r, self.z, eps_abs, eps_rel)
        prim_ok = dual_ok = prim_inf = dual_inf = False
        if self.m == 0 or info.pri_res < eps_prim:
            prim_ok = True
        else:
            prim_inf = self._is_primal_infeasible(epi)
        if info.dua_res < eps_dual:
            dual_ok = True
        else:
            dual_inf = self._is_dual_infeasible(edi)
        if prim_ok and dual_ok:
            info.status_val = OSQP_SOLVED_INACCURATE if approximate else OSQP_SOLVED
            return True
        if prim_inf:
            info.status_val = OSQP_PRIMAL_INFEASIBLE_INACCURATE if approximate else OSQP_PRIMAL_INFEASIBLE
            if self._unscaled():
                self.delta_y = self.E * self.delta_y
            info.obj_val = OSQP_INFTY
            return True
        if dual_inf:
            info.status_val = OSQP_DUAL_INFEASIBLE_INACCURATE if approximate else OSQP_DUAL_INFEASIBLE
            if self._unscaled():
                self.delta_x = self.D * self.delta_x
            info.obj_val = -OSQP_INFTY
            return True
        return False

    def _rho_estimate(self, r):
        """reference src/auxil.c:13-52 (scaled residuals and norms)"""
        pri = _norm_inf(r.pri_vec) if self.m else 0.0
        dua = _norm_inf(r.dua_vec)
        pri /= (max(_norm_inf(self.z), _norm_inf(r.Ax)) if self.m else 0.0) + OSQP_DIVISION_TOL
        dua /= max(_norm_inf(self.q), _norm_inf(r.Aty), _norm_inf(r.Px)) + OSQP_DIVISION_TOL
        est = self.s.rho * np.sqrt(pri / dua) if dua > 0 else RHO_MAX
        return float(min(max(est, RHO_MIN), RHO_MAX))

    def update_rho(self, rho_new):
        """reference src/osqp.c:1281-1332"""
        self.s.rho = min(max(rho_new, RHO_MIN), RHO_MAX)
        self.rho_vec = np.where(self.ctype == 0, self.s.rho,
                                np.where(self.ctype == 1, RHO_EQ_OVER_RHO_INEQ * self.s.rho, self.rho_vec))
        self.kkt.update_rho_vec(self.rho_vec)

    # ---- the ADMM loop, reference src/osqp.c:288-654 + src/auxil.c:161-225
    def solve(self, adaptive_interval_if_auto=None):
        s, n, m = self.s, self.n, self.m
        info = SimpleNamespace(iter=0, status_val=OSQP_UNSOLVED, obj_val=0.0, pri_res=0.0,
                               dua_res=0.0, rho_updates=0, status_polish=0, rho_estimate=s.rho)
        if not s.warm_start:
            self.x[:], self.z[:], self.y[:] = 0.0, 0.0, 0.0
        interval = s.adaptive_rho_interval
        if s.adaptive_rho and not interval:
            # the reference picks this from wall-clock (osqp.c:459-485); the oracle is
            # deterministic and uses the non-PROFILING rule (osqp.c:486-498)
            interval = adaptive_interval_if_auto or (4 * s.check_termination if s.check_termination else 100)
        self.delta_x, self.delta_y = np.zeros(n), np.zeros(m)
        can_check, r, it = False, None, 0
        for it in range(1, s.max_iter + 1):
            x_prev, z_prev = self.x, self.z
            rho_inv = 1.0 / self.rho_vec
            rhs = np.concatenate([s.sigma * x_prev - self.q, z_prev - rho_inv * self.y])   # auxil.c:161-175
            xz = self.kkt.solve(rhs)
            xt, zt = xz[:n], xz[n:]
            self.x = s.alpha * xt + (1 - s.alpha) * x_prev                                 # auxil.c:185-198
            self.delta_x = self.x - x_prev
            zr = s.alpha * zt + (1 - s.alpha) * z_prev
            self.z = np.minimum(np.maximum(zr + rho_inv * self.y, self.l), self.u)         # auxil.c:200-212
            self.delta_y = self.rho_vec * (zr - self.z)                                    # auxil.c:214-225
            self.y = self.y + self.delta_y
            can_check = bool(s.check_termination) and it % s.check_termination == 0
            r = None
            if can_check:
                r = self.residuals(self.x, self.z, self.y)
                info.iter, info.pri_res, info.dua_res = it, r.pri_res, r.dua_res
                if self._check_termination(info, r, False):
                    break
            if s.adaptive_rho and interval and it % interval == 0:
                if r is None:
                    r = self.residuals(self.x, self.z, self.y)
                    info.iter, info.pri_res, info.dua_res = it, r.pri_res, r.dua_res
                est = self._rho_estimate(r)
                info.rho_estimate = est
                if est > s.rho * s.adaptive_rho_tolerance or est < s.rho / s.adaptive_rho_tolerance:
                    self.update_rho(est)
                    info.rho_updates += 1
        if not can_check:
            r = self.residuals(self.x, self.z, self.y)
            info.iter, info.pri_res, info.dua_res = it, r.pri_res, r.dua_res
            self._check_termination(info, r, False)
        if r is None:
            r = self.residuals(self.x, self.z, self.y)
        if info.status_val not in (OSQP_PRIMAL_INFEASIBLE, OSQP_DUAL_INFEASIBLE, OSQP_NON_CVX):
            info.obj_val = r.obj
        if info.status_val == OSQP_UNSOLVED:
            if not self._check_termination(info, r, True):
                info.status_val = OSQP_MAX_ITER_REACHED
        info.rho_estimate = self._rho_estimate(r)
        if s.polish and info.status_val == OSQP_SOLVED:
            self._polish(info)
        # store_solution, reference src/auxil.c:524-562
        if info.status_val in (OSQP_PRIMAL_INFEASIBLE, OSQP_PRIMAL_INFEASIBLE_INACCURATE,
                               OSQP_DUAL_INFEASIBLE, OSQP_DUAL_INFEASIBLE_INACCURATE, OSQP_NON_CVX):
            info.x, info.y = np.full(n, OSQP_NAN), np.full(m, OSQP_NAN)
            if info.status_val in (OSQP_PRIMAL_INFEASIBLE, OSQP_PRIMAL_INFEASIBLE_INACCURATE):
                self.delta_y = self.delta_y / _norm_inf(self.delta_y)
            if info.status_val in (OSQP_DUAL_INFEASIBLE, OSQP_DUAL_INFEASIBLE_INACCURATE):
                self.delta_x = self.delta_x / _norm_inf(self.delta_x)
            self.x, self.z, self.y = np.zeros(n), np.zeros(m), np.zeros(m)
        else:
            info.x = self.D * self.x                     # scaling.c:177-192
            info.y = self.cinv * self.E * self.y
        info.delta_x, info.delta_y = self.delta_x, self.delta_y
        info.rho = s.rho
        return info

    def _polish(self, info):
        """reference src/polish.c:212-350"""
        s, n = self.s, self.n
        low = self.z - self.l < -self.y               # polish.c:34 (lower-active)
        upp = self.u - self.z < self.y                # polish.c:43 (upper-active; disjoint from low as l <= u)
        idx_low, idx_upp = np.where(low)[0], np.where(upp)[0]
        Ared = sp.vstack([self.A[idx_low], self.A[idx_upp]], format="csc") if (len(idx_low) + len(idx_upp)) \
            else sp.csc_matrix((0, n))
        mred = Ared.shape[0]
        kkt = KKT(self.P, Ared, s.delta, None, polish=True)
        rhs = np.concatenate([-self.q, self.l[idx_low], self.u[idx_upp]])
        sol = kkt.solve(rhs)
        K = sp.bmat([[self.P, Ared.T if mred else None], [Ared if mred else None, None]], format="csc") \
            if mred else self.P
        for _ in range(s.polish_refine_iter):          # polish.c:134-181
            sol = sol + kkt.solve(rhs - K @ sol)
        px = sol[:n]
        pz = self.A @ px
        py = np.zeros(self.m)
        py[idx_low] = sol[n:n + len(idx_low)]
        py[idx_upp] = sol[n + len(idx_low):]
        t = pz + py                                    # proj.c:16-29
        pz = np.minimum(np.maximum(t, self.l), self.u)
        py = t - pz
        r = self.residuals(px, pz, py)
        ok = ((r.pri_res < info.pri_res and r.dua_res < info.dua_res) or
              (r.pri_res < info.pri_res and info.dua_res < 1e-10) or
              (r.dua_res < info.dua_res and info.pri_res < 1e-10))
        if ok:
            info.obj_val, info.pri_res, info.dua_res, info.status_polish = r.obj, r.pri_res, r.dua_res, 1
            self.x, self.z, self.y = px, pz, py
        else:
            info.status_polish = -1

    # ---- updates, reference src/osqp.c:765-1279
    def update_lin_cost(self, q_new):
        self.q0 = np.asarray(q_new, float)
        self.q = self.c * self.D * self.q0 if self.s.scaling else self.q0.copy()

    def update_bounds(self, l_new, u_new):
        self.l0, self.u0 = np.asarray(l_new, float), np.asarray(u_new, float)
        self.l, self.u = self.E * self.l0, self.E * self.u0
        ctype, rho_vec = rho_vectors(self.l, self.u, self.s.rho)
        if np.any(ctype != self.ctype):
            self.ctype, self.rho_vec = ctype, rho_vec
            self.kkt.update_rho_vec(self.rho_vec)

    def update_matrices(self, P_new=None, A_new=None):
        if P_new is not None:
            P = sp.csc_matrix(P_new, dtype=float)
            self.P0 = sp.csc_matrix(sp.triu(P) + sp.triu(P, 1).T)
        if A_new is not None:
            self.A0 = sp.csc_matrix(A_new, dtype=float)
        self._scale()
        self.kkt = KKT(self.P, self.A, self.s.sigma, self.rho_vec)
        return 0 if self.kkt.positive == self.n else -1

    def warm_start(self, x=None, y=None):
        """reference src/osqp.c:942-1007"""
        if x is not None:
            self.x = self.Dinv * np.asarray(x, float)
            self.z = self.A @ self.x
        if y is not None:
            self.y = self.c * self.Einv * np.asarray(y, float)


def solve(P, q, A, l, u, **settings):
    """One-shot: setup + solve.  Returns info with x, y, status_val, iter, obj_val, ..."""
    w = Workspace(P, q, A, l, u, **settings)
    info = w.solve()
    info.setup_flag = w.setup_flag
    return info


def solve_kkt(P, A, sigma, rho, rhs):
    """The plugin's `solve` contract on host arrays (qdldl_interface.c:350-376):
    returns [x~ ; z~] for rhs = [rhs_x ; rhs_z]."""
    P = sp.csc_matrix(P, dtype=float)
    P = sp.triu(P) + sp.triu(P, 1).T
    A = sp.csc_matrix(A, dtype=float)
    rho_vec = np.full(A.shape[0], float(rho)) if np.isscalar(rho) else np.asarray(rho, float)
    return KKT(sp.csc_matrix(P), A, sigma, rho_vec).solve(np.asarray(rhs, float))
