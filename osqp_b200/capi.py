"""ctypes view of the OSQP C ABI (include/osqp.h).

The struct layouts below are those of reference include/types.h:21-288 as its CMake
defaults configure them (DLONG, double, PRINTING, PROFILING), which are also the
layouts of this repo's include/osqp.h.  The same Python classes therefore drive

  * osqp_b200/lib/libosqp_b200.so  -- the product (C host + CUDA plugin), and
  * any other library exporting the same ABI (the tests and bench.py pass the path
    of the reference compiled by oracle/Makefile as the checker / CPU baseline).

Nothing here computes: it marshals numpy arrays into `csc` / `OSQPData`, calls
osqp_setup / osqp_solve / osqp_update_* / osqp_cleanup and reads back
`work->info` and `work->solution`.  There is no CPU fallback: if the CUDA plugin
or a B200 is missing, osqp_setup fails with OSQP_LINSYS_SOLVER_LOAD_ERROR.
"""
import ctypes as C
import os

import numpy as np

c_int = C.c_longlong
c_float = C.c_double
c_int_p = C.POINTER(c_int)
c_float_p = C.POINTER(c_float)

OSQP_INFTY = 1e30
OSQP_NAN = float(0x7FC00000)

# status values, reference include/constants.h:17-29
OSQP_DUAL_INFEASIBLE_INACCURATE = 4
OSQP_PRIMAL_INFEASIBLE_INACCURATE = 3
OSQP_SOLVED_INACCURATE = 2
OSQP_SOLVED = 1
OSQP_MAX_ITER_REACHED = -2
OSQP_PRIMAL_INFEASIBLE = -3
OSQP_DUAL_INFEASIBLE = -4
OSQP_SIGINT = -5
OSQP_TIME_LIMIT_REACHED = -6
OSQP_NON_CVX = -7
OSQP_UNSOLVED = -10

QDLDL_SOLVER = 0
MKL_PARDISO_SOLVER = 1
CUDA_PCG_SOLVER = 2

# error codes, reference include/constants.h:42-50
OSQP_DATA_VALIDATION_ERROR = 1
OSQP_SETTINGS_VALIDATION_ERROR = 2
OSQP_LINSYS_SOLVER_LOAD_ERROR = 3
OSQP_LINSYS_SOLVER_INIT_ERROR = 4
OSQP_NONCVX_ERROR = 5
OSQP_MEM_ALLOC_ERROR = 6
OSQP_WORKSPACE_NOT_INIT_ERROR = 7


class csc(C.Structure):
    _fields_ = [("nzmax", c_int), ("m", c_int), ("n", c_int), ("p", c_int_p),
                ("i", c_int_p), ("x", c_float_p), ("nz", c_int)]


class OSQPScaling(C.Structure):
    _fields_ = [("c", c_float), ("D", c_float_p), ("E", c_float_p), ("cinv", c_float),
                ("Dinv", c_float_p), ("Einv", c_float_p)]


class OSQPSolution(C.Structure):
    _fields_ = [("x", c_float_p), ("y", c_float_p)]


class OSQPInfo(C.Structure):
    _fields_ = [("iter", c_int), ("status", C.c_char * 32), ("status_val", c_int),
                ("status_polish", c_int), ("obj_val", c_float), ("pri_res", c_float),
                ("dua_res", c_float), ("setup_time", c_float), ("solve_time", c_float),
                ("update_time", c_float), ("polish_time", c_float), ("run_time", c_float),
                ("rho_updates", c_int), ("rho_estimate", c_float)]


class OSQPData(C.Structure):
    _fields_ = [("n", c_int), ("m", c_int), ("P", C.POINTER(csc)), ("A", C.POINTER(csc)),
                ("q", c_float_p), ("l", c_float_p), ("u", c_float_p)]


class OSQPSettings(C.Structure):
    _fields_ = [("rho", c_float), ("sigma", c_float), ("scaling", c_int),
                ("adaptive_rho", c_int), ("adaptive_rho_interval", c_int),
                ("adaptive_rho_tolerance", c_float), ("adaptive_rho_fraction", c_float),
                ("max_iter", c_int), ("eps_abs", c_float), ("eps_rel", c_float),
                ("eps_prim_inf", c_float), ("eps_dual_inf", c_float), ("alpha", c_float),
                ("linsys_solver", C.c_int), ("delta", c_float), ("polish", c_int),
                ("polish_refine_iter", c_int), ("verbose", c_int),
                ("scaled_termination", c_int), ("check_termination", c_int),
                ("warm_start", c_int), ("time_limit", c_float)]


class OSQPPolish(C.Structure):
    _fields_ = [("Ared", C.POINTER(csc)), ("n_low", c_int), ("n_upp", c_int),
                ("A_to_Alow", c_int_p), ("A_to_Aupp", c_int_p), ("Alow_to_A", c_int_p),
                ("Aupp_to_A", c_int_p), ("x", c_float_p), ("z", c_float_p), ("y", c_float_p),
                ("obj_val", c_float), ("pri_res", c_float), ("dua_res", c_float)]


class LinSysSolver(C.Structure):
    pass


LinSysSolver._fields_ = [
    ("type", C.c_int),
    ("solve", C.CFUNCTYPE(c_int, C.POINTER(LinSysSolver), c_float_p)),
    ("free", C.CFUNCTYPE(None, C.POINTER(LinSysSolver))),
    ("update_matrices", C.CFUNCTYPE(c_int, C.POINTER(LinSysSolver), C.POINTER(csc), C.POINTER(csc))),
    ("update_rho_vec", C.CFUNCTYPE(c_int, C.POINTER(LinSysSolver), c_float_p)),
    ("nthreads", c_int),
]


class OSQPWorkspace(C.Structure):
    _fields_ = [("data", C.POINTER(OSQPData)), ("linsys_solver", C.POINTER(LinSysSolver)),
                ("pol", C.POINTER(OSQPPolish)),
                ("rho_vec", c_float_p), ("rho_inv_vec", c_float_p), ("constr_type", c_int_p),
                ("x", c_float_p), ("y", c_float_p), ("z", c_float_p), ("xz_tilde", c_float_p),
                ("x_prev", c_float_p), ("z_prev", c_float_p), ("Ax", c_float_p),
                ("Px", c_float_p), ("Aty", c_float_p), ("delta_y", c_float_p),
                ("Atdelta_y", c_float_p), ("delta_x", c_float_p), ("Pdelta_x", c_float_p),
                ("Adelta_x", c_float_p), ("D_temp", c_float_p), ("D_temp_A", c_float_p),
                ("E_temp", c_float_p),
                ("settings", C.POINTER(OSQPSettings)), ("scaling", C.POINTER(OSQPScaling)),
                ("solution", C.POINTER(OSQPSolution)), ("info", C.POINTER(OSQPInfo)),
                ("timer", C.c_void_p), ("first_run", c_int), ("clear_update_time", c_int),
                ("rho_update_from_solve", c_int), ("summary_printed", c_int)]


class OSQPB200Stats(C.Structure):
    _fields_ = [("admm_iters", c_int), ("pcg_iters", c_int), ("kernel_launches", c_int),
                ("rho_updates", c_int), ("pcg_maxed", c_int), ("pcg_breakdowns", c_int),
                ("sell_launches", c_int), ("tile_launches", c_int), ("pcg_noncvx", c_int),
                ("slack_vars", c_int)]


_HERE = os.path.dirname(os.path.abspath(__file__))
PRODUCT_LIB = os.path.join(_HERE, "lib", "libosqp_b200.so")
PLUGIN_LIB = os.environ.get("OSQP_CUDA_PCG_LIB") or os.path.join(_HERE, "lib", "libosqp_cuda_pcg.so")


def _declare(lib):
    W = C.POINTER(OSQPWorkspace)
    lib.osqp_set_default_settings.argtypes = [C.POINTER(OSQPSettings)]
    lib.osqp_set_default_settings.restype = None
    lib.osqp_setup.argtypes = [C.POINTER(W), C.POINTER(OSQPData), C.POINTER(OSQPSettings)]
    lib.osqp_setup.restype = c_int
    for name in ("osqp_solve", "osqp_cleanup"):
        f = getattr(lib, name)
        f.argtypes = [W]
        f.restype = c_int
    for name in ("osqp_update_lin_cost", "osqp_update_lower_bound", "osqp_update_upper_bound",
                 "osqp_warm_start_x", "osqp_warm_start_y"):
        f = getattr(lib, name)
        f.argtypes = [W, c_float_p]
        f.restype = c_int
    for name in ("osqp_update_bounds", "osqp_warm_start"):
        f = getattr(lib, name)
        f.argtypes = [W, c_float_p, c_float_p]
        f.restype = c_int
    for name in ("osqp_update_P", "osqp_update_A"):
        f = getattr(lib, name)
        f.argtypes = [W, c_float_p, c_int_p, c_int]
        f.restype = c_int
    lib.osqp_update_P_A.argtypes = [W, c_float_p, c_int_p, c_int, c_float_p, c_int_p, c_int]
    lib.osqp_update_P_A.restype = c_int
    lib.osqp_update_rho.argtypes = [W, c_float]
    lib.osqp_update_rho.restype = c_int
    if hasattr(lib, "osqp_solve_batch"):          # additive entry point of the product library
        lib.osqp_solve_batch.argtypes = [C.POINTER(W), c_int, c_int]
        lib.osqp_solve_batch.restype = c_int
    for name, t in (("max_iter", c_int), ("eps_abs", c_float), ("eps_rel", c_float),
                    ("eps_prim_inf", c_float), ("eps_dual_inf", c_float), ("alpha", c_float),
                    ("warm_start", c_int), ("scaled_termination", c_int),
                    ("check_termination", c_int), ("delta", c_float), ("polish", c_int),
                    ("polish_refine_iter", c_int), ("verbose", c_int), ("time_limit", c_float)):
        f = getattr(lib, "osqp_update_" + name)
        f.argtypes = [W, t]
        f.restype = c_int
    return lib


_LIBS = {}


def load_library(path=None):
    """dlopen an OSQP-ABI library (default: the product) and declare its prototypes."""
    path = os.path.abspath(path or PRODUCT_LIB)
    if path not in _LIBS:
        if not os.path.exists(path):
            raise OSError(
                f"{path} is missing: build it first (python -c 'import __graft_entry__ as g; "
                "g.build()' or make -C osqp_b200/csrc). There is no CPU fallback.")
        _LIBS[path] = _declare(C.CDLL(path, mode=C.RTLD_LOCAL))
    return _LIBS[path]


def _f64(a):
    return np.ascontiguousarray(a, dtype=np.float64)


def _i64(a):
    return np.ascontiguousarray(a, dtype=np.int64)


def _fp(a):
    return a.ctypes.data_as(c_float_p) if a is not None else None


def _ip(a):
    return a.ctypes.data_as(c_int_p) if a is not None else None


class CscHolder:
    """A `csc` struct plus the numpy arrays it borrows (kept alive here)."""

    def __init__(self, M, upper=False):
        import scipy.sparse as sp
        M = sp.csc_matrix(M)
        if upper:
            M = sp.triu(M, format="csc")
        M.sort_indices()
        self._wrap(M.shape, _i64(M.indptr), _i64(M.indices), _f64(M.data))

    def _wrap(self, shape, p, i, x):
        self.shape = tuple(shape)
        self.p, self.i, self.x = p, i, x
        self.nnz = int(self.p[-1])
        self.struct = csc(self.nnz, shape[0], shape[1], _ip(self.p), _ip(self.i),
                          _fp(self.x), -1)

    @classmethod
    def from_arrays(cls, shape, p, i, x):
        """Borrow caller-owned int64 / float64 arrays as they are (e.g. pinned host buffers)."""
        assert p.dtype == np.int64 and i.dtype == np.int64 and x.dtype == np.float64
        self = cls.__new__(cls)
        self._wrap(shape, p, i, x)
        return self


DEFAULT_SETTING_NAMES = [f[0] for f in OSQPSettings._fields_]


class Solver:
    """One OSQP workspace behind the C API.  Method names follow the C entry points.

    lib_path=None -> the product library (CUDA).  Settings are keyword arguments with
    the names of OSQPSettings (reference include/types.h:139-176).
    """

    def __init__(self, P, q, A, l, u, lib_path=None, **settings):
        self.lib = load_library(lib_path)
        self.is_product = lib_path is None or os.path.abspath(lib_path) == os.path.abspath(PRODUCT_LIB)
        self.n = int(len(q))
        self.m = int(len(l))
        self._P = P if isinstance(P, CscHolder) else CscHolder(P, upper=True)
        self._A = A if isinstance(A, CscHolder) else CscHolder(A)
        self._q, self._l, self._u = _f64(q), _f64(l), _f64(u)
        self._data = OSQPData(self.n, self.m, C.pointer(self._P.struct), C.pointer(self._A.struct),
                              _fp(self._q), _fp(self._l), _fp(self._u))
        self.settings = OSQPSettings()
        self.lib.osqp_set_default_settings(C.byref(self.settings))
        self.settings.verbose = 0
        for k, v in settings.items():
            if k not in DEFAULT_SETTING_NAMES:
                raise KeyError(k)
            setattr(self.settings, k, v)
        self.work = C.POINTER(OSQPWorkspace)()
        self.setup_flag = int(self.lib.osqp_setup(C.byref(self.work), C.byref(self._data),
                                                  C.byref(self.settings)))
        if self.setup_flag != 0:
            # reference tests call osqp_cleanup on a failed setup (tests/non_cvx/test_non_cvx.h:32-38)
            if self.work:
                self.lib.osqp_cleanup(self.work)
            self.work = C.POINTER(OSQPWorkspace)()

    # -- lifetime
    def cleanup(self):
        if self.work:
            self.lib.osqp_cleanup(self.work)
            self.work = C.POINTER(OSQPWorkspace)()

    def __del__(self):
        try:
            self.cleanup()
        except Exception:
            pass

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.cleanup()

    # -- solve + results
    def solve(self):
        flag = int(self.lib.osqp_solve(self.work))
        return flag

    @property
    def info(self):
        return self.work.contents.info.contents

    @property
    def status(self):
        return self.info.status.decode()

    @property
    def x(self):
        return np.ctypeslib.as_array(self.work.contents.solution.contents.x, (self.n,)).copy()

    @property
    def y(self):
        if self.m == 0:
            return np.zeros(0)
        return np.ctypeslib.as_array(self.work.contents.solution.contents.y, (self.m,)).copy()

    def data_vec(self, name):
        """work->data->{q,l,u}: the (scaled) host mirrors the reference tests read."""
        d = self.work.contents.data.contents
        ln = self.n if name == "q" else self.m
        if ln == 0:
            return np.zeros(0)
        return np.ctypeslib.as_array(getattr(d, name), (ln,)).copy()

    def certificate(self, which):
        w = self.work.contents
        if which == "delta_y":
            return np.ctypeslib.as_array(w.delta_y, (self.m,)).copy()
        return np.ctypeslib.as_array(w.delta_x, (self.n,)).copy()

    def current_setting(self, name):
        return getattr(self.work.contents.settings.contents, name)

    # -- updates (reference src/osqp.c:765-1332)
    def update_lin_cost(self, q):
        q = _f64(q)
        return int(self.lib.osqp_update_lin_cost(self.work, _fp(q)))

    def update_bounds(self, l, u):
        l, u = _f64(l), _f64(u)
        return int(self.lib.osqp_update_bounds(self.work, _fp(l), _fp(u)))

    def update_lower_bound(self, l):
        l = _f64(l)
        return int(self.lib.osqp_update_lower_bound(self.work, _fp(l)))

    def update_upper_bound(self, u):
        u = _f64(u)
        return int(self.lib.osqp_update_upper_bound(self.work, _fp(u)))

    def warm_start(self, x=None, y=None):
        if x is not None and y is not None:
            x, y = _f64(x), _f64(y)
            return int(self.lib.osqp_warm_start(self.work, _fp(x), _fp(y)))
        if x is not None:
            x = _f64(x)
            return int(self.lib.osqp_warm_start_x(self.work, _fp(x)))
        y = _f64(y)
        return int(self.lib.osqp_warm_start_y(self.work, _fp(y)))

    def update_P(self, Px, idx=None):
        Px = _f64(Px)
        idx = _i64(idx) if idx is not None else None
        return int(self.lib.osqp_update_P(self.work, _fp(Px), _ip(idx), len(Px)))

    def update_A(self, Ax, idx=None):
        Ax = _f64(Ax)
        idx = _i64(idx) if idx is not None else None
        return int(self.lib.osqp_update_A(self.work, _fp(Ax), _ip(idx), len(Ax)))

    def update_P_A(self, Px, Pidx, Ax, Aidx):
        Px, Ax = _f64(Px), _f64(Ax)
        Pidx = _i64(Pidx) if Pidx is not None else None
        Aidx = _i64(Aidx) if Aidx is not None else None
        return int(self.lib.osqp_update_P_A(self.work, _fp(Px), _ip(Pidx), len(Px),
                                            _fp(Ax), _ip(Aidx), len(Ax)))

    def update_rho(self, rho):
        return int(self.lib.osqp_update_rho(self.work, rho))

    def update_setting(self, name, value):
        return int(getattr(self.lib, "osqp_update_" + name)(self.work, value))

    # -- additive, product only
    def stats(self):
        if not self.is_product:
            raise RuntimeError("stats() exists in the product library only")
        f = self.lib.osqp_b200_get_stats
        f.argtypes = [C.POINTER(OSQPWorkspace), C.POINTER(OSQPB200Stats)]
        f.restype = c_int
        s = OSQPB200Stats()
        f(self.work, C.byref(s))
        return {k: int(getattr(s, k)) for k, _ in OSQPB200Stats._fields_}

    def iterates(self):
        """Scaled device iterates (x, z, y) copied to the host -- product only."""
        f = self.lib.osqp_b200_get_iterates
        f.argtypes = [C.POINTER(OSQPWorkspace), c_float_p, c_float_p, c_float_p]
        f.restype = c_int
        x, z, y = np.zeros(self.n), np.zeros(max(self.m, 1)), np.zeros(max(self.m, 1))
        f(self.work, _fp(x), _fp(z), _fp(y))
        return x, z[:self.m], y[:self.m]

    def engine_handle(self):
        f = self.lib.osqp_b200_engine
        f.argtypes = [C.POINTER(OSQPWorkspace)]
        f.restype = C.c_void_p
        return f(self.work)


def solve_batch(solvers, nthreads=16):
    """osqp_solve_batch over a list of Solver objects of the product library (BASELINE config 4):
    returns the number of failed solves; results are read from each Solver as after solve()."""
    if not solvers:
        return 0
    lib = solvers[0].lib
    W = C.POINTER(OSQPWorkspace)
    arr = (W * len(solvers))(*[s.work for s in solvers])
    return int(lib.osqp_solve_batch(arr, len(solvers), int(nthreads)))


class Plugin:
    """ctypes view of the engine-level entry points of libosqp_cuda_pcg.so that bench.py and
    the tests use for device timing and per-kernel profiling (include/osqp_cuda_pcg.h)."""

    KERNELS = ["spmv_A(rho.*A p)", "spmv_M([P|A'][p;w])", "rhs([P|A'][x;v])", "pcg_update",
               "pcg_dir", "admm_update(A x~ + x,z,y)", "resid_m", "resid_n",
               "tile_M([P|A'] product only)", "tile_A(A product only)",
               "allreduce([P|A'] partial sums, NCCL)", "apply_epilogue(reduced sums)"]

    def __init__(self, path=None):
        path = os.path.abspath(path or PLUGIN_LIB)
        if not os.path.exists(path):
            raise OSError(f"{path} is missing: build it first; there is no CPU fallback")
        self.lib = C.CDLL(path)
        L = self.lib
        L.osqp_cuda_engine_profile.argtypes = [C.c_void_p, c_int]
        L.osqp_cuda_engine_profile.restype = c_int
        L.osqp_cuda_engine_profile_get.argtypes = [C.c_void_p, c_int, c_int_p, c_float_p, c_float_p]
        L.osqp_cuda_engine_profile_get.restype = c_int
        L.osqp_cuda_engine_timer_start.argtypes = [C.c_void_p]
        L.osqp_cuda_engine_timer_start.restype = c_int
        L.osqp_cuda_engine_timer_stop.argtypes = [C.c_void_p, c_float_p]
        L.osqp_cuda_engine_timer_stop.restype = c_int
        L.osqp_cuda_engine_set_pcg.argtypes = [C.c_void_p, c_float, c_float, c_int]
        L.osqp_cuda_engine_set_pcg.restype = c_int
        L.osqp_cuda_engine_set_pcg_lambda.argtypes = [C.c_void_p, c_float]
        L.osqp_cuda_engine_set_pcg_lambda.restype = c_int
        L.osqp_cuda_engine_sync.argtypes = [C.c_void_p]
        L.osqp_cuda_engine_sync.restype = c_int
        L.osqp_cuda_nccl_unique_id.argtypes = [C.c_char_p]
        L.osqp_cuda_nccl_unique_id.restype = c_int
        L.osqp_cuda_set_partition.argtypes = [c_int, c_int, C.c_char_p]
        L.osqp_cuda_set_partition.restype = c_int
        L.osqp_cuda_get_partition.argtypes = [c_int_p, c_int_p, c_int_p]
        L.osqp_cuda_get_partition.restype = c_int

    # ---- row-partitioned mode bootstrap (include/osqp_cuda_pcg.h)
    def nccl_unique_id(self):
        buf = C.create_string_buffer(128)
        if self.lib.osqp_cuda_nccl_unique_id(buf):
            raise RuntimeError("cannot create an NCCL unique id")
        return buf.raw

    def set_partition(self, rank, world, uid=None):
        if self.lib.osqp_cuda_set_partition(rank, world, uid):
            raise RuntimeError("osqp_cuda_set_partition failed")

    def get_partition(self):
        r, w, a = c_int(0), c_int(0), c_int(0)
        self.lib.osqp_cuda_get_partition(C.byref(r), C.byref(w), C.byref(a))
        return dict(rank=int(r.value), world=int(w.value), allreduces=int(a.value))

    def profile(self, engine, enable):
        return int(self.lib.osqp_cuda_engine_profile(engine, 1 if enable else 0))

    def profile_get(self, engine):
        out = []
        for k, name in enumerate(self.KERNELS):
            n, ms, b = c_int(0), c_float(0), c_float(0)
            self.lib.osqp_cuda_engine_profile_get(engine, k, C.byref(n), C.byref(ms), C.byref(b))
            out.append(dict(kernel=name, launches=int(n.value), total_ms=float(ms.value),
                            bytes_per_launch=float(b.value)))
        return out

    def timer_start(self, engine):
        return int(self.lib.osqp_cuda_engine_timer_start(engine))

    def timer_stop(self, engine):
        ms = c_float(0)
        if self.lib.osqp_cuda_engine_timer_stop(engine, C.byref(ms)):
            raise RuntimeError("device timer failed")
        return float(ms.value)

    def set_pcg(self, engine, rel_tol=0.0, abs_floor=0.0, max_iter=0):
        return int(self.lib.osqp_cuda_engine_set_pcg(engine, rel_tol, abs_floor, max_iter))

    def set_pcg_lambda(self, engine, lam):
        return int(self.lib.osqp_cuda_engine_set_pcg_lambda(engine, lam))

    def sync(self, engine):
        return int(self.lib.osqp_cuda_engine_sync(engine))
