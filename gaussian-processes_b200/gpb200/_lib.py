"""ctypes binding of libgpb200.so (include/gpb200.h).

The shared library is the product; this module only marshals numpy buffers across the C ABI.
There is no CPU fallback: if the library is missing, or no B200 is visible, calls raise.
"""
import ctypes as C
import os

import numpy as np

_PKG = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
# GPB200_LIB selects an alternative build of the SAME library (tuning variants); never a fallback
LIB_PATH = os.environ.get("GPB200_LIB") or os.path.join(_PKG, "libgpb200.so")

KIND_SE, KIND_WHITE, KIND_MATERN = 0, 1, 2
FORMULA_AS_PARSED, FORMULA_GPML = 0, 1
GRAD_AS_EXECUTED, GRAD_ANALYTIC = 0, 1

EXPORTS = [
    "gpb_ctx_create", "gpb_ctx_destroy", "gpb_dataset_create", "gpb_dataset_destroy",
    "gpb_loglik", "gpb_loglik_batch", "gpb_ll_value", "gpb_predict", "gpb_distance_matrix",
    "gpb_build_cov", "gpb_potrf", "gpb_trsm_lower", "gpb_trsm_upper", "gpb_mh_run",
    "gpb_philox_streams", "gpb_strerror", "gpb_last_cuda_error", "gpb_launch_count",
    "gpb_last_timing", "gpb_set_sched", "gpb_set_trace", "gpb_get_trace", "gpb_version",
    "gpb_plane_stats", "gpb_sample_plane", "gpb_draw", "gpb_mll_gradient",
    "gpb_schedule_task", "gpb_build_cov_hot", "gpb_mh_run_ex", "gpb_mh_run_host", "gpb_philox_streams_device",
    "gpb_comm_unique_id", "gpb_comm_init_rank", "gpb_comm_init_all", "gpb_comm_destroy", "gpb_comm_info",
    "gpb_comm_shard", "gpb_gather", "gpb_loglik_noalloc",
]
RNG_HOST, RNG_PHILOX = 0, 1
SMALL_AUTO, SMALL_OFF = 0, 1
E_NOTPD, E_NOCOMM = -7, -8
COMM_ID_BYTES = 128


class Term(C.Structure):
    _fields_ = [("kind", C.c_int), ("p0", C.c_double), ("p1", C.c_double), ("p2", C.c_double)]


class LL(C.Structure):
    _fields_ = [("Q", C.c_double), ("U1", C.c_double), ("S", C.c_double), ("n", C.c_int),
                ("info", C.c_int)]


class MhOpts(C.Structure):
    _fields_ = [("formula", C.c_int), ("on_notpd", C.c_int), ("delta", C.c_double),
                ("nugget", C.c_double)]


class MhConfig(C.Structure):
    _fields_ = [("struct_size", C.c_int), ("formula", C.c_int), ("on_notpd", C.c_int), ("rng", C.c_int),
                ("delta", C.c_double), ("nugget", C.c_double), ("seed", C.c_ulonglong), ("chain0", C.c_int),
                ("mean_step", C.c_int), ("mean_prior_mean", C.c_double), ("mean_prior_var", C.c_double),
                ("mean_nugget", C.c_double), ("flush_every", C.c_int), ("gather", C.c_int),
                ("total_chains", C.c_int), ("small_n", C.c_int)]


class MhResult(C.Structure):
    _fields_ = [("info", C.c_int), ("stopped_chains", C.c_int), ("path", C.c_int), ("flushes", C.c_int),
                ("device_ms", C.c_float)]


PROPOSE_FN = C.CFUNCTYPE(None, C.c_void_p, C.c_int, C.c_int, C.POINTER(C.c_double), C.POINTER(C.c_double))
LOG_PRIOR_FN = C.CFUNCTYPE(None, C.c_void_p, C.c_int, C.POINTER(C.c_double), C.POINTER(C.c_double))


class GpbError(RuntimeError):
    def __init__(self, status, msg):
        super().__init__(f"gpb200 status {status}: {msg}")
        self.status = status


_lib = None
_dp = C.POINTER(C.c_double)
_ip = C.POINTER(C.c_int)


def _prefer_bundled_nccl():
    """libgpb200.so binds NCCL at run time (gpb_comm_*).  In a Python process that also imports torch,
    both must end up on the SAME libnccl.so.2: a system NCCL loaded first would shadow the (newer) one
    torch is linked against.  Point GPB_NCCL_LIB at the wheel's copy unless the caller chose one."""
    if os.environ.get("GPB_NCCL_LIB"):
        return
    try:
        import importlib.util
        spec = importlib.util.find_spec("nvidia.nccl")
        for loc in (spec.submodule_search_locations if spec else []):
            cand = os.path.join(loc, "lib", "libnccl.so.2")
            if os.path.exists(cand):
                os.environ["GPB_NCCL_LIB"] = cand
                return
    except Exception:
        pass


def load():
    """Load libgpb200.so; raises (never falls back) when it has not been built."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise ImportError(
            f"{LIB_PATH} not found: build it with `make -C gaussian-processes_b200/csrc` "
            "(or __graft_entry__.build()); gpb200 has no CPU fallback")
    _prefer_bundled_nccl()
    L = C.CDLL(LIB_PATH)
    L.gpb_strerror.restype = C.c_char_p
    L.gpb_last_cuda_error.restype = C.c_char_p
    L.gpb_last_cuda_error.argtypes = [C.c_void_p]
    L.gpb_version.restype = C.c_char_p
    L.gpb_launch_count.restype = C.c_longlong
    L.gpb_launch_count.argtypes = [C.c_void_p]
    L.gpb_ll_value.restype = C.c_double
    L.gpb_ll_value.argtypes = [C.POINTER(LL), C.c_int]
    L.gpb_ctx_create.argtypes = [C.c_int, C.POINTER(C.c_void_p)]
    L.gpb_ctx_destroy.argtypes = [C.c_void_p]
    L.gpb_ctx_destroy.restype = None
    L.gpb_dataset_create.argtypes = [C.c_void_p, C.c_int, C.c_int, _dp, _dp, C.POINTER(C.c_void_p)]
    L.gpb_dataset_destroy.argtypes = [C.c_void_p]
    L.gpb_dataset_destroy.restype = None
    L.gpb_loglik_batch.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.POINTER(Term), C.c_int, _dp,
                                   C.c_int, C.c_double, C.POINTER(LL)]
    L.gpb_loglik.argtypes = [C.c_void_p, C.c_void_p, C.POINTER(Term), C.c_int, _dp, C.c_int,
                             C.c_double, C.POINTER(LL)]
    L.gpb_loglik_noalloc.argtypes = [C.c_int, C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_void_p,
                                     C.POINTER(Term), C.c_int, _dp, C.c_int, C.c_double, C.POINTER(LL)]
    L.gpb_predict.argtypes = [C.c_void_p, C.c_void_p, C.POINTER(Term), C.c_int, C.c_double, C.c_int,
                              _dp, _dp, _dp, _ip]
    L.gpb_distance_matrix.argtypes = [C.c_void_p, C.c_int, C.c_int, _dp, _dp]
    L.gpb_build_cov.argtypes = [C.c_void_p, C.c_int, C.c_int, _dp, C.POINTER(Term), C.c_int,
                                C.c_double, _dp]
    L.gpb_potrf.argtypes = [C.c_void_p, C.c_int, _dp, _ip]
    L.gpb_trsm_lower.argtypes = [C.c_void_p, C.c_int, C.c_int, _dp, _dp]
    L.gpb_trsm_upper.argtypes = [C.c_void_p, C.c_int, C.c_int, _dp, _dp]
    L.gpb_mh_run.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.POINTER(Term), C.c_int,
                             _dp, C.c_int, _dp, C.POINTER(MhOpts), _dp, _dp, _dp, _dp, _ip, _ip]
    L.gpb_philox_streams.argtypes = [C.c_ulonglong, C.c_int, C.c_int, C.c_int, C.c_int, _dp, _dp]
    L.gpb_philox_streams_device.argtypes = [C.c_void_p, C.c_ulonglong, C.c_int, C.c_int, C.c_int, C.c_int, _dp, _dp]
    L.gpb_mh_run_ex.argtypes = [C.c_void_p, C.c_void_p, C.POINTER(MhConfig), C.c_int, C.c_int, C.POINTER(Term),
                                C.c_int, _dp, C.c_int, _dp, _dp, _dp, _dp, _dp, _dp, _ip, _ip, C.POINTER(MhResult)]
    L.gpb_mh_run_host.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.POINTER(Term), C.c_int, _dp, C.c_int,
                                  C.POINTER(MhOpts), PROPOSE_FN, LOG_PRIOR_FN, C.c_void_p, _dp, _dp, _dp, _ip, _ip]
    L.gpb_build_cov_hot.argtypes = [C.c_void_p, C.c_int, C.c_int, _dp, C.POINTER(Term), C.c_int, C.c_double, _dp]
    L.gpb_comm_unique_id.argtypes = [C.c_void_p]
    L.gpb_comm_init_rank.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_void_p]
    L.gpb_comm_init_all.argtypes = [C.POINTER(C.c_void_p), C.c_int]
    L.gpb_comm_destroy.argtypes = [C.c_void_p]
    L.gpb_comm_info.argtypes = [C.c_void_p, _ip, _ip]
    L.gpb_comm_shard.argtypes = [C.c_int, C.c_int, C.c_int, _ip, _ip]
    L.gpb_gather.argtypes = [C.c_void_p, _dp, C.c_int, _dp]
    L.gpb_plane_stats.argtypes = [C.c_void_p, C.c_void_p, C.POINTER(Term), C.c_int, C.c_double, C.c_int,
                                  _dp, _dp, _dp, _ip]
    L.gpb_sample_plane.argtypes = [C.c_void_p, C.c_void_p, C.POINTER(Term), C.c_int, C.c_double, C.c_int,
                                   _dp, C.c_double, C.c_double, _dp, _dp, _ip]
    L.gpb_draw.argtypes = [C.c_void_p, C.c_int, C.c_int, _dp, C.POINTER(Term), C.c_int, C.c_double,
                           _dp, _dp, _dp, _ip]
    L.gpb_mll_gradient.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.POINTER(Term), C.c_int, _dp, C.c_int,
                                   C.c_double, C.c_int, _dp, _ip]
    L.gpb_schedule_task.argtypes = [C.c_int, C.c_int, C.c_int, C.c_longlong, _ip]
    L.gpb_schedule_task.restype = C.c_longlong
    L.gpb_last_timing.argtypes = [C.c_void_p, C.POINTER(C.c_float)]
    L.gpb_set_sched.argtypes = [C.c_void_p, C.c_int]
    L.gpb_set_trace.argtypes = [C.c_void_p, C.c_int]
    L.gpb_get_trace.argtypes = [C.c_void_p, C.c_void_p, C.c_longlong]
    L.gpb_get_trace.restype = C.c_longlong
    _lib = L
    return L


def _ptr(a):
    return a.ctypes.data_as(_dp)


def as_f64(a):
    return np.ascontiguousarray(a, dtype=np.float64)


def x_soa(x):
    """(n,) or (n, dim) array of locations -> contiguous SoA [dim][n], dim"""
    x = np.asarray(x, dtype=np.float64)
    if x.ndim == 1:
        x = x[:, None]
    return np.ascontiguousarray(x.T), x.shape[1]


def terms_array(term_lists):
    """term_lists: list (batch) of lists of (kind, p0, p1) -> (Term * (batch*nterms)), nterms"""
    nterms = len(term_lists[0])
    arr = (Term * (len(term_lists) * nterms))()
    k = 0
    for tl in term_lists:
        if len(tl) != nterms:
            raise ValueError("all proposals of a batch must have the same kernel structure")
        for kind, p0, p1 in tl:
            arr[k] = Term(int(kind), float(p0), float(p1), 0.0)
            k += 1
    return arr, nterms


NOALLOC_MAX_N = 512


def loglik_noalloc(device, stream, n, dim, d_x, d_y, terms, beta=None, nugget=1e-3):
    """gpb_loglik_noalloc: d_x / d_y are DEVICE addresses (ints) the caller owns; `stream` a cudaStream_t
    handle (0 / None = the default stream).  Returns (status, LL); allocates nothing on the device."""
    L = load()
    arr, nterms = terms_array([terms])
    if beta is None or len(beta) == 0:
        bptr, nbeta = None, 0
    else:
        bv = as_f64(beta)
        bptr, nbeta = _ptr(bv), bv.size
    out = LL()
    st = L.gpb_loglik_noalloc(int(device), C.c_void_p(int(stream or 0)), int(n), int(dim), C.c_void_p(int(d_x)),
                              C.c_void_p(int(d_y)), arr, nterms, bptr, nbeta, float(nugget), C.byref(out))
    return st, out


class Context:
    """One device + stream + workspace (gpb_ctx).  Not thread-safe; one per host thread."""

    def __init__(self, device=0):
        self.L = load()
        h = C.c_void_p()
        self._check(self.L.gpb_ctx_create(int(device), C.byref(h)), ctx=None)
        self.h = h
        self.device = device

    def _check(self, status, ctx="self"):
        if status != 0:
            msg = self.L.gpb_strerror(status).decode()
            if ctx is not None and getattr(self, "h", None):
                extra = self.L.gpb_last_cuda_error(self.h).decode()
                if extra:
                    msg += " [" + extra + "]"
            raise GpbError(status, msg)

    def close(self):
        if getattr(self, "h", None):
            self.L.gpb_ctx_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def set_sched(self, mode):
        self._check(self.L.gpb_set_sched(self.h, int(mode)))

    def set_trace(self, on):
        self._check(self.L.gpb_set_trace(self.h, int(on)))

    def get_trace(self):
        """(tasks, 16) uint64 schedule trace of the last factorisation (see gpb200.h)."""
        n = int(self.L.gpb_get_trace(self.h, None, 0))
        out = np.zeros((n, 16), dtype=np.uint64)
        if n:
            self.L.gpb_get_trace(self.h, out.ctypes.data_as(C.c_void_p), n)
        return out

    def launch_count(self):
        return int(self.L.gpb_launch_count(self.h))

    def last_timing(self):
        t = (C.c_float * 3)()
        self._check(self.L.gpb_last_timing(self.h, t))
        return [float(v) for v in t]

    def dataset(self, x, y=None):
        return Dataset(self, x, y)

    # ---- matrix level ----
    def distance_matrix(self, x):
        xs, dim = x_soa(x)
        n = xs.shape[1]
        out = np.empty((n, n), dtype=np.float64)
        self._check(self.L.gpb_distance_matrix(self.h, n, dim, _ptr(xs), _ptr(out)))
        return out.T

    def build_cov(self, x, terms, nugget=0.0):
        xs, dim = x_soa(x)
        n = xs.shape[1]
        arr, nterms = terms_array([terms])
        out = np.empty((n, n), dtype=np.float64)
        self._check(self.L.gpb_build_cov(self.h, n, dim, _ptr(xs), arr, nterms, float(nugget), _ptr(out)))
        return out.T

    def build_cov_hot(self, x, terms, nugget=0.0):
        """The matrix in the form the factorisation consumes (k_build hot-path mode): lower triangle."""
        xs, dim = x_soa(x)
        n = xs.shape[1]
        arr, nterms = terms_array([terms])
        out = np.empty((n, n), dtype=np.float64)
        self._check(self.L.gpb_build_cov_hot(self.h, n, dim, _ptr(xs), arr, nterms, float(nugget), _ptr(out)))
        return out.T

    # ---- communicator (one rank per Context) ----
    def comm_init_rank(self, nranks, rank, uid):
        buf = C.create_string_buffer(bytes(uid), COMM_ID_BYTES)
        self._check(self.L.gpb_comm_init_rank(self.h, int(nranks), int(rank), buf))

    def comm_destroy(self):
        self._check(self.L.gpb_comm_destroy(self.h))

    def comm_info(self):
        """(rank, nranks, has_communicator)"""
        r, n = C.c_int(0), C.c_int(1)
        st = self.L.gpb_comm_info(self.h, C.byref(r), C.byref(n))
        return r.value, n.value, st == 0

    def gather(self, values):
        """All-gather of a vector of doubles through the context's communicator: (nranks, count)."""
        v = as_f64(values).ravel()
        _, nranks, ok = self.comm_info()
        out = np.empty((nranks, v.size))
        self._check(self.L.gpb_gather(self.h, _ptr(v), int(v.size), _ptr(out)))
        return out

    def philox_streams_device(self, seed, chain0, chains, iters, P):
        """The streams exactly as the device generator of mh_run_ex(rng=RNG_PHILOX) consumes them."""
        z = np.empty((iters, chains, P))
        u = np.empty((iters, chains))
        self._check(self.L.gpb_philox_streams_device(self.h, int(seed), int(chain0), int(chains), int(iters), int(P),
                                                     _ptr(z), _ptr(u)))
        return z, u

    def potrf(self, K):
        K = np.asarray(K, dtype=np.float64)
        n = K.shape[0]
        a = np.array(K.T, dtype=np.float64, order="C")  # buffer == column-major K
        info = C.c_int(0)
        self._check(self.L.gpb_potrf(self.h, n, _ptr(a), C.byref(info)))
        return a.T.copy(), info.value

    def _trsm(self, fn, T, B):
        T = np.asarray(T, dtype=np.float64)
        B = np.asarray(B, dtype=np.float64)
        vec = B.ndim == 1
        Bm = B[:, None] if vec else B
        n, nrhs = Bm.shape
        t = np.array(T.T, dtype=np.float64, order="C")
        b = np.array(Bm.T, dtype=np.float64, order="C")
        self._check(fn(self.h, n, nrhs, _ptr(t), _ptr(b)))
        out = b.T.copy()
        return out[:, 0] if vec else out

    def draw(self, x, terms, z, mu=None, nugget=1e-3):
        """gpb_draw: mu + L z with K + nugget I = L L^T.  Returns (sample, info)."""
        xs, dim = x_soa(x)
        n = xs.shape[1]
        arr, nterms = terms_array([terms])
        zv = as_f64(z)
        muv = as_f64(mu) if mu is not None else None
        out = np.empty(n)
        info = C.c_int(0)
        self._check(self.L.gpb_draw(self.h, n, dim, _ptr(xs), arr, nterms, float(nugget),
                                    _ptr(muv) if muv is not None else None, _ptr(zv), _ptr(out),
                                    C.byref(info)))
        return out, info.value

    def trsm_lower(self, L, B):
        return self._trsm(self.L.gpb_trsm_lower, L, B)

    def trsm_upper(self, U, B):
        return self._trsm(self.L.gpb_trsm_upper, U, B)


class Dataset:
    """Observations resident on the device (gpb_dataset)."""

    def __init__(self, ctx, x, y=None):
        self.ctx = ctx
        xs, dim = x_soa(x)
        self.n, self.dim = xs.shape[1], dim
        yv = as_f64(y) if y is not None else None
        if yv is not None and yv.shape != (self.n,):
            raise ValueError("y must have one value per location")
        h = C.c_void_p()
        ctx._check(ctx.L.gpb_dataset_create(ctx.h, self.n, dim, _ptr(xs),
                                            _ptr(yv) if yv is not None else None, C.byref(h)))
        self.h = h

    def close(self):
        if getattr(self, "h", None) and getattr(self.ctx, "h", None):
            self.ctx.L.gpb_dataset_destroy(self.h)
        self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def loglik_batch(self, term_lists, betas=None, nugget=1e-3):
        """term_lists: [batch][nterms] of (kind, p0, p1).  Returns list of dicts Q,U1,S,n,info."""
        ctx = self.ctx
        arr, nterms = terms_array(term_lists)
        batch = len(term_lists)
        if betas is None or len(betas) == 0 or len(betas[0]) == 0:
            bptr, nbeta = None, 0
        else:
            bv = as_f64(betas)
            if bv.shape[0] != batch:
                raise ValueError("one beta vector per proposal")
            bptr, nbeta = _ptr(bv), bv.shape[1]
        out = (LL * batch)()
        ctx._check(ctx.L.gpb_loglik_batch(ctx.h, self.h, batch, arr, nterms, bptr, nbeta,
                                          float(nugget), out))
        return out

    def mll_gradient(self, term_lists, betas=None, nugget=1e-3, mode=GRAD_AS_EXECUTED):
        """term_lists: [batch][nterms] of (kind, p0, p1) holding the UNCONSTRAINED (log) parameters.
        Returns (grad [batch][P], info [batch])."""
        ctx = self.ctx
        arr, nterms = terms_array(term_lists)
        batch = len(term_lists)
        if betas is None or len(betas) == 0 or len(betas[0]) == 0:
            bptr, nbeta = None, 0
        else:
            bv = as_f64(betas)
            if bv.shape[0] != batch:
                raise ValueError("one beta vector per proposal")
            bptr, nbeta = _ptr(bv), bv.shape[1]
        P = sum(2 if t[0] == 0 else 1 for t in term_lists[0])
        grad = np.empty((batch, P))
        info = np.zeros(batch, dtype=np.int32)
        ctx._check(ctx.L.gpb_mll_gradient(ctx.h, self.h, batch, arr, nterms, bptr, nbeta, float(nugget),
                                          int(mode), _ptr(grad), info.ctypes.data_as(_ip)))
        return grad, info

    def predict(self, terms, xstar, nugget=1e-6):
        ctx = self.ctx
        arr, nterms = terms_array([terms])
        xt, dim = x_soa(xstar)
        if dim != self.dim:
            raise ValueError("Dimension mismatch")  # Location.scala:32
        m = xt.shape[1]
        mean = np.empty(m)
        var = np.empty(m)
        info = C.c_int(0)
        ctx._check(ctx.L.gpb_predict(ctx.h, self.h, arr, nterms, float(nugget), m, _ptr(xt),
                                     _ptr(mean), _ptr(var), C.byref(info)))
        return mean, var, info.value


    def plane_stats(self, terms, X, nugget=1e-6):
        """gpb_plane_stats: X (n, d) design matrix -> (G = X^T K^-1 X, w = X^T L^-1 y, info)."""
        ctx = self.ctx
        arr, nterms = terms_array([terms])
        Xc = np.ascontiguousarray(np.asarray(X, dtype=np.float64).T)   # [d][n]
        d = Xc.shape[0]
        G = np.empty((d, d)); w = np.empty(d)
        info = C.c_int(0)
        ctx._check(ctx.L.gpb_plane_stats(ctx.h, self.h, arr, nterms, float(nugget), d, _ptr(Xc),
                                         _ptr(G), _ptr(w), C.byref(info)))
        return G, w, info.value

    def sample_plane(self, terms, X, prior_mean, prior_var, z, nugget=1e-6):
        """gpb_sample_plane: one Gibbs draw of the plane coefficients.  Returns (beta, info)."""
        ctx = self.ctx
        arr, nterms = terms_array([terms])
        Xc = np.ascontiguousarray(np.asarray(X, dtype=np.float64).T)
        d = Xc.shape[0]
        zv = as_f64(z)
        beta = np.empty(d)
        info = C.c_int(0)
        ctx._check(ctx.L.gpb_sample_plane(ctx.h, self.h, arr, nterms, float(nugget), d, _ptr(Xc),
                                          float(prior_mean), float(prior_var), _ptr(zv), _ptr(beta),
                                          C.byref(info)))
        return beta, info.value

    def mh_run(self, init_term_lists, ig, delta, z, u, beta=(), nugget=1e-3,
               formula=FORMULA_AS_PARSED, on_notpd=0):
        """Device-resident MH for len(init_term_lists) chains (gpb_mh_run).
        z: (iters, chains, P) standard normals, u: (iters, chains) uniforms, ig: (P, 2).
        Returns states (iters, chains, P), ll (iters, chains), accept (iters, chains), info."""
        ctx = self.ctx
        arr, nterms = terms_array(init_term_lists)
        chains = len(init_term_lists)
        zv, uv, igv = as_f64(z), as_f64(u), as_f64(ig)
        iters = uv.shape[0]
        P = igv.shape[0]
        if uv.shape != (iters, chains) or zv.shape != (iters, chains, P) or igv.shape != (P, 2):
            raise ValueError("stream / prior shapes do not match (iters, chains, P)")
        bv = as_f64(beta) if len(beta) else None
        opts = MhOpts(int(formula), int(on_notpd), float(delta), float(nugget))
        states = np.empty((iters, chains, P))
        ll = np.empty((iters, chains))
        acc = np.empty((iters, chains), dtype=np.int32)
        info = C.c_int(0)
        st = ctx.L.gpb_mh_run(ctx.h, self.h, chains, iters, arr, nterms,
                              _ptr(bv) if bv is not None else None, 0 if bv is None else bv.shape[0],
                              _ptr(igv), C.byref(opts), _ptr(zv), _ptr(uv), _ptr(states), _ptr(ll),
                              acc.ctypes.data_as(_ip), C.byref(info))
        if st == -7:  # GPB_E_NOTPD: the reference throws from cholesky here
            raise GpbError(st, f"proposal covariance not positive definite (info={info.value})")
        ctx._check(st)
        return states, ll, acc, info.value


    def mh_run_ex(self, init_term_lists, ig, delta, iters, z=None, u=None, beta0=None, design=None,
                  nugget=1e-3, formula=FORMULA_AS_PARSED, on_notpd=0, rng=RNG_HOST, seed=0, chain0=0,
                  mean_step=False, mean_prior=(0.0, 1.0), mean_nugget=1e-6, flush_every=0, gather=False,
                  total_chains=0, small_n=SMALL_AUTO, raise_on_notpd=True):
        """gpb_mh_run_ex: the Mcmc.sample iteration for len(init_term_lists) chains on this device.
        z: (iters, chains, P [+ d]) / u: (iters, chains) for rng=RNG_HOST; beta0: (chains, nbeta);
        design: (n, d) design matrix (mean_step).  Returns a dict: states (iters, C, W), ll, accept
        (iters, C), stop (C,), info, stopped_chains, path, flushes, device_ms, status."""
        ctx = self.ctx
        arr, nterms = terms_array(init_term_lists)
        chains = len(init_term_lists)
        igv = as_f64(ig)
        P = igv.shape[0]
        if igv.shape != (P, 2) or P != sum(2 if t[0] == KIND_SE else 1 for t in init_term_lists[0]):
            raise ValueError("ig must hold (shape, scale) for each of the P kernel parameters")
        bv = as_f64(beta0) if beta0 is not None and len(beta0) else None
        nbeta = 0 if bv is None else bv.shape[1]
        if bv is not None and bv.shape != (chains, nbeta):
            raise ValueError("beta0 must be (chains, nbeta)")
        d = nbeta if mean_step else 0
        W = P + d
        zv = uv = None
        if rng == RNG_HOST:
            zv, uv = as_f64(z), as_f64(u)
            if uv.shape != (iters, chains) or zv.shape != (iters, chains, P + d):
                raise ValueError("stream shapes do not match (iters, chains, P + d)")
        dv = None
        if mean_step:
            dv = np.ascontiguousarray(np.asarray(design, dtype=np.float64).T)   # [d][n]
            if dv.shape != (d, self.n):
                raise ValueError("design must be (n, nbeta)")
        cfg = MhConfig(C.sizeof(MhConfig), int(formula), int(on_notpd), int(rng), float(delta), float(nugget),
                       int(seed), int(chain0), int(bool(mean_step)), float(mean_prior[0]), float(mean_prior[1]),
                       float(mean_nugget), int(flush_every), int(bool(gather)), int(total_chains), int(small_n))
        Cn = total_chains if gather else chains
        states = np.empty((iters, Cn, W)); ll = np.empty((iters, Cn))
        acc = np.empty((iters, Cn), dtype=np.int32); stop = np.empty(Cn, dtype=np.int32)
        res = MhResult()
        st = ctx.L.gpb_mh_run_ex(ctx.h, self.h, C.byref(cfg), chains, int(iters), arr, nterms,
                                 _ptr(bv) if bv is not None else None, nbeta, _ptr(igv),
                                 _ptr(zv) if zv is not None else None, _ptr(uv) if uv is not None else None,
                                 _ptr(dv) if dv is not None else None, _ptr(states), _ptr(ll),
                                 acc.ctypes.data_as(_ip), stop.ctypes.data_as(_ip), C.byref(res))
        if st == E_NOTPD and raise_on_notpd:  # the reference throws from cholesky on that chain
            raise GpbError(st, f"proposal covariance not positive definite (info={res.info}, "
                               f"{res.stopped_chains} chain(s) stopped)")
        if st != E_NOTPD:
            ctx._check(st)
        return {"states": states, "ll": ll, "accept": acc, "stop": stop, "info": res.info,
                "stopped_chains": res.stopped_chains, "path": res.path, "flushes": res.flushes,
                "device_ms": res.device_ms, "status": st}

    def mh_run_host(self, init_term_lists, propose, log_prior, u, beta=(), nugget=1e-3,
                    formula=FORMULA_AS_PARSED, on_notpd=0):
        """gpb_mh_run_host: arbitrary host closures.  propose(iter, cur (chains, P)) -> (chains, P);
        log_prior(params (chains, P)) -> (chains,).  u: (iters, chains)."""
        ctx = self.ctx
        arr, nterms = terms_array(init_term_lists)
        chains = len(init_term_lists)
        P = sum(2 if t[0] == KIND_SE else 1 for t in init_term_lists[0])
        uv = as_f64(u)
        iters = uv.shape[0]
        bv = as_f64(beta) if len(beta) else None

        def _prop(_user, it, nch, cur, nxt):
            a = np.ctypeslib.as_array(cur, shape=(nch, P))
            np.ctypeslib.as_array(nxt, shape=(nch, P))[:] = np.asarray(propose(it, a.copy()), dtype=np.float64)

        def _prior(_user, nch, par, out):
            a = np.ctypeslib.as_array(par, shape=(nch, P))
            np.ctypeslib.as_array(out, shape=(nch,))[:] = np.asarray(log_prior(a.copy()), dtype=np.float64)

        opts = MhOpts(int(formula), int(on_notpd), 0.0, float(nugget))
        states = np.empty((iters, chains, P)); ll = np.empty((iters, chains)); acc = np.empty((iters, chains), dtype=np.int32)
        info = C.c_int(0)
        st = ctx.L.gpb_mh_run_host(ctx.h, self.h, chains, iters, arr, nterms, _ptr(bv) if bv is not None else None,
                                   0 if bv is None else bv.shape[0], C.byref(opts), PROPOSE_FN(_prop), LOG_PRIOR_FN(_prior),
                                   None, _ptr(uv), _ptr(states), _ptr(ll), acc.ctypes.data_as(_ip), C.byref(info))
        if st == E_NOTPD:
            raise GpbError(st, f"proposal covariance not positive definite (info={info.value})")
        ctx._check(st)
        return states, ll, acc, info.value


def comm_unique_id():
    """gpb_comm_unique_id: the 128-byte id rank 0 hands to the other ranks."""
    buf = C.create_string_buffer(COMM_ID_BYTES)
    st = load().gpb_comm_unique_id(buf)
    if st != 0:
        raise GpbError(st, load().gpb_strerror(st).decode())
    return buf.raw


def comm_init_all(contexts):
    """gpb_comm_init_all: communicators for a list of Contexts (one per device) from one thread."""
    arr = (C.c_void_p * len(contexts))(*[c.h for c in contexts])
    st = load().gpb_comm_init_all(arr, len(contexts))
    if st != 0:
        raise GpbError(st, load().gpb_strerror(st).decode())


def comm_shard(total, nranks, rank):
    f, c = C.c_int(0), C.c_int(0)
    st = load().gpb_comm_shard(int(total), int(nranks), int(rank), C.byref(f), C.byref(c))
    if st != 0:
        raise GpbError(st, "bad shard arguments")
    return f.value, c.value


def schedule(n, appended=1, batch=1):
    """Ticket order of the factorisation kernel: (tickets, 4) int array of {proposal, tile row, tile
    column, kind} (gpb_schedule_task; host only)."""
    L = load()
    probe = (C.c_int * 4)()
    total = int(L.gpb_schedule_task(n, appended, batch, -1, probe))
    if total < 0:
        raise GpbError(total, "gpb_schedule_task")
    out = np.zeros((total, 4), dtype=np.int32)
    for t in range(total):
        L.gpb_schedule_task(n, appended, batch, t, out[t].ctypes.data_as(_ip))
    return out


def philox_streams(seed, chain0, chains, iters, P):
    """gpb_philox_streams: z (iters, chains, P) normals and u (iters, chains) uniforms (host)."""
    z = np.empty((iters, chains, P))
    u = np.empty((iters, chains))
    st = load().gpb_philox_streams(int(seed), int(chain0), int(chains), int(iters), int(P), _ptr(z), _ptr(u))
    if st != 0:
        raise GpbError(st, load().gpb_strerror(st).decode())
    return z, u


def ll_value(r, formula=FORMULA_AS_PARSED):
    return load().gpb_ll_value(C.byref(r), int(formula))
