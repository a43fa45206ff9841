"""ctypes binding of the CPU oracle (oracle/gp_oracle.c).

TEST INFRASTRUCTURE ONLY -- see the header of gp_oracle.c.  Imported by tests/, by
__graft_entry__.smoke() and by bench.py's cpu_baseline / --impl reference legs, never by the
product package.  Parity unpinned (the reference ships no golden vectors for this path).
"""
import ctypes as C
import math
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "libgp_oracle.so")

SE, WHITE = 0, 1


class Term(C.Structure):
    _fields_ = [("kind", C.c_int), ("p0", C.c_double), ("p1", C.c_double), ("p2", C.c_double)]


class LL(C.Structure):
    _fields_ = [("Q", C.c_double), ("U1", C.c_double), ("S", C.c_double), ("ll_ref", C.c_double),
                ("ll_gpml", C.c_double), ("n", C.c_int), ("info", C.c_int)]


def build():
    src = os.path.join(_HERE, "gp_oracle.c")
    if not os.path.exists(_SO) or os.path.getmtime(_SO) < os.path.getmtime(src):
        subprocess.check_call(["make", "-C", _HERE, "-s"])


_lib = None
_dp = C.POINTER(C.c_double)
_ip = C.POINTER(C.c_int)


def lib():
    global _lib
    if _lib is None:
        build()
        L = C.CDLL(_SO)
        L.orc_cov_fn.restype = C.c_double
        L.orc_cov_fn.argtypes = [C.POINTER(Term), C.c_int, C.c_double]
        L.orc_prior.restype = C.c_double
        L.orc_invgamma_logpdf.restype = C.c_double
        L.orc_invgamma_logpdf.argtypes = [C.c_double] * 3
        _lib = L
    return _lib


def terms_array(terms):
    """terms: list of ("se", h, sigma) | ("white", sigma)"""
    arr = (Term * len(terms))()
    for i, t in enumerate(terms):
        if t[0] == "se":
            arr[i] = Term(SE, float(t[1]), float(t[2]), 0.0)
        elif t[0] == "white":
            arr[i] = Term(WHITE, float(t[1]), 0.0, 0.0)
        else:
            raise ValueError(t[0])
    return arr


def _d(a):
    a = np.ascontiguousarray(a, dtype=np.float64)
    return a, a.ctypes.data_as(_dp)


def _xsoa(x):
    """x: (n,) or (n, dim) -> SoA [dim][n] contiguous, dim"""
    x = np.asarray(x, dtype=np.float64)
    if x.ndim == 1:
        x = x[:, None]
    return np.ascontiguousarray(x.T), x.shape[1]


def cov_fn(terms, d):
    return lib().orc_cov_fn(terms_array(terms), len(terms), float(d))


def build_cov(x, terms, nugget):
    xs, dim = _xsoa(x)
    n = xs.shape[1]
    K = np.empty((n, n), dtype=np.float64)
    lib().orc_build_cov(C.c_int(n), C.c_int(dim), xs.ctypes.data_as(_dp), terms_array(terms),
                        C.c_int(len(terms)), C.c_double(nugget), K.ctypes.data_as(_dp))
    return K.T  # column-major buffer viewed as (row, col)


def potrf_lower(K):
    """K (n,n) symmetric -> (L lower with zero upper, info)"""
    n = K.shape[0]
    a = np.array(K.T, dtype=np.float64, order="C")  # buffer = column-major of K
    info = lib().orc_potrf_lower(C.c_int(n), a.ctypes.data_as(_dp), C.c_int(n))
    lib().orc_zero_upper(C.c_int(n), a.ctypes.data_as(_dp), C.c_int(n))
    return a.T.copy(), info


def trsv_lower(L, b):
    n = L.shape[0]
    a = np.array(L.T, dtype=np.float64, order="C")
    xv = np.array(b, dtype=np.float64)
    lib().orc_trsv_lower(C.c_int(n), a.ctypes.data_as(_dp), C.c_int(n), xv.ctypes.data_as(_dp))
    return xv


def trsv_upper(U, b):
    n = U.shape[0]
    a = np.array(U.T, dtype=np.float64, order="C")
    xv = np.array(b, dtype=np.float64)
    lib().orc_trsv_upper(C.c_int(n), a.ctypes.data_as(_dp), C.c_int(n), xv.ctypes.data_as(_dp))
    return xv


def trsm_upper(U, B):
    m, nrhs = B.shape
    a = np.array(U.T, dtype=np.float64, order="C")
    bb = np.array(B.T, dtype=np.float64, order="C")
    lib().orc_trsm_upper(C.c_int(m), C.c_int(nrhs), a.ctypes.data_as(_dp), C.c_int(m),
                         bb.ctypes.data_as(_dp), C.c_int(m))
    return bb.T.copy()


def loglik(x, y, terms, beta=(), nugget=1e-3, long_double=False):
    xs, dim = _xsoa(x)
    n = xs.shape[1]
    yv, yp = _d(y)
    bv, bp = _d(np.asarray(beta, dtype=np.float64) if len(beta) else np.zeros(1))
    out = LL()
    if long_double:
        rc = lib().orc_loglik_ld(C.c_int(n), C.c_int(dim), xs.ctypes.data_as(_dp), yp,
                                 terms_array(terms), C.c_int(len(terms)), bp, C.c_int(len(beta)),
                                 C.c_double(nugget), C.byref(out))
    else:
        work = np.empty(n * n + n, dtype=np.float64)
        rc = lib().orc_loglik(C.c_int(n), C.c_int(dim), xs.ctypes.data_as(_dp), yp,
                              terms_array(terms), C.c_int(len(terms)), bp, C.c_int(len(beta)),
                              C.c_double(nugget), work.ctypes.data_as(_dp), C.byref(out))
    assert rc == 0
    return {k: getattr(out, k) for k, _ in LL._fields_}


def predict(x, y, terms, xstar, nugget=1e-6):
    xs, dim = _xsoa(x)
    n = xs.shape[1]
    xt, dim2 = _xsoa(xstar)
    assert dim == dim2
    m = xt.shape[1]
    yv, yp = _d(y)
    mean = np.empty(m)
    var = np.empty(m)
    work = np.empty(n * n + 2 * n, dtype=np.float64)
    info = C.c_int(0)
    lib().orc_predict(C.c_int(n), C.c_int(dim), xs.ctypes.data_as(_dp), yp, terms_array(terms),
                      C.c_int(len(terms)), C.c_double(nugget), C.c_int(m), xt.ctypes.data_as(_dp),
                      mean.ctypes.data_as(_dp), var.ctypes.data_as(_dp), work.ctypes.data_as(_dp),
                      C.byref(info))
    return mean, var, info.value


def nparams(terms):
    return sum(2 if t[0] == "se" else 1 for t in terms)


def prior(terms, ig):
    igv, igp = _d(ig)
    return lib().orc_prior(terms_array(terms), C.c_int(len(terms)), igp)


def mh_run(x, y, init_terms, ig, delta, z, u, beta=(), nugget=1e-3, formula=0, evals_per_iter=1,
           with_prob=False):
    """z: (iters, P) standard normals; u: (iters,) uniforms; ig: (P, 2) inverse-gamma (shape, scale).
    with_prob: also return the acceptance probabilities a_i = min(1, exp(new - old)) and the number
    of iterations completed (< iters when a proposal was not positive definite: rc = LAPACK info)."""
    xs, dim = _xsoa(x)
    n = xs.shape[1]
    yv, yp = _d(y)
    P = nparams(init_terms)
    zv, zp = _d(z)
    uv, up = _d(u)
    iters = uv.shape[0]
    assert zv.shape == (iters, P)
    igv, igp = _d(ig)
    assert igv.shape == (P, 2)
    bv, bp = _d(np.asarray(beta, dtype=np.float64) if len(beta) else np.zeros(1))
    states = np.empty((iters, P))
    ll = np.empty(iters)
    acc = np.empty(iters, dtype=np.int32)
    prob = np.full(iters, np.nan)
    done = C.c_int(0)
    rc = lib().orc_mh_run_a(C.c_int(n), C.c_int(dim), xs.ctypes.data_as(_dp), yp,
                            terms_array(init_terms), C.c_int(len(init_terms)), bp,
                            C.c_int(len(beta)), C.c_double(nugget), igp, C.c_double(delta),
                            C.c_int(iters), zp, up, C.c_int(formula), C.c_int(evals_per_iter),
                            states.ctypes.data_as(_dp), ll.ctypes.data_as(_dp),
                            acc.ctypes.data_as(_ip), prob.ctypes.data_as(_dp), C.byref(done))
    if with_prob:
        return states, ll, acc, rc, prob, done.value
    return states, ll, acc, rc


def mcmc_sample(x, y, init_terms, beta0, ig, delta, z, u, prior_mean, prior_var, nugget=1e-3,
                mean_nugget=1e-6, formula=0):
    """Mcmc.sample (Mcmc.scala:63-76), host-stepped: step = KernelParameters.sample compose
    MeanParameters.sample, i.e. every iteration first draws the plane coefficients from their full
    conditional (samplePlane, MeanParameters.scala:74-103, at the CURRENT kernel parameters), then
    runs Breeze's metropolis on the kernel parameters with ll(next) and ll(t) both evaluated at the
    new coefficients (KernelParameters.scala:231-246).  Proposal / prior as in mh_run.
    z: (iters, P + d) -- [0, P) the proposal normals, [P, P + d) the Gibbs normals; u: (iters,).
    Returns states (iters, P + d) [kernel parameters, beta], ll (iters,), accept (iters,)."""
    P = nparams(init_terms)
    beta = np.array(beta0, dtype=np.float64)
    d = len(beta)
    cur = [tuple(t) for t in init_terms]
    iters = len(u)
    states = np.empty((iters, P + d)); lls = np.empty(iters); acc = np.zeros(iters, dtype=np.int32)

    def total(terms, b):
        r = loglik(x, y, terms, beta=tuple(b), nugget=nugget)
        assert r["info"] == 0
        return prior(terms, ig) + (r["ll_gpml"] if formula else r["ll_ref"])

    for it in range(iters):
        beta, _, _ = sample_plane(x, y, cur, prior_mean, prior_var, z[it, P:], nugget=mean_nugget)
        nxt, p = [], 0
        for t in cur:
            if t[0] == "se":
                nxt.append(("se", t[1] * math.exp(delta * z[it, p]), t[2] * math.exp(delta * z[it, p + 1]))); p += 2
            else:
                nxt.append(("white", t[1] * math.exp(delta * z[it, p]))); p += 1
        new_ll, old_ll = total(nxt, beta), total(cur, beta)
        dl = new_ll - old_ll
        a = min(1.0, math.exp(dl)) if dl == dl else float("nan")
        if u[it] < a:
            cur, old_ll, acc[it] = nxt, new_ll, 1
        flat = []
        for t in cur:
            flat += list(t[1:])
        states[it, :P] = flat; states[it, P:] = beta; lls[it] = old_ll
    return states, lls, acc


# ---- "next" rows of the scope table: numpy restatements, small cases only -------
def design_matrix(x):
    """MeanParameters.makeDesignMatrix (MeanParameters.scala:55-65), INCLUDING its layout quirk:
    `new DenseMatrix(n, m, xs.flatten.toArray)` reads the row-major flattened coordinates as a
    column-major n x m matrix, so for 2-D locations the coordinate columns are interleaved."""
    x = np.asarray(x, dtype=np.float64)
    if x.ndim == 1:
        x = x[:, None]
    n, m = x.shape
    res = x.reshape(-1).reshape((m, n)).T          # column-major view of the flattened data
    return np.hstack([np.ones((n, 1)), res])


def sample_plane(x, y, terms, prior_mean, prior_var, z, nugget=1e-6):
    """MeanParameters.samplePlane (MeanParameters.scala:74-103), line by line; `\\` on a square
    Breeze matrix is an LU solve (np.linalg.solve).  z: the standard normals of `Gaussian(0,1).draw`.
    Returns (beta, prec, mean)."""
    X = design_matrix(x)
    K = build_cov(x, terms, nugget)
    L, _info = potrf_lower(K)
    m = X.shape[1]
    prior_prec = np.eye(m) / prior_var
    prec = X.T @ np.linalg.solve(K, X) + prior_prec
    u = trsv_lower(L, np.asarray(y, dtype=np.float64))
    mean = np.linalg.solve(prec, prior_prec @ np.full(m, float(prior_mean)) + X.T @ u)
    root = np.linalg.cholesky(prec)
    beta = mean + np.linalg.solve(root, np.asarray(z, dtype=np.float64))
    return beta, prec, mean


def draw(x, terms, z, beta=(), nugget=1e-3):
    """GaussianProcess.draw (GaussianProcess.scala:67-83): mu + V sqrt(Lambda) z from eigSym.  The
    device path uses mu + L z (Cholesky): equal in distribution, not bitwise, so
    tests compare covariances / the Cholesky form, which this also returns."""
    xx = np.asarray(x, dtype=np.float64)
    if xx.ndim == 1:
        xx = xx[:, None]
    K = build_cov(x, terms, nugget)
    mu = np.zeros(len(xx))
    if len(beta):
        mu = beta[0] + xx @ np.asarray(beta[1:], dtype=np.float64)
    w, V = np.linalg.eigh(K)
    zz = np.asarray(z, dtype=np.float64)
    return mu + V @ (np.sqrt(w) * zz), mu + potrf_lower(K)[0] @ zz


def distance_matrix(x):
    xx = np.asarray(x, dtype=np.float64)
    if xx.ndim == 1:
        xx = xx[:, None]
    d2 = (xx[:, None, 0] - xx[None, :, 0]) ** 2
    if xx.shape[1] == 2:
        d2 = d2 + (xx[:, None, 1] - xx[None, :, 1]) ** 2
    return np.sqrt(d2)


def mll_gradient(x, y, theta, beta=(), nugget=1e-3, mode="as_executed"):
    """GaussianProcess.mllGradient (GaussianProcess.scala:166-187) with KernelParameters.tangentMatrix /
    gradient (KernelParameters.scala:270-300), line by line.  theta: terms holding the parameters as
    sampleHmc passes them (unconstrained / log scale).
      as_executed: kxx = buildCov(theta AS IS) + nugget (:171-176); tangent entries exp(-d^2/s^2) + 1,
                   -(2h/s^3) exp(-d^2/s^2) + 1 and 2 (white), h = exp(hu), s = exp(su) (:272-281);
                   alpha = kxx \\ ys (LU); grad = 0.5 sum(diag(alpha alpha^T g - kxx \\ g)).
      analytic:    the gradient of ll_gpml with respect to the log parameters (kxx from exp(theta))."""
    D = distance_matrix(x)
    n = D.shape[0]
    xx = np.asarray(x, dtype=np.float64).reshape(n, -1)
    ys = np.asarray(y, dtype=np.float64).copy()
    if len(beta):
        ys = ys - (beta[0] + xx @ np.asarray(beta[1:], dtype=np.float64))
    if mode == "as_executed":
        kterms = list(theta)
    else:
        kterms = [(t[0], math.exp(t[1]), math.exp(t[2])) if t[0] == "se" else (t[0], math.exp(t[1])) for t in theta]
    K = build_cov(x, kterms, nugget)
    tangents = []
    for t in theta:
        if t[0] == "se":
            h, s = math.exp(t[1]), math.exp(t[2])
            E = np.exp(-(D * D) / (s * s))
            if mode == "as_executed":
                tangents += [E + 1.0, -(2 * h / (s * s * s)) * E + 1.0]
            else:
                tangents += [h * E, h * E * 2.0 * (D * D) / (s * s)]
        else:
            if mode == "as_executed":
                tangents.append(np.full((n, n), 2.0))
            else:
                tangents.append(math.exp(t[1]) * (D == 0.0))
    alpha = np.linalg.solve(K, ys)
    return np.array([0.5 * np.sum(np.diag(np.outer(alpha, alpha) @ g - np.linalg.solve(K, g))) for g in tangents])
