"""Mirror of the reference's objects for the hot path (see package docstring).

Citations are file:line under /root/reference/core/src/main/scala/gp/.
"""
import math
from dataclasses import dataclass
from typing import Callable, List, Sequence, Tuple

import numpy as np

from . import _lib as G

# --------------------------------------------------------------------------------------------
# Location.scala:5-33
# --------------------------------------------------------------------------------------------


@dataclass(frozen=True)
class One:
    x: float

    def toVector(self):
        return [self.x]


@dataclass(frozen=True)
class Two:
    x: float
    y: float

    def toVector(self):
        return [self.x, self.y]


class Location:
    tol = 1e-3  # Location.scala:16

    @staticmethod
    def euclidean(a, b) -> float:
        """Location.scala:27-33.  Scalar host helper (used for single pairs, e.g. kyy);
        matrices of distances are built on the device by GaussianProcess.distanceMatrix."""
        if isinstance(a, One) and isinstance(b, One):
            return math.sqrt((a.x - b.x) ** 2)
        if isinstance(a, Two) and isinstance(b, Two):
            return math.sqrt((a.x - b.x) ** 2 + (a.y - b.y) ** 2)
        raise Exception("Dimension mismatch")  # Location.scala:32


def _require_euclidean(dist):
    # dist is an arbitrary closure in the reference; only Location.euclidean ships.
    if dist is not Location.euclidean and getattr(dist, "__func__", dist) is not Location.euclidean:
        raise NotImplementedError(
            "only Location.euclidean is supported on the device (no CPU fallback by design)")


def _pack_locations(xs: Sequence) -> np.ndarray:
    if len(xs) == 0:
        raise ValueError("empty location vector")
    if all(isinstance(p, One) for p in xs):
        return np.array([[p.x] for p in xs], dtype=np.float64)
    if all(isinstance(p, Two) for p in xs):
        return np.array([[p.x, p.y] for p in xs], dtype=np.float64)
    raise Exception("Dimension mismatch")  # Location.scala:32


# --------------------------------------------------------------------------------------------
# KernelParameters.scala:11-85  (ADT only; HMC drivers are out of scope)
# --------------------------------------------------------------------------------------------


@dataclass(frozen=True)
class SquaredExp:
    h: float
    sigma: float

    def map(self, f):
        return SquaredExp(f(self.h), f(self.sigma))

    def toList(self):
        return [self.h, self.sigma]

    def toMap(self):
        return {"h": self.h, "sigma": self.sigma}

    def add(self, that):
        if isinstance(that, SquaredExp):
            return SquaredExp(self.h + that.h, self.sigma + that.sigma)
        raise Exception()


@dataclass(frozen=True)
class White:
    sigma: float

    def map(self, f):
        return White(f(self.sigma))

    def toList(self):
        return [self.sigma]

    def toMap(self):
        return {"sigma_y": self.sigma}

    def add(self, that):
        if isinstance(that, White):
            return White(self.sigma + that.sigma)
        raise Exception()


@dataclass(frozen=True)
class Matern:
    sigma: float
    nu: float
    l: float

    def map(self, f):
        return Matern(f(self.sigma), f(self.nu), f(self.l))

    def toList(self):
        return [self.sigma, self.nu, self.l]

    def toMap(self):
        return {"sigma": self.sigma, "nu": self.nu, "l": self.l}


def _terms(ps) -> List[Tuple[int, float, float]]:
    out = []
    for p in ps:
        if isinstance(p, SquaredExp):
            out.append((G.KIND_SE, p.h, p.sigma))
        elif isinstance(p, White):
            out.append((G.KIND_WHITE, p.sigma, 0.0))
        elif isinstance(p, Matern):
            out.append((G.KIND_MATERN, p.sigma, p.nu))  # -> GPB_E_UNSUPPORTED at the boundary
        else:
            raise TypeError(p)
    return out


# --------------------------------------------------------------------------------------------
# MeanParameters.scala:8-46, MeanFunction.scala:7-17
# --------------------------------------------------------------------------------------------


@dataclass(frozen=True)
class _Zero:
    def map(self, f):
        return self

    def toList(self):
        return []

    def toMap(self):
        return {}


Zero = _Zero()


@dataclass(frozen=True)
class Plane:
    beta: tuple

    def __init__(self, beta):
        object.__setattr__(self, "beta", tuple(float(b) for b in beta))

    def map(self, f):
        return Plane([f(b) for b in self.beta])

    def toList(self):
        return list(self.beta)

    def toMap(self):
        return {f"beta_{i}": b for i, b in enumerate(self.beta)}


class MeanParameters:
    zero = Zero

    @staticmethod
    def plane(beta):
        return Plane(beta)

    @staticmethod
    def vectorToParams(p, pActual):
        return Zero if p is Zero else Plane(pActual)

    @staticmethod
    def makeDesignMatrix(xs):
        """MeanParameters.scala:55-65, including its layout: `new DenseMatrix(n, m, xs.flatten)`
        reads the row-major flattened coordinates column-major (interleaved columns for Two)."""
        x = _pack_locations(xs)
        if x.ndim == 1:
            x = x[:, None]
        n, m = x.shape
        res = x.reshape(-1).reshape((m, n)).T
        return np.hstack([np.ones((n, 1)), res])

    @staticmethod
    def samplePlane(prior, obs, xs, dist, p, rng=None):
        """MeanParameters.scala:74-103 -> gpb_sample_plane (build + Cholesky + the design-matrix and
        observation rows solved in the same pass on the device; the (dim+1)^2 algebra on the host)."""
        _require_euclidean(dist)
        if len(xs) != len(obs):
            raise Exception("Dimension mismatch")  # Breeze: kxx is |xs| x |xs|, y has |obs| entries
        rng = np.random.default_rng() if rng is None else rng
        X = MeanParameters.makeDesignMatrix(xs)
        ds = default_context().dataset(_pack_locations(xs), np.array([d.y for d in obs], dtype=np.float64))
        try:
            z = rng.standard_normal(X.shape[1])
            beta, info = ds.sample_plane(_terms(p.kernelParameters), X, prior.mean, prior.variance, z,
                                         nugget=1e-6)
        finally:
            ds.close()
        if info != 0:
            raise NotConvergedException(info)
        return Plane(beta)

    @staticmethod
    def sample(ys, xs, dist, prior):
        """MeanParameters.scala:111-121.  Zero => identity.  Plane => Gibbs step samplePlane."""
        def step(p, rng=None):
            if p.meanParameters is Zero:
                return p
            return p.copy(meanParameters=MeanParameters.samplePlane(prior, ys, xs, dist, p, rng))
        return step


class MeanFunction:
    @staticmethod
    def plane(beta):
        def f(loc):
            if isinstance(loc, One):
                return beta[0] + loc.x * beta[1]
            return beta[0] + loc.x * beta[1] + loc.y * beta[2]
        return f

    @staticmethod
    def zero(loc):
        return 0.0

    @staticmethod
    def apply(p):
        return MeanFunction.zero if p is Zero else MeanFunction.plane(p.beta)


# --------------------------------------------------------------------------------------------
# device context shared by the module-level API (one per process; one process per GPU)
# --------------------------------------------------------------------------------------------
_ctx = None
_device = 0


def set_default_device(device: int):
    global _ctx, _device
    if _ctx is not None:
        _ctx.close()
        _ctx = None
    _device = int(device)


def default_context() -> G.Context:
    global _ctx
    if _ctx is None:
        _ctx = G.Context(_device)
    return _ctx


class NotConvergedException(Exception):
    """Stands in for breeze.linalg.NotConvergedException thrown by `cholesky` on a non-PD matrix
    (SURVEY.md 0.2l): a bad proposal raises, it is not silently rejected."""

    def __init__(self, info):
        super().__init__(f"Cholesky: leading minor {info} is not positive definite")
        self.info = info


@dataclass(frozen=True)
class Gaussian:
    """breeze.stats.distributions.Gaussian(mu, sigma) as returned by Predict.fit."""
    mean: float
    sigma: float

    @property
    def variance(self):
        return self.sigma * self.sigma


# --------------------------------------------------------------------------------------------
# KernelFunction.scala
# --------------------------------------------------------------------------------------------


class _CovFn:
    """The Double => Double closure KernelFunction.apply returns (KernelFunction.scala:47-55).
    Callable on a scalar distance like the reference's; carries its parameters so that the
    matrix-level builders can run it on the device."""

    def __init__(self, ps):
        self.ps = list(ps)
        self.terms = _terms(ps)

    def __call__(self, d: float) -> float:
        s = 0.0
        for p in self.ps:
            if isinstance(p, SquaredExp):
                s = s + p.h * math.exp(-(d * d) / (p.sigma * p.sigma))
            elif isinstance(p, White):
                s = s + (p.sigma if d == 0 else 0.0)
            else:
                raise NotImplementedError("Matern kernel is out of scope (SURVEY.md 2)")
        return s


class KernelFunction:
    @staticmethod
    def squaredExponential(h, sigma):
        return _CovFn([SquaredExp(h, sigma)])

    @staticmethod
    def white(sigma):
        return _CovFn([White(sigma)])

    @staticmethod
    def apply(ps):
        return _CovFn(ps)

    @staticmethod
    def buildCov(xs, covFn, dist):
        """KernelFunction.scala:84-89 -> gpb_build_cov.  Full symmetric n x n ndarray."""
        _require_euclidean(dist)
        if not isinstance(covFn, _CovFn):
            raise NotImplementedError("covFn must come from KernelFunction.apply (no CPU fallback)")
        return default_context().build_cov(_pack_locations(xs), covFn.terms, 0.0)


# --------------------------------------------------------------------------------------------
# GaussianProcess.scala
# --------------------------------------------------------------------------------------------


@dataclass(frozen=True)
class Parameters:
    meanParameters: object
    kernelParameters: tuple

    def __init__(self, meanParameters, kernelParameters):
        object.__setattr__(self, "meanParameters", meanParameters)
        object.__setattr__(self, "kernelParameters", tuple(kernelParameters))

    def map(self, f):
        return Parameters(self.meanParameters.map(f), [k.map(f) for k in self.kernelParameters])

    def toList(self):  # GaussianProcess.scala:24-25: mean first, then kernel terms in order
        out = list(self.meanParameters.toList())
        for k in self.kernelParameters:
            out.extend(k.toList())
        return out

    def toMap(self):
        out = dict(self.meanParameters.toMap())
        for k in self.kernelParameters:
            out.update(k.toMap())
        return out

    def copy(self, kernelParameters=None, meanParameters=None):
        return Parameters(self.meanParameters if meanParameters is None else meanParameters,
                          self.kernelParameters if kernelParameters is None else kernelParameters)


@dataclass(frozen=True)
class Data:
    x: object
    y: float


class GaussianProcess:
    Parameters = Parameters
    Data = Data
    formula = G.FORMULA_AS_PARSED  # what GaussianProcess.scala:126 executes

    @staticmethod
    def distanceMatrix(xs, dist):
        """GaussianProcess.scala:50-61 -> gpb_distance_matrix."""
        _require_euclidean(dist)
        return default_context().distance_matrix(_pack_locations(xs))

    @staticmethod
    def draw(xs, dist, p, rng=None):
        """GaussianProcess.scala:67-83 -> gpb_draw: mu + root z, nugget 1e-3.  Cholesky root instead
        of the reference's eigSym root: equal in distribution, not bitwise."""
        _require_euclidean(dist)
        rng = np.random.default_rng() if rng is None else rng
        x = _pack_locations(xs)
        mf = MeanFunction.apply(p.meanParameters)
        mu = np.array([mf(loc) for loc in xs], dtype=np.float64)
        out, info = default_context().draw(x, _terms(p.kernelParameters), rng.standard_normal(len(xs)), mu, nugget=1e-3)
        if info != 0:
            raise NotConvergedException(info)
        return out

    @staticmethod
    def loglikelihood(observed, dist):
        """GaussianProcess.scala:109-127.  Curried like the reference: the observations are
        uploaded when the outer function is applied; the returned closure ships only the
        hyper-parameters and reads back 3 doubles."""
        _require_euclidean(dist)
        x = _pack_locations([d.x for d in observed])
        y = np.array([d.y for d in observed], dtype=np.float64)
        ds = default_context().dataset(x, y)

        def ll(p: Parameters, formula=None) -> float:
            return GaussianProcess._ll_batch(ds, [p], formula)[0]

        ll.dataset = ds
        ll.batch = lambda ps, formula=None: GaussianProcess._ll_batch(ds, ps, formula)
        return ll

    gradient_mode = G.GRAD_AS_EXECUTED  # what GaussianProcess.scala:166-187 executes; G.GRAD_ANALYTIC = the true gradient

    @staticmethod
    def mllGradient(observed, dist):
        """GaussianProcess.scala:166-187 -> gpb_mll_gradient.  Curried like the reference; p holds the
        UNCONSTRAINED kernel parameters (that is what sampleHmc passes, KernelParameters.scala:141-144).
        Returns the gradient per kernel parameter in toList order."""
        _require_euclidean(dist)
        x = _pack_locations([d.x for d in observed])
        y = np.array([d.y for d in observed], dtype=np.float64)
        ds = default_context().dataset(x, y)

        def batch(ps, mode=None):
            mode = GaussianProcess.gradient_mode if mode is None else mode
            betas = [list(p.meanParameters.toList()) for p in ps]
            g, info = ds.mll_gradient([_terms(p.kernelParameters) for p in ps], betas, nugget=1e-3, mode=mode)
            for i in info:
                if i != 0:
                    raise NotConvergedException(int(i))  # the reference's LU would carry on (include/gpb200.h)
            return g

        def grad(p: Parameters, mode=None):
            return batch([p], mode)[0]

        grad.dataset = ds
        grad.batch = batch
        return grad

    @staticmethod
    def _ll_batch(ds, ps, formula=None):
        formula = GaussianProcess.formula if formula is None else formula
        betas = [list(p.meanParameters.toList()) for p in ps]
        rs = ds.loglik_batch([_terms(p.kernelParameters) for p in ps], betas, nugget=1e-3)
        out = []
        for r in rs:
            if r.info != 0:
                raise NotConvergedException(r.info)
            out.append(G.ll_value(r, formula))
        return out


# --------------------------------------------------------------------------------------------
# KernelParameters object: paramsToArray / arrayToParams / sample (MH)
# --------------------------------------------------------------------------------------------


class KernelParameters:
    se = staticmethod(lambda h, sigma: SquaredExp(h, sigma))
    white = staticmethod(lambda sigma: White(sigma))
    matern = staticmethod(lambda sigma, nu, l: Matern(sigma, nu, l))

    @staticmethod
    def paramsToArray(p: Parameters):  # KernelParameters.scala:88-89
        return np.array(p.toList(), dtype=np.float64)

    @staticmethod
    def vectorToParams(ps, values):  # KernelParameters.scala:306-...
        out, k = [], 0
        for p in ps:
            n = len(p.toList())
            out.append(type(p)(*values[k:k + n]))
            k += n
        return out

    @staticmethod
    def arrayToParams(p: Parameters, pActual):  # KernelParameters.scala:94-101
        n = len(p.meanParameters.toList())
        vals = list(pActual)
        return Parameters(MeanParameters.vectorToParams(p.meanParameters, vals[:n]),
                          KernelParameters.vectorToParams(p.kernelParameters, vals[n:]))

    @staticmethod
    def unconstrainParams(ps):  # KernelParameters.scala:251-255
        return [p.map(math.log) for p in ps]

    @staticmethod
    def constrainParams(ps):  # KernelParameters.scala:260-264
        return [p.map(math.exp) for p in ps]

    @staticmethod
    def sampleHmc(*args, **kwargs):
        """KernelParameters.scala:121-156 -- out of the hot-path scope: host control flow kept,
        frozen, in gpb200/extras_hmc.py around the in-scope device calls (loglikelihood, mllGradient)."""
        from . import extras_hmc
        return extras_hmc.sampleHmc(*args, **kwargs)

    @staticmethod
    def sampleEhmc(*args, **kwargs):
        """KernelParameters.scala:171-198 -- see sampleHmc."""
        from . import extras_hmc
        return extras_hmc.sampleEhmc(*args, **kwargs)

    @staticmethod
    def sample(obs, dist, prior: Callable, prop: Callable):
        """KernelParameters.scala:231-246 + Breeze metropolis: returns step(p, rng) -> p'.
        prior / prop stay host closures (prop(kernelParameters, rng) -> kernelParameters).
        ll(next) and ll(t) are both evaluated every step like Breeze does -- as ONE batch of two
        proposals on the device."""
        llf = GaussianProcess.loglikelihood(obs, dist)

        def step(p: Parameters, rng: np.random.Generator) -> Parameters:
            nxt = p.copy(kernelParameters=prop(p.kernelParameters, rng))
            lln, llo = llf.batch([nxt, p])
            new_ll = prior(nxt.kernelParameters) + lln
            old_ll = prior(p.kernelParameters) + llo
            # Breeze metropolis on the JVM: scala.math.min propagates NaN and `u < NaN` is false, so a NaN
            # on either side rejects (Python's min(1.0, nan) would return 1.0 and always accept)
            dl = new_ll - old_ll
            a = min(1.0, math.exp(min(dl, 700.0))) if dl == dl else float("nan")
            u = rng.uniform()
            return nxt if u < a else p

        return step


# --------------------------------------------------------------------------------------------
# Predict.scala
# --------------------------------------------------------------------------------------------


class Predict:
    @staticmethod
    def forwardSolve(A, y):
        """Predict.scala:40-45 (dtrsv L,N,N) -> gpb_trsm_lower."""
        return default_context().trsm_lower(A, y)

    @staticmethod
    def fit(newXs, observed, dist, p: Parameters):
        """Predict.scala:57-94 -> gpb_predict.  Vector[(Location, Gaussian(mean, sqrt(var)))]."""
        _require_euclidean(dist)
        x = _pack_locations([d.x for d in observed])
        y = np.array([d.y for d in observed], dtype=np.float64)
        ds = default_context().dataset(x, y)
        try:
            mean, var, info = ds.predict(_terms(p.kernelParameters), _pack_locations(newXs), nugget=1e-6)
        finally:
            ds.close()
        if info != 0:
            raise NotConvergedException(info)
        return [(xs, Gaussian(float(m), math.sqrt(v) if v >= 0 else float("nan")))
                for xs, m, v in zip(newXs, mean, var)]

    @staticmethod
    def buildData(xs, ys):
        return [Data(x, y) for x, y in zip(xs, ys)]


# --------------------------------------------------------------------------------------------
# Mcmc.scala
# --------------------------------------------------------------------------------------------


class Mcmc:
    @staticmethod
    def backSolve(A, y):
        """Mcmc.scala:35-41 (vector) / 52-58 (matrix) -> gpb_trsm_upper."""
        return default_context().trsm_upper(A, y)

    @staticmethod
    def sample(ys, priorKern, priorMean, propKern, dist, initP, rng=None):
        """Mcmc.scala:63-76: MarkovChain(initP)(kernelStep compose meanStep) as a generator.
        The mean step runs first (Kleisli compose)."""
        rng = np.random.default_rng() if rng is None else rng
        xs = []
        for d in ys:  # ys.map(_.x).distinct
            if d.x not in xs:
                xs.append(d.x)
        mean_step = MeanParameters.sample(ys, xs, dist, priorMean)
        kern_step = KernelParameters.sample(ys, dist, priorKern, propKern)

        def chain():
            p = initP
            while True:
                yield p
                p = kern_step(mean_step(p, rng), rng)

        return chain()
