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
    def sampleHmc(ys, dist, init: Parameters, prior, m, l, eps, rng=None):
        """KernelParameters.scala:121-156: HMC over the unconstrained parameters.  ll = prior + device
        log-likelihood of the constrained parameters; grad = device mllGradient + prior gradient (zipped
        like the reference).  Generator of theta arrays (MarkovChain(theta))."""
        rng = np.random.default_rng() if rng is None else rng
        ll, grad, theta = KernelParameters._hmc_target(ys, dist, init, prior)

        def chain():
            t = theta
            while True:
                yield t
                t, _ = Hmc.step(eps, l, m, ll, grad)(t, rng)

        return chain()

    @staticmethod
    def sampleEhmc(ys, dist, init: Parameters, prior, l0, warmupIterations, rng=None, k=2000):
        """KernelParameters.scala:171-198: empirical HMC (step size by dual averaging, leapfrog counts from
        the empirical longest-batch distribution); ll and grad as in sampleHmc.  k = Ehmc.sample's k."""
        ll, grad, theta = KernelParameters._hmc_target(ys, dist, init, prior)
        return Ehmc.sample(l0, np.eye(len(theta)), ll, grad, warmupIterations, theta, k=k, rng=rng)

    @staticmethod
    def _hmc_target(ys, dist, init: Parameters, prior):
        llf = GaussianProcess.loglikelihood(ys, dist)
        gradf = GaussianProcess.mllGradient(ys, dist)

        def ll(p):
            params = KernelParameters.arrayToParams(init, p)
            kp = KernelParameters.constrainParams(params.kernelParameters)
            return sum(pr.logPdf(pa) for pr, pa in zip(prior, p)) + llf(params.copy(kernelParameters=kp))

        last = {}

        def grad(p):
            # Hmc.leapfrog asks for the gradient at the end of one step and again at the start of the next
            # (Hmc.scala:30-39): the same point, and the device call is deterministic -- one-entry memo
            key = np.asarray(p, dtype=np.float64).tobytes()
            if last.get("key") != key:
                params = KernelParameters.arrayToParams(init, p)
                evalPrior = [pr.gradLogPdf(pa) for pr, pa in zip(prior, p)]
                last["key"], last["val"] = key, np.array([a + b for a, b in zip(gradf(params), evalPrior)])
                last["calls"] = last.get("calls", 0) + 1
            return last["val"].copy()

        grad.stats = last
        kp = KernelParameters.unconstrainParams(init.kernelParameters)
        return ll, grad, KernelParameters.paramsToArray(init.copy(kernelParameters=kp))

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
# Hmc.scala / GradDist.scala: host control flow around the device log-likelihood and gradient
# --------------------------------------------------------------------------------------------


class GradDist:
    """GradDist.scala:5-24: a univariate prior with logPdf and gradLogPdf (breeze Gamma(shape, scale) /
    Gaussian(mu, sigma)); the normal's gradient divides by sigma, as the reference does (:22)."""

    def __init__(self, logPdf, gradLogPdf):
        self.logPdf, self.gradLogPdf = logPdf, gradLogPdf

    @staticmethod
    def gamma(shape, scale):
        return GradDist(lambda x: (shape - 1.0) * math.log(x) - x / scale - (math.lgamma(shape) + shape * math.log(scale)),
                        lambda x: (shape - 1.0) / x - 1.0 / scale)

    @staticmethod
    def normal(mu, sigma):
        return GradDist(lambda x: -(x - mu) ** 2 / (2.0 * sigma * sigma) - math.log(sigma * math.sqrt(2.0 * math.pi)),
                        lambda x: -(x - mu) / sigma)


class Hmc:
    @staticmethod
    def priorPhi(m, rng):  # Hmc.scala:8-11: MultivariateGaussian(0, m) = chol(m) z
        m = np.asarray(m, dtype=np.float64)
        return np.linalg.cholesky(m) @ rng.standard_normal(m.shape[1])

    @staticmethod
    def phiHalfStep(eps, theta, phi, gradient):  # :13-17
        return phi + np.asarray(gradient(theta)) * 0.5 * eps

    @staticmethod
    def thetaStep(eps, theta, phi):  # :19-20
        return theta + eps * phi

    @staticmethod
    def leapfrog(eps, gradient, theta, phi):  # :30-39
        p1 = Hmc.phiHalfStep(eps, theta, phi, gradient)
        t1 = Hmc.thetaStep(eps, theta, p1)
        return t1, Hmc.phiHalfStep(eps, t1, p1, gradient)

    @staticmethod
    def leapfrogs(eps, gradient, l, theta, phi):  # :44-56
        with np.errstate(over="ignore", invalid="ignore"):
            for _ in range(l):
                theta, phi = Hmc.leapfrog(eps, gradient, theta, phi)
        return theta, phi

    @staticmethod
    def kinetic(phi):  # :59-68
        return 0.5 * float(np.sum(np.asarray(phi) ** 2))

    @staticmethod
    def logDensity(theta, phi, pos):  # :70-73
        return pos(theta) - Hmc.kinetic(phi)

    @staticmethod
    def logAcceptance(propTheta, propPhi, theta, phi, pos):  # :78-85
        a = Hmc.logDensity(propTheta, propPhi, pos) - Hmc.logDensity(theta, phi, pos)
        return -math.inf if math.isnan(a) else min(a, 0.0)

    @staticmethod
    def step(eps, l, m, pos, gradient):  # :87-105
        def run(theta, rng):
            theta = np.asarray(theta, dtype=np.float64)
            phi = Hmc.priorPhi(m, rng)
            t, p = Hmc.leapfrogs(eps, gradient, l, theta, phi)
            a = Hmc.logAcceptance(t, p, theta, phi, pos)
            u = rng.uniform()
            return (t, a) if math.log(u) < a else (theta, a)
        return run


def _jexp(x):
    """math.exp with the JVM's IEEE behaviour (Infinity instead of an exception)."""
    try:
        return math.exp(x)
    except OverflowError:
        return math.inf


class DualAverageState:
    """DualAveraging.scala:7-11."""

    def __init__(self, iter, theta, logeps, logepsbar, hm):
        self.iter, self.theta, self.logeps, self.logepsbar, self.hm = iter, theta, logeps, logepsbar, hm


class DualAverage:
    """DualAveraging.scala:19-127 (host control flow; pos / gradient are the device calls).
    as_executed = True mirrors the Scala literally: `step` builds its next state positionally, which puts
    (hm, logEps, logEpsBar) into the fields (logeps, logepsbar, hm) (:64), and `tuneStepsize` returns the
    log step size that Ehmc.sample then uses as the step size (:110-111) -- the tuner diverges and the
    chain stops moving.  as_executed = False is what the algorithm (Hoffman & Gelman dual averaging)
    intends: named fields, and exp(logepsbar) as the tuned step size."""
    as_executed = True

    @staticmethod
    def updateEps(m, mu, delta, acceptProb, k=0.75, gamma=0.05, t0=10.0):  # :26-46
        def run(hm0, logeps0, logepsbar0):
            md = float(m)
            ra = 1 / (md + t0)
            hm = (1 - ra) * hm0 + ra * (delta - acceptProb)
            logeps1 = mu - ((math.sqrt(md) * hm) / gamma)
            power = math.pow(md, -k)
            return hm, logeps1, power * logeps1 + (1.0 - power) * logepsbar0
        return run

    @staticmethod
    def step(l0, mu, m, pos, gradient, delta=0.65):  # :48-65
        def run(s, rng):
            eps = _jexp(s.logeps)
            t, a = Hmc.step(eps, l0, m, pos, gradient)(s.theta, rng)
            acceptProb = min(1.0, _jexp(a))
            hm, logEps, logEpsBar = DualAverage.updateEps(s.iter, mu, delta, acceptProb)(s.hm, s.logeps, s.logepsbar)
            if DualAverage.as_executed:
                return DualAverageState(s.iter + 1, t, hm, logEps, logEpsBar)
            return DualAverageState(s.iter + 1, t, logeps=logEps, logepsbar=logEpsBar, hm=hm)
        return run

    @staticmethod
    def findReasonableEpsilon(theta, phi, pos, gradient):  # :70-100
        eps = 1.0
        initTheta, initPhi = Hmc.leapfrog(eps, gradient, theta, phi)

        def prop(propTheta, propPhi):
            return Hmc.logAcceptance(propTheta, propPhi, theta, phi, pos)

        a = 1.0 if prop(initTheta, initPhi) > math.log(0.5) else -1.0
        thetaP, phiP, curEps, count = initTheta, initPhi, eps, 0
        while a * prop(thetaP, phiP) > -a * math.log(2.0) and count < 100:
            thetaP, phiP = Hmc.leapfrog(curEps, gradient, theta, phi)
            curEps, count = math.pow(2.0, a) * curEps, count + 1
        return curEps

    @staticmethod
    def tuneStepsize(k, params, l0, m, pos, gradient, delta=0.65, rng=None):  # :108-126
        rng = np.random.default_rng() if rng is None else rng
        phi = Hmc.priorPhi(m, rng)
        eps0 = DualAverage.findReasonableEpsilon(np.asarray(params, dtype=np.float64), phi, pos, gradient)
        mu = math.log(10 * eps0)
        s = DualAverageState(1, np.asarray(params, dtype=np.float64), math.log(eps0), 0.0, 0.0)
        kernel = DualAverage.step(l0, mu, m, pos, gradient, delta)
        for _ in range(k + 1):  # MarkovChain(init)(step).steps.drop(k).next: the (k+1)-th transition
            s = kernel(s, rng)
        return s.logepsbar if DualAverage.as_executed else _jexp(s.logepsbar)


class Ehmc:
    """Ehmc.scala:8-122.  as_executed = True keeps the argument order of the U-turn test in longestBatch
    (:40: isUTurn(t, p, theta)); False tests (t - theta) . p < 0 on the running trajectory."""
    as_executed = True

    @staticmethod
    def isUTurn(theta, thetaPlus, phiPlus):  # :9-26
        out = float(np.sum((np.asarray(thetaPlus) - np.asarray(theta)) * np.asarray(phiPlus)))
        return True if math.isnan(out) else out < 0

    @staticmethod
    def longestBatch(theta, phi, eps, gradient, L):  # :28-52
        thetaOut, phiOut, t, p, l = theta, phi, theta, phi, 1
        while not (Ehmc.isUTurn(t, p, theta) if Ehmc.as_executed else Ehmc.isUTurn(theta, t, p)):
            t1, p1 = Hmc.leapfrog(eps, gradient, t, p)
            if l == L:
                thetaOut, phiOut = t1, p1
            t, p, l = t1, p1, l + 1
        return thetaOut, phiOut, l

    @staticmethod
    def longestBatchStep(eps, l0, m, gradient, pos):  # :54-78
        def run(thetal, rng):
            phi = Hmc.priorPhi(m, rng)
            theta = np.asarray(thetal[0], dtype=np.float64)
            t, p, l = Ehmc.longestBatch(theta, phi, eps, gradient, l0)
            propTheta, propPhi = Hmc.leapfrogs(eps, gradient, l0 - l, t, p) if l < l0 else (t, p)
            u = rng.uniform()
            a = Hmc.logAcceptance(propTheta, propPhi, theta, phi, pos)
            return (propTheta, l) if math.log(u) < a else (theta, l)
        return run

    @staticmethod
    def discreteUniform(lo, hi, rng):  # :80-83
        return int(math.floor(rng.uniform() * (hi - lo) + lo))

    @staticmethod
    def empiricalLongestStep(eps, l0, m, pos, gradient, theta, k=2000, rng=None):  # :85-99
        rng = np.random.default_rng() if rng is None else rng
        kernel = Ehmc.longestBatchStep(eps, l0, m, gradient, pos)
        state, out = (np.asarray(theta, dtype=np.float64), l0), []
        for _ in range(k):
            state = kernel(state, rng)
            out.append(state[1])
        return out

    @staticmethod
    def sample(l0, m, pos, gradient, warmupIterations, initTheta, delta=0.65, k=2000, rng=None):  # :101-121
        rng = np.random.default_rng() if rng is None else rng
        initTheta = np.asarray(initTheta, dtype=np.float64)
        m = np.eye(initTheta.size)  # the reference shadows its argument with the identity (:111)
        eps = DualAverage.tuneStepsize(warmupIterations, initTheta, l0, m, pos, gradient, delta, rng)
        empiricalL = Ehmc.empiricalLongestStep(eps, l0, m, pos, gradient, initTheta, k, rng)

        def chain():
            theta = initTheta
            while True:
                yield theta
                l = empiricalL[Ehmc.discreteUniform(0, k, rng)]
                theta, _ = Hmc.step(eps, l, m, pos, gradient)(theta, rng)

        return chain()


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
