"""OUT OF SCOPE for the hot path (SURVEY.md 2 marks Hmc.scala, Ehmc.scala, DualAveraging.scala, GradDist.scala
and the sampleHmc / sampleEhmc drivers out of scope) -- kept, FROZEN, for callers of
GaussianProcess.mllGradient (scope row f2), which is the only in-scope piece these samplers use.

Pure host control flow around two device calls (log-likelihood and gradient through the C ABI); nothing
here is part of the measured path, and it is not meant to grow.  Citations are file:line under
/root/reference/core/src/main/scala/gp/.
"""
import math
from dataclasses import dataclass

import numpy as np

from .api import GaussianProcess, KernelParameters, Parameters  # noqa: F401


def sampleHmc(ys, dist, init: Parameters, prior, m, l, eps, rng=None):
    """KernelParameters.scala:121-156: HMC over the unconstrained parameters.  ll = prior + device
    log-likelihood of the constrained parameters; grad = device mllGradient + prior gradient (zipped
    like the reference).  Generator of theta arrays (MarkovChain(theta))."""
    rng = np.random.default_rng() if rng is None else rng
    ll, grad, theta = _hmc_target(ys, dist, init, prior)

    def chain():
        t = theta
        while True:
            yield t
            t, _ = Hmc.step(eps, l, m, ll, grad)(t, rng)

    return chain()

def sampleEhmc(ys, dist, init: Parameters, prior, l0, warmupIterations, rng=None, k=2000):
    """KernelParameters.scala:171-198: empirical HMC (step size by dual averaging, leapfrog counts from
    the empirical longest-batch distribution); ll and grad as in sampleHmc.  k = Ehmc.sample's k."""
    ll, grad, theta = _hmc_target(ys, dist, init, prior)
    return Ehmc.sample(l0, np.eye(len(theta)), ll, grad, warmupIterations, theta, k=k, rng=rng)

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


