"""gpb200 -- host-side mirror of the reference's Scala API for the GP inference hot path.

Same object / method names, argument meaning and error behaviour as
core/src/main/scala/gp/{Location,KernelParameters,MeanParameters,MeanFunction,KernelFunction,
GaussianProcess,Predict,Mcmc,Hmc,GradDist,Ehmc,DualAveraging}.scala, so that tests read like the reference's own.  Every numeric
method crosses the C ABI (include/gpb200.h) into libgpb200.so; nothing here computes the hot path
on the CPU.  The Scala binding a maintainer would add is in INTEGRATION.md / scala/.
"""
from .api import (  # noqa: F401
    One, Two, Location, SquaredExp, White, Matern, KernelParameters, Zero, Plane, MeanParameters,
    MeanFunction, KernelFunction, GaussianProcess, Predict, Mcmc, Gaussian, NotConvergedException,
    default_context, set_default_device,
)
# out-of-scope samplers around the in-scope gradient (frozen; see extras_hmc.py)
from .extras_hmc import Hmc, GradDist, Ehmc, DualAverage, DualAverageState  # noqa: F401,E402
from ._lib import GpbError, FORMULA_AS_PARSED, FORMULA_GPML, GRAD_AS_EXECUTED, GRAD_ANALYTIC  # noqa: F401
