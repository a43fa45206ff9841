"""Timing of gpb_mll_gradient by phase (build / factor + L^-T rows / alpha + contraction + finalize)."""
import math
import sys

sys.path.insert(0, "gaussian-processes_b200")
import numpy as np
from gpb200 import _lib as G

ctx = G.Context(0)
cases = ((4096, 1), (4096, 8), (8192, 1)) if len(sys.argv) < 2 else ((int(sys.argv[1]), int(sys.argv[2])),)
for n, batch in cases:
    rng = np.random.default_rng(1)
    x = np.sort(rng.uniform(0, n / 10, n)); y = rng.standard_normal(n)
    ds = ctx.dataset(x, y)
    th = [[(0, math.log(3.0), math.log(5.5)), (1, math.log(0.5), 0.0)]] * batch
    for _ in range(3):
        g, info = ds.mll_gradient(th, mode=G.GRAD_ANALYTIC)
    t = ctx.last_timing()
    print(f"n={n} batch={batch} info={info[0]} build={t[0]:.3f} ms factor+V={t[1]:.3f} ms grad kernels={t[2]:.3f} ms  "
          f"factor TF/s={(2*batch*n**3/3)/t[1]/1e9:.2f} contract TF/s={(batch*n**3/3)/t[2]/1e9:.2f}")
    ds.close()
