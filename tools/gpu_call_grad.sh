#!/bin/bash
# scratch: gradient tests + a quick timing of the gradient at n=4096
python -m pytest tests/test_gpu_parity.py -x -q -m gpu -k "mll_gradient" 2>&1 | tail -15
python - <<'PY'
import sys, time, math
sys.path.insert(0, "gaussian-processes_b200")
import numpy as np
from gpb200 import _lib as G
ctx = G.Context(0)
for n, batch in ((4096, 1), (4096, 8), (8192, 1)):
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
PY
