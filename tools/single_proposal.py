"""One-proposal log-likelihood calls at n=4096 (the chain-bound regime), for profiling."""
import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "gaussian-processes_b200"))
from gpb200 import _lib as G
ctx = G.Context(0)
rng = np.random.default_rng(3)
n = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
x = rng.uniform(-10, 10, size=(n, 2)); y = rng.standard_normal(n)
ds = ctx.dataset(x, y)
for _ in range(4):
    r = ds.loglik_batch([[(0, 3.0, 5.5), (1, 0.5, 0.0)]])
print("timing", ctx.last_timing(), "info", r[0].info)
