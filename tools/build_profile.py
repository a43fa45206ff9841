"""A few batched log-likelihood calls (n, batch from argv) for profiling k_build."""
import os, sys, math
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "gaussian-processes_b200"))
from gpb200 import _lib as G
ctx = G.Context(0)
rng = np.random.default_rng(3)
n = int(sys.argv[1]) if len(sys.argv) > 1 else 1024
B = int(sys.argv[2]) if len(sys.argv) > 2 else 256
x = rng.uniform(-10, 10, size=(n, 2)); y = rng.standard_normal(n)
ds = ctx.dataset(x, y)
props = [[(0, 3.0 * math.exp(0.05 * rng.standard_normal()), 5.5), (1, 0.5, 0.0)] for _ in range(B)]
for _ in range(3):
    r = ds.loglik_batch(props)
print("timing", ctx.last_timing(), "info", r[0].info)
