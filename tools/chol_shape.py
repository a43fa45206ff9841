"""A few log-likelihood batches at one shape (n, batch) -- the launch to profile with ncu."""
import math, os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "gaussian-processes_b200"))
from gpb200 import _lib as G
n, B = int(sys.argv[1]), int(sys.argv[2])
reps = int(sys.argv[3]) if len(sys.argv) > 3 else 4
ctx = G.Context(0)
rng = np.random.default_rng(3)
x = rng.uniform(-10, 10, size=(n, 2)); y = rng.standard_normal(n)
ds = ctx.dataset(x, y)
props = [[(0, 3.0 * math.exp(0.05 * rng.standard_normal()), 5.5), (1, 0.5, 0.0)] for _ in range(B)]
ts = []
for _ in range(reps):
    r = ds.loglik_batch(props); ts.append(ctx.last_timing())
print(f"n={n} B={B} timing(ms) build/chol/epi", np.round(np.median(ts, axis=0), 4).tolist(), "info", max(v.info for v in r))
