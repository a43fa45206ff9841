import sys, os, math
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "gaussian-processes_b200"))
from gpb200 import _lib as G
ctx = G.Context(0)
rng = np.random.default_rng(11)
for n, B, reps in [(4096, 1, 4), (1024, 64, 4), (4096, 8, 3)]:
    x = rng.uniform(-10, 10, size=(n, 2)); y = rng.standard_normal(n)
    ds = ctx.dataset(x, y)
    props = [[(0, 3.0 * math.exp(0.05 * rng.standard_normal()), 5.5), (1, 0.5, 0.0)] for _ in range(B)]
    bad = 0
    for rep in range(reps):
        got = [r.info for r in ds.loglik_batch(props)]
        bad += any(g != 0 for g in got)
    print(f"n={n} B={B}: {bad}/{reps} runs with info != 0, chol ms {ctx.last_timing()[1]:.3f}")
    ds.close()
