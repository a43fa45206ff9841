"""Scratch: phase timing (build / factor+solve / epilogue) for batched configs."""
import sys, os, math, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "gaussian-processes_b200"))
from gpb200 import _lib as G
ctx = G.Context(0)
rng = np.random.default_rng(3)
for n, B in [(1024, 256), (1024, 32), (1024, 8), (1024, 1), (4096, 1), (4096, 2), (4096, 8), (8192, 1), (8192, 8), (300, 256), (300, 2)]:
    x = rng.uniform(-10, 10, size=(n, 2)); y = rng.standard_normal(n)
    ds = ctx.dataset(x, y)
    props = [[(0, 3.0 * math.exp(0.05 * rng.standard_normal()), 5.5), (1, 0.5, 0.0)] for _ in range(B)]
    for _ in range(3): ds.loglik_batch(props)
    ts = []
    t0 = time.perf_counter()
    for _ in range(5):
        ds.loglik_batch(props); ts.append(ctx.last_timing())
    wall = (time.perf_counter() - t0) / 5
    t = np.mean(ts, axis=0)
    fl = B * (n**3 / 3 + n * n)
    print(f"n={n} B={B}: build {t[0]:.3f} ms  chol {t[1]:.3f} ms ({fl/t[1]*1e-9:.2f} TF)  epi {t[2]:.3f} ms  wall {wall*1e3:.3f} ms")
    ds.close()
