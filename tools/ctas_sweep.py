"""Scratch: factor+solve time for chain-bound vs work-bound launches (GPB_CTAS_PER_SM_RT=1|2 forces the mode)."""
import sys, os, math
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "gaussian-processes_b200"))
from gpb200 import _lib as G
ctx = G.Context(0)
rng = np.random.default_rng(3)
out = []
for n, B in [(300, 2), (300, 32), (300, 256), (1024, 1), (1024, 8), (1024, 16), (1024, 32), (1024, 64), (2048, 1), (2048, 4), (2048, 8), (2048, 16),
             (4096, 1), (4096, 2), (4096, 3), (4096, 4), (8192, 1), (8192, 2)]:
    x = rng.uniform(-10, 10, size=(n, 2)); y = rng.standard_normal(n)
    ds = ctx.dataset(x, y)
    props = [[(0, 3.0 * math.exp(0.05 * rng.standard_normal()), 5.5), (1, 0.5, 0.0)] for _ in range(B)]
    for _ in range(3): ds.loglik_batch(props)
    ts = []
    for _ in range(7):
        ds.loglik_batch(props); ts.append(ctx.last_timing()[1])
    out.append(f"{n}x{B}:{np.median(ts):.3f}")
    ds.close()
print("ctas", os.environ.get("GPB_CTAS_PER_SM_RT", "auto"), " ".join(out))
