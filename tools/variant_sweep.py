"""Factor+solve time per shape for the library selected by GPB200_LIB (one process per variant)."""
import math, os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "gaussian-processes_b200"))
from gpb200 import _lib as G
SHAPES = [(1024, 1), (1024, 32), (1024, 64), (1024, 256), (300, 256), (2048, 8), (2048, 64), (4096, 1), (4096, 2), (4096, 8), (8192, 8)]
if len(sys.argv) > 1:
    SHAPES = [tuple(int(v) for v in a.split("x")) for a in sys.argv[1:]]
ctx = G.Context(0)
rng = np.random.default_rng(3)
out = []
for n, B in SHAPES:
    x = rng.uniform(-10, 10, size=(n, 2)); y = rng.standard_normal(n)
    ds = ctx.dataset(x, y)
    props = [[(0, 3.0 * math.exp(0.05 * rng.standard_normal()), 5.5), (1, 0.5, 0.0)] for _ in range(B)]
    for _ in range(3): r = ds.loglik_batch(props)
    ts = []
    for _ in range(9):
        r = ds.loglik_batch(props); ts.append(ctx.last_timing()[1])
    out.append(f"{n}x{B}:{np.median(ts):.3f}ms(Q={sum(v.Q for v in r):.10e})")
    ds.close()
print(os.path.basename(os.environ.get("GPB200_LIB", "default")), " ".join(out))
