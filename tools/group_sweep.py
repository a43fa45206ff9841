"""Ticket-group sweep for one (n, batch): run with GPB_TICKET_GROUP set per process."""
import sys, os, math, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "gaussian-processes_b200"))
from gpb200 import _lib as G
n, B = int(sys.argv[1]), int(sys.argv[2])
ctx = G.Context(0)
rng = np.random.default_rng(3)
x = rng.uniform(-15, 15, size=(n, 2)); y = rng.standard_normal(n)
ds = ctx.dataset(x, y)
props = [[(0, 3.0 * math.exp(0.05 * rng.standard_normal()), 5.5), (0, 1.0, 0.8), (1, 0.5, 0.0)] for _ in range(B)]
ds.loglik_batch(props)
ts = []
for _ in range(3):
    rs = ds.loglik_batch(props); ts.append(ctx.last_timing())
t = np.mean(ts, axis=0)
fl = B * (n**3 / 3 + n * n)
print(f"G={os.environ.get('GPB_TICKET_GROUP','-')} n={n} B={B}: build {t[0]:.3f} ms chol {t[1]:.3f} ms ({fl/t[1]*1e-9:.2f} TF) bad={sum(r.info!=0 for r in rs)}")
