"""How close the device log-likelihood is to the CPU oracle (relative), for a few sizes."""
import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "gaussian-processes_b200"))
from gpb200 import _lib as G
from oracle import oracle as o
o.build()
ctx = G.Context(0)
T = [("se", 3.0, 5.5), ("white", 0.5)]
for n, dim in [(300, 1), (1024, 2), (2048, 2)]:
    rng = np.random.default_rng(n)
    x = rng.uniform(-10, 10, size=(n, dim))
    y = np.linalg.cholesky(o.build_cov(x, T, 1e-3)) @ rng.standard_normal(n)
    ds = ctx.dataset(x, y)
    r = ds.loglik_batch([[(0, 3.0, 5.5), (1, 0.5, 0.0)]])[0]
    ro = o.loglik(x, y, T)
    print(f"n={n}: rel err ll_gpml {abs(G.ll_value(r,1)-ro['ll_gpml'])/abs(ro['ll_gpml']):.2e}  ll_as_parsed {abs(G.ll_value(r,0)-ro['ll_ref'])/abs(ro['ll_ref']):.2e}  (tolerance 1e-10)")
    ds.close()
