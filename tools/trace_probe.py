"""Where a block column of the chain goes (fused tasks of one launch): solve, X X^T, factorisation, publication."""
import sys, os, math
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "gaussian-processes_b200")); sys.path.insert(0, os.path.join(ROOT, "tools"))
import trace_analyze as TA
from gpb200 import _lib as G
n, B = int(sys.argv[1]), int(sys.argv[2])
ctx = G.Context(0)
rng = np.random.default_rng(3)
x = rng.uniform(-10, 10, size=(n, 2)); y = rng.standard_normal(n)
ds = ctx.dataset(x, y)
props = [[(0, 3.0, 5.5), (1, 0.5, 0.0)] for _ in range(B)]
for _ in range(2): ds.loglik_batch(props)
ctx.set_trace(1); ds.loglik_batch(props); print("timing", ctx.last_timing())
tr = ctx.get_trace().astype(np.int64); ctx.set_trace(0)
npad = (n + 63) // 64 * 64; nt = npad // 64; rt = nt + 1
dec = np.array([TA.decode(B, nt, rt, t) for t in range(len(tr))])
fused = (dec[:, 1] == dec[:, 2] + 1) & (dec[:, 1] < nt)
us = lambda a: a / 1965.0
t = tr[fused]
print("fused tasks", fused.sum())
for name, a, b in (("L^-T arrived -> solve done", 4, 5), ("solve done -> factor start (X->Cs, XX^T, barriers)", 5, 14), ("diag_factor", 14, 15), ("factor end -> release (publish, L write)", 15, 6)):
    d = us(t[:, b] - t[:, a]); print(f"  {name}: p50 {np.median(d):.2f} mean {d.mean():.2f} min {d.min():.2f}")
