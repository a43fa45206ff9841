"""Scratch: timeline of the rhs-row tasks (tile row rt-1) of a single proposal, from the schedule trace."""
import sys, os, math
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "gaussian-processes_b200"))
sys.path.insert(0, os.path.join(ROOT, "tools"))
from gpb200 import _lib as G
from trace_analyze import decode
n, B = int(sys.argv[1]), int(sys.argv[2])
ghz = 1.965
ctx = G.Context(0)
rng = np.random.default_rng(3)
x = rng.uniform(-10, 10, size=(n, 2)); y = rng.standard_normal(n)
ds = ctx.dataset(x, y)
props = [[(0, 3.0 * math.exp(0.05 * rng.standard_normal()), 5.5), (1, 0.5, 0.0)] for _ in range(B)]
for _ in range(3): ds.loglik_batch(props)
ts = []
for _ in range(5):
    ds.loglik_batch(props); ts.append(ctx.last_timing()[1])
print("light", os.environ.get("GPB_LIGHT_ROWS", "1"), f"n={n} B={B} chol ms", np.mean(ts))
ctx.set_trace(1)
ds.loglik_batch(props)
print("traced timing", ctx.last_timing())
tr = ctx.get_trace().astype(np.int64)
ctx.set_trace(0)
npad = (n + 63) // 64 * 64; nt = npad // 64; rt = nt + 1
dec = np.array([decode(B, nt, rt, t) for t in range(len(tr))])
g0 = tr[:, 0].min()
us = lambda a: a / (ghz * 1e3)
end = (tr[:, 7].max() - g0) / 1e3
print("kernel span us", end)
for t in range(len(tr)):
    b, i, j = dec[t]
    if b == 0 and (i == rt - 1 or i == j) and (j % 8 == 7 or j >= nt - 3):
        r = tr[t]
        print(f"({i},{j}) recv {(r[0]-g0)/1e3:8.1f} first {us(r[2]-r[1]):6.1f} mainloop {us(r[3]-r[2]):7.1f} dep {us(r[4]-r[3]):7.1f} solve {us(r[5]-r[4]):6.1f} rel {us(r[6]-r[5]):5.1f} released at {(r[7]-g0)/1e3:8.1f} notready {r[8]}")
