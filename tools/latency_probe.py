"""Per-call latency of the reference-facing path for small problems: wall time per gpb_loglik_batch call vs
the device time of its three kernels (build / factor+solve / epilogue)."""
import math, os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "gaussian-processes_b200"))
from gpb200 import _lib as G
ctx = G.Context(0)
rng = np.random.default_rng(3)
for n, B in [(64, 1), (300, 1), (300, 2), (1000, 1), (1000, 2), (2048, 1), (4096, 1)]:
    x = np.sort(rng.uniform(-10, 10, n)); y = rng.standard_normal(n)
    ds = ctx.dataset(x, y)
    props = [[(0, 3.0, 5.5), (1, 0.5, 0.0)]] * B
    for _ in range(20): ds.loglik_batch(props)
    N = 200
    dev = np.zeros(3)
    t0 = time.perf_counter()
    for _ in range(N):
        ds.loglik_batch(props); dev += ctx.last_timing()
    wall = (time.perf_counter() - t0) / N * 1e6
    dev = dev / N * 1e3
    print(f"n={n:5d} B={B}: wall {wall:7.1f} us/call | device build {dev[0]:6.1f} chol {dev[1]:7.1f} epi {dev[2]:5.1f} us | host+sync overhead {wall - dev.sum():6.1f} us | nt={(n+63)//64} -> {dev[1]/((n+63)//64):.1f} us per block column")
    ds.close()
