"""Factor+solve time per shape with the chain pause forced off / on (GPB_CHAIN_PAUSE=0|1 per process)."""
import math, os, subprocess, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SHAPES = [(1024, 8), (1024, 16), (1024, 32), (1024, 64), (1024, 128), (1024, 256), (300, 32), (300, 256), (2048, 4), (2048, 8),
          (2048, 16), (2048, 64), (4096, 2), (4096, 3), (4096, 4), (4096, 8), (8192, 2), (8192, 8)]
if len(sys.argv) > 1 and sys.argv[1] == "child":
    sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "gaussian-processes_b200"))
    from gpb200 import _lib as G
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
        out.append((n, B, float(np.median(ts)), sum(v.Q for v in r)))
        ds.close()
    print(repr(out))
else:
    res = {}
    for mode in ("0", "1"):
        env = dict(os.environ, GPB_CHAIN_PAUSE=mode)
        res[mode] = eval(subprocess.check_output([sys.executable, __file__, "child"], env=env).decode().strip().splitlines()[-1])
    for a, b in zip(res["0"], res["1"]):
        same = "bit-identical" if a[3] == b[3] else "RESULTS DIFFER"
        print(f"n={a[0]:5d} B={a[1]:4d}: pause off {a[2]:8.3f} ms   on {b[2]:8.3f} ms   ({100 * (b[2] / a[2] - 1):+.1f} %)  {same}")
