import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "gaussian-processes_b200"))
import bench as B
from gpb200 import _lib as G
ctx = G.Context(0)
n = int(sys.argv[1]) if len(sys.argv) > 1 else 20
x, y = B.synth_dataset(ctx, n, 1, -10, 10, 1, B.TRUE_TERMS)
ds = ctx.dataset(x, y)
r = ds.mh_run_ex([B.gterms(B.TRUE_TERMS)] * 2, B.IG3, 0.05, 1024, rng=G.RNG_PHILOX, seed=2, on_notpd=1)
print(r["device_ms"])
