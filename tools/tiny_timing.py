"""One-launch small-n MH kernel: iterations/s by system size and chain count."""
import os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "gaussian-processes_b200"))
import bench as B
from gpb200 import _lib as G
ctx = G.Context(0)
for n in (20, 32, 64):
    x, y = B.synth_dataset(ctx, n, 1, -10, 10, 1, B.TRUE_TERMS)
    ds = ctx.dataset(x, y)
    for ch in (2, 148, 1184):
        inits = [B.gterms(B.TRUE_TERMS)] * ch
        ds.mh_run_ex(inits, B.IG3, 0.05, 200, rng=G.RNG_PHILOX, seed=1, on_notpd=1)
        t0 = time.perf_counter(); r = ds.mh_run_ex(inits, B.IG3, 0.05, 20000, rng=G.RNG_PHILOX, seed=2, on_notpd=1); w = time.perf_counter() - t0
        print(f"n={n} chains={ch}: {20000 / w:.0f} it/s e2e, {1e3 * r['device_ms'] / 20000:.2f} us per iteration (device), path {r['path']}, accept {r['accept'].mean():.3f}")
    ds.close()
