"""A short device-resident MH run at one shape (n, chains, iters) -- for launch lists / profiles of the loop."""
import os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "gaussian-processes_b200"))
import bench as B
from gpb200 import _lib as G
n, chains, iters = int(sys.argv[1]), int(sys.argv[2]), int(sys.argv[3])
inits_kind = sys.argv[4] if len(sys.argv) > 4 else "prior"
ctx = G.Context(0)
x, y = B.synth_dataset(ctx, n, 2, -10, 10, 3, B.TRUE_TERMS)
ds = ctx.dataset(x, y)
inits = B.prior_draws(B.TRUE_TERMS, B.IG3, chains, 20) if inits_kind == "prior" else [B.gterms(B.TRUE_TERMS)] * chains
ds.mh_run_ex(inits, B.IG3, 0.05, 5, rng=G.RNG_PHILOX, seed=1, on_notpd=1)
t0 = time.perf_counter()
r = ds.mh_run_ex(inits, B.IG3, 0.05, iters, rng=G.RNG_PHILOX, seed=2, on_notpd=1)
w = time.perf_counter() - t0
print(f"n={n} chains={chains} iters={iters} inits={inits_kind}: {iters / w:.1f} it/s e2e, device {r['device_ms'] / (iters + 1):.4f} ms per evaluation, accept {r['accept'].mean():.3f}")
