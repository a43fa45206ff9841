"""Race hunt: the factorisation is schedule-independent by construction, so repeated runs of the
same batch must agree bit for bit (and with the per-phase debug scheduler)."""
import sys, os, math
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "gaussian-processes_b200"))
from gpb200 import _lib as G
ctx = G.Context(0)
rng = np.random.default_rng(11)
bad_total = 0
for n, B, reps in [(4096, 8, 12), (4096, 1, 12), (4096, 2, 12), (4096, 3, 8), (2048, 8, 12), (1024, 32, 20), (1024, 64, 20), (8192, 2, 4), (300, 16, 20)]:
    x = rng.uniform(-10, 10, size=(n, 2)); y = rng.standard_normal(n)
    ds = ctx.dataset(x, y)
    props = [[(0, 3.0 * math.exp(0.05 * rng.standard_normal()), 5.5), (1, 0.5, 0.0)] for _ in range(B)]
    ctx.set_sched(1)
    ref = [(r.Q, r.U1, r.S, r.info) for r in ds.loglik_batch(props)]
    ctx.set_sched(0)
    bad = 0
    for rep in range(reps):
        got = [(r.Q, r.U1, r.S, r.info) for r in ds.loglik_batch(props)]
        if got != ref:
            bad += 1
            if bad <= 2:
                diffs = [(i, g, r) for i, (g, r) in enumerate(zip(got, ref)) if g != r]
                print("   mismatch rep", rep, diffs[:2])
    print(f"n={n} B={B}: {bad}/{reps} runs differ from the per-phase scheduler; infos {sorted(set(r[3] for r in ref))}")
    bad_total += bad
    ds.close()
print("pause", os.environ.get("GPB_CHAIN_PAUSE", "auto"), "TOTAL_BAD", bad_total)
