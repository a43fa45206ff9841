"""Race / schedule hunt over ragged shapes: likelihood batches and predictive systems (light and full
last tile rows), production scheduler vs the per-phase debug scheduler, bit for bit, repeated.
GPB_CTAS_PER_SM_RT=1|2 forces the launch mode (run once per mode)."""
import sys, os, math
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "gaussian-processes_b200"))
from gpb200 import _lib as G
ctx = G.Context(0)
rng = np.random.default_rng(int(os.environ.get("SEED", "5")))
bad_total = runs = 0
shapes = [(1, 1, 0), (2, 3, 0), (63, 7, 0), (64, 9, 0), (65, 33, 0), (100, 300, 0), (300, 64, 0), (640, 40, 0), (1000, 17, 0), (1023, 5, 0), (1025, 3, 0),
          (2049, 2, 0), (3000, 1, 0), (3000, 3, 0), (150, 1, 69), (150, 1, 64), (150, 1, 57), (700, 1, 1000), (1024, 1, 7), (2048, 1, 2055)]
for n, B, m in shapes:
    dim = 1 + (n % 2)
    x = rng.uniform(-10, 10, size=(n, dim)); y = rng.standard_normal(n)
    ds = ctx.dataset(x, y)
    props = [[(0, 3.0 * math.exp(0.05 * rng.standard_normal()), 5.5), (1, 0.5, 0.0)] for _ in range(B)]
    xs = rng.uniform(-10, 10, size=(m, dim)) if m else None
    def run():
        out = [(r.Q, r.U1, r.S, r.info) for r in ds.loglik_batch(props)]
        if m:
            mean, var, info = ds.predict(props[0], xs, nugget=1e-3)
            out.append((mean.tobytes(), var.tobytes(), info))
        return out
    ctx.set_sched(1); ref = run(); ctx.set_sched(0)
    reps = 12 if n <= 1100 else 5
    bad = sum(run() != ref for _ in range(reps))
    runs += reps; bad_total += bad
    print(f"n={n} dim={dim} B={B} m={m}: {bad}/{reps} differ; info {sorted(set(r[3] for r in ref[:B]))}", flush=True)
    ds.close()
print("mode", os.environ.get("GPB_CTAS_PER_SM_RT", "auto"), "RUNS", runs, "TOTAL_BAD", bad_total)
