"""Scratch: MH iterations/s when the chains of one GPU run as G concurrent chain groups, each group
with its own gpb_ctx (stream + workspace) on its own host thread.  The persistent factorisation
kernel of one group releases SMs as its tickets run out (chain-bound tail); the other groups'
kernels fill them."""
import sys, os, math, time, threading
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "gaussian-processes_b200"))
from gpb200 import _lib as G

IG3 = [[3.0, 4.0], [3.0, 6.0], [10.0, 6.0]]
cases = [(1024, 256, 60), (1024, 128, 60), (1024, 64, 100), (1024, 32, 100)]
if len(sys.argv) > 1:
    cases = [tuple(int(v) for v in a.split(",")) for a in sys.argv[1:]]
rng = np.random.default_rng(3)
for n, chains, iters in cases:
    x = rng.uniform(-10, 10, size=(n, 2)); y = rng.standard_normal(n) * 1.8
    for groups in (1, 2, 4, 8):
        if chains % groups: continue
        per = chains // groups
        ctxs = [G.Context(0) for _ in range(groups)]
        dss = [c.dataset(x, y) for c in ctxs]
        inits = [[(0, 3.0 * math.exp(0.05 * rng.standard_normal()), 5.5), (1, 0.5, 0.0)] for _ in range(chains)]
        res = [None] * groups

        def work(gi, nit, seed):
            z, u = G.philox_streams(seed, gi * per, per, nit, 3)
            res[gi] = dss[gi].mh_run(inits[gi * per:(gi + 1) * per], IG3, 0.05, z, u, on_notpd=1)

        def run(nit, seed):
            th = [threading.Thread(target=work, args=(gi, nit, seed)) for gi in range(groups)]
            t0 = time.perf_counter()
            for t in th: t.start()
            for t in th: t.join()
            return time.perf_counter() - t0

        run(5, 100)
        best = min(run(iters, 101) for _ in range(3))
        acc = np.mean([r[2].mean() for r in res])
        print(f"n={n} chains={chains} groups={groups}: {iters / best:9.1f} it/s  ({best / iters * 1e3:.3f} ms/iter, accept {acc:.3f})", flush=True)
        for d in dss: d.close()
        for c in ctxs: c.close()
