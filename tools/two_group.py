"""Experiment: the chains of one GPU as G groups, one gpb_ctx + host thread per group, so that their
persistent factorisation kernels run side by side (GPB_CTAS_PER_SM_RT=2 GPB_GRID_CAP=1: one CTA per SM
each, two-CTA register budget).  Prints chain-iterations/s against one context running all chains."""
import os, sys, threading, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "gaussian-processes_b200"))
import bench as B
from gpb200 import _lib as G
n, chains, iters, groups = int(sys.argv[1]), int(sys.argv[2]), int(sys.argv[3]), int(sys.argv[4])
delay_us = float(sys.argv[5]) if len(sys.argv) > 5 else 0.0
ctxs = [G.Context(0) for _ in range(groups)]
x, y = B.synth_dataset(ctxs[0], n, 2, -10, 10, 3, B.TRUE_TERMS)
dss = [c.dataset(x, y) for c in ctxs]
inits = [B.gterms(B.TRUE_TERMS)] * (chains // groups)
def run(g, nit, seed):
    return dss[g].mh_run_ex(inits, B.IG3, 0.05, nit, rng=G.RNG_PHILOX, seed=seed, chain0=g * len(inits), on_notpd=1)
for g in range(groups): run(g, 10, 1)
bar = threading.Barrier(groups)
res = [None] * groups
def worker(g):
    bar.wait()
    if g: 
        t = time.perf_counter()
        while (time.perf_counter() - t) * 1e6 < delay_us * g: pass
    res[g] = run(g, iters, 2)
th = [threading.Thread(target=worker, args=(g,)) for g in range(groups)]
t0 = time.perf_counter()
[t.start() for t in th]; [t.join() for t in th]
w = time.perf_counter() - t0
print(f"n={n} chains={chains} groups={groups} delay={delay_us}us cap={os.environ.get('GPB_GRID_CAP','-')} ctas={os.environ.get('GPB_CTAS_PER_SM_RT','-')}: "
      f"{iters / w:.1f} it/s of all {chains} chains; device ms per iteration per group {[round(r['device_ms'] / (iters + 1), 4) for r in res]}")
