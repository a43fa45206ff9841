"""Scratch bring-up check (not a test): CUDA path vs oracle on a few sizes, both schedulers."""
import sys, time, os
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "gaussian-processes_b200"))
from oracle import oracle as o
from gpb200 import _lib as G

ctx = G.Context(0)
terms_o = [("se", 3.0, 5.5), ("white", 0.5)]
terms_g = [(0, 3.0, 5.5), (1, 0.5, 0.0)]
rng = np.random.default_rng(2)
modes = [int(a) for a in sys.argv[1].split(",")] if len(sys.argv) > 1 else [1, 0]
sizes = [int(a) for a in sys.argv[2].split(",")] if len(sys.argv) > 2 else [1, 2, 20, 64, 65, 300, 1024]
for mode in modes:
    ctx.set_sched(mode)
    for n in sizes:
        dim = 1 if n <= 300 else 2
        x = rng.uniform(-10, 10, size=(n, dim))
        K = o.build_cov(x, terms_o, 1e-3)
        y = np.linalg.cholesky(K) @ rng.standard_normal(n)
        ds = ctx.dataset(x, y)
        t0 = time.time(); r = ds.loglik_batch([terms_g])[0]; dt = time.time() - t0
        ro = o.loglik(x, y, terms_o)
        ref, gp = G.ll_value(r, 0), G.ll_value(r, 1)
        print(f"mode={mode} n={n} info={r.info} ll_ref={ref:.12g} (oracle {ro['ll_ref']:.12g}) "
              f"rel={abs(ref-ro['ll_ref'])/abs(ro['ll_ref']):.2e} gpml rel={abs(gp-ro['ll_gpml'])/abs(ro['ll_gpml']):.2e} "
              f"t={dt*1e3:.2f}ms timing={ctx.last_timing()}", flush=True)
        if n <= 300:
            Kg = ctx.build_cov(x, terms_g, 1e-3)
            print("   build_cov max abs diff", np.abs(Kg - K).max(), " dist diff",
                  np.abs(ctx.distance_matrix(x) - o.build_cov(x, [("se", 1.0, 1.0)], 0.0) * 0).max() * 0)
            Lg, info = ctx.potrf(K)
            Lo, _ = o.potrf_lower(K)
            print("   potrf info", info, "max abs diff", np.abs(Lg - Lo).max())
            xs = rng.uniform(-10, 10, size=(70, dim))
            mg, vg, ig = ds.predict(terms_g, xs)
            mo, vo, io = o.predict(x, y, terms_o, xs)
            print("   predict info", ig, "mean diff", np.abs(mg - mo).max(), "var diff", np.abs(vg - vo).max())
        ds.close()
# batch
n = 1024
x = rng.uniform(-10, 10, size=(n, 2)); K = o.build_cov(x, terms_o, 1e-3); y = np.linalg.cholesky(K) @ rng.standard_normal(n)
ds = ctx.dataset(x, y)
B = 32
tl = [[(0, 3.0 * np.exp(0.05 * rng.standard_normal()), 5.5 * np.exp(0.05 * rng.standard_normal())), (1, 0.5, 0.0)] for _ in range(B)]
for mode in modes:
    ctx.set_sched(mode)
    for rep in range(3):
        t0 = time.time(); rs = ds.loglik_batch(tl); dt = time.time() - t0
    errs = []
    for b in (0, 7, B - 1):
        ro = o.loglik(x, y, [("se", tl[b][0][1], tl[b][0][2]), ("white", 0.5)])
        errs.append(abs(G.ll_value(rs[b], 1) - ro["ll_gpml"]) / abs(ro["ll_gpml"]))
    print(f"batch={B} n={n} mode={mode} t={dt*1e3:.2f}ms timing={ctx.last_timing()} rel errs {errs}")
print("launches", ctx.launch_count())
