import sys, os, math
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "gaussian-processes_b200"))
from gpb200 import _lib as G
ctx = G.Context(0)
rng = np.random.default_rng(5)
n = int(sys.argv[1]) if len(sys.argv) > 1 else 1024
x = rng.uniform(-10, 10, size=(n, 2))
d2 = ((x[:, None, :] - x[None, :, :]) ** 2).sum(-1)
K = 3.0 * np.exp(-d2 / 30.25) + 0.501 * np.eye(n)
Lref = np.linalg.cholesky(K)
for rep in range(3):
    L, info = ctx.potrf(K)
    err = np.abs(L - Lref)
    err[np.isnan(L)] = np.inf
    nt = n // 64
    tmap = np.zeros((nt, nt))
    for i in range(nt):
        for j in range(i + 1):
            tmap[i, j] = err[i*64:(i+1)*64, j*64:(j+1)*64].max()
    bad = np.argwhere(tmap > 1e-9)
    print(f"rep {rep}: info {info}, bad tiles {len(bad)} of {nt*(nt+1)//2}")
    if len(bad):
        # earliest in dependency order: smallest column, then row
        order = sorted(map(tuple, bad), key=lambda t: (t[1], t[0]))
        print("  first bad tiles (i,j):", order[:12])
        i, j = order[0]
        e = err[i*64:(i+1)*64, j*64:(j+1)*64]
        print("  per 8x8 sub-block max err of the first bad tile (rows=r-blocks, cols=c-blocks):")
        for rb in range(8):
            print("   ", " ".join("%8.1e" % e[rb*8:(rb+1)*8, cb*8:(cb+1)*8].max() for cb in range(8)))
        got = L[i*64:(i+1)*64, j*64:(j+1)*64]; ref = Lref[i*64:(i+1)*64, j*64:(j+1)*64]
        print("  sample got/ref [0,0]:", got[0,0], ref[0,0], " nan count", int(np.isnan(got).sum()))
