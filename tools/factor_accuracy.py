"""Accuracy of the device factorisation against a long-double Cholesky of the same matrix (numpy longdouble, column by
column): forward error of L and of log det, backward residual -- for well- and ill-conditioned covariance matrices."""
import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "gaussian-processes_b200"))
from gpb200 import _lib as G

def chol_ld(K):
    A = K.astype(np.longdouble).copy(); n = A.shape[0]
    L = np.zeros_like(A)
    for j in range(n):
        v = A[j:, j] - L[j:, :j] @ L[j, :j]
        L[j:, j] = v / np.sqrt(v[0])
    return L

ctx = G.Context(0)
print(os.path.basename(os.environ.get("GPB200_LIB", "default")))
for n, ls, nug in [(192, 5.5, 1e-3), (192, 5.5, 1e-6), (320, 9.0, 1e-6), (448, 3.0, 1e-3), (448, 20.0, 1e-7)]:
    rng = np.random.default_rng(n)
    x = rng.uniform(-10, 10, size=(n, 2))
    K = ctx.build_cov(x, [(0, 3.0, ls), (1, 0.5 if nug > 1e-4 else 0.0, 0.0)], nug)
    L, info = ctx.potrf(K)
    Lr = chol_ld(K)
    cond = np.linalg.cond(K)
    fe = float(np.max(np.abs(L - Lr)) / np.max(np.abs(Lr)))
    ld = float(abs(np.sum(np.log(np.diag(L))) - np.sum(np.log(np.diag(Lr).astype(np.float64)))))
    res = float(np.linalg.norm((L.astype(np.longdouble) @ L.T.astype(np.longdouble) - K).astype(np.float64)) / np.linalg.norm(K))
    Ls = np.linalg.cholesky(K)
    fe_lapack = float(np.max(np.abs(Ls - Lr)) / np.max(np.abs(Lr)))
    print(f"n={n} ls={ls} nugget={nug:g} cond={cond:.1e} info={info}: max|L-L_ld|/max|L| {fe:.2e} (LAPACK {fe_lapack:.2e})  |dlogdet| {ld:.2e}  ||LL^T-K||/||K|| {res:.2e}")
