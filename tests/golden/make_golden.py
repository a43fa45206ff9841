"""Generates tests/golden/gp_golden.json: seeded inputs + the oracle's outputs for the hot path.

The reference ships no golden vectors for this path and cannot run here (Scala/JVM),
so these vectors freeze the CPU oracle (oracle/gp_oracle.c) after cross-checking every case
against an independent evaluation (numpy/scipy LAPACK) to 1e-11: "parity unpinned" still applies
with respect to the real Breeze build.  Run:  python tests/golden/make_golden.py
"""
import json
import os
import sys

import numpy as np
import scipy.linalg as sl

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import oracle as o  # noqa: E402

KERNELS = {
    "se_w": [("se", 3.0, 5.5), ("white", 0.5)],              # SimulatedGp.scala:30-33
    "se_se_w": [("se", 3.0, 5.5), ("se", 1.0, 0.8), ("white", 0.5)],
    "w_se": [("white", 0.25), ("se", 2.0, 1.5)],             # order matters for the left fold
}


def scipy_twin(x, y, terms, beta, nugget):
    x = np.asarray(x, float).reshape(len(y), -1)
    d = np.sqrt(((x[:, None, :] - x[None, :, :]) ** 2).sum(-1))
    K = np.zeros_like(d)
    for t in terms:
        K = K + (t[1] * np.exp(-(d * d) / (t[2] * t[2])) if t[0] == "se" else np.where(d == 0, t[1], 0.0))
    K = K + nugget * np.eye(len(y))
    mu = np.zeros(len(y)) if len(beta) == 0 else beta[0] + x @ np.asarray(beta[1:])
    L = np.linalg.cholesky(K)
    u = sl.solve_triangular(L, y - mu, lower=True)
    return K, L, u


def main():
    cases = []
    rng = np.random.default_rng(20240611)
    for n, dim, kname, plane in [(1, 1, "se_w", False), (2, 1, "se_w", False), (3, 2, "se_se_w", False),
                                 (20, 1, "se_w", False), (20, 2, "w_se", True), (64, 2, "se_w", False),
                                 (65, 1, "se_se_w", True), (130, 2, "se_se_w", False)]:
        terms = KERNELS[kname]
        x = np.round(rng.uniform(-10, 10, size=(n, dim)), 6)
        if n >= 20:
            x[7] = x[3]  # exact duplicate location: the white kernel adds sigma OFF the diagonal too
        beta = list(np.round(rng.normal(size=dim + 1), 3)) if plane else []
        K = o.build_cov(x, terms, 1e-3)
        y = np.round(np.linalg.cholesky(K) @ rng.standard_normal(n), 6)
        r = o.loglik(x, y, terms, beta=beta, nugget=1e-3)
        Kt, Lt, ut = scipy_twin(x, y, terms, beta, 1e-3)
        S = np.log(np.diag(Lt)).sum()
        C = n * 0.5 * np.log(2 * np.pi)
        assert np.abs(Kt - K).max() < 1e-13
        for got, want in ((r["Q"], ut @ ut), (r["S"], S), (r["ll_gpml"], -0.5 * ut @ ut - S - C),
                          (r["ll_ref"], -0.5 * ut @ ut + 0.5 * (S + C) * ut.sum())):
            assert abs(got - want) <= 1e-11 * max(1.0, abs(want)), (n, got, want)
        xs = np.round(rng.uniform(-10, 10, size=(5, dim)), 6)
        xs[0] = x[0]  # a test point on top of a training point
        mean, var, info = o.predict(x, y, terms, xs, nugget=1e-6)
        assert info == 0
        cases.append({"n": n, "dim": dim, "kernel": kname, "terms": terms, "beta": beta,
                      "x": x.tolist(), "y": y.tolist(), "nugget_ll": 1e-3, "ll": r,
                      "xstar": xs.tolist(), "nugget_pred": 1e-6, "pred_mean": mean.tolist(),
                      "pred_var": var.tolist()})
    # one short MH chain (SimulatedGp.scala priors / proposal), explicit streams
    n = 20
    x = np.round(np.sort(rng.uniform(-10, 10, size=n)), 6)
    terms = KERNELS["se_w"]
    y = np.round(np.linalg.cholesky(o.build_cov(x, terms, 1e-3)) @ rng.standard_normal(n), 6)
    iters = 40
    z = np.round(rng.standard_normal((iters, 3)), 8)
    u = np.round(rng.uniform(size=iters), 8)
    ig = [[3.0, 4.0], [3.0, 6.0], [10.0, 6.0]]  # priorh, priorSigma, priorSigmaY
    mh = {}
    for formula in (0, 1):
        st, ll, acc, rc = o.mh_run(x, y, terms, ig, 0.05, z, u, formula=formula, evals_per_iter=2)
        assert rc == 0
        mh[str(formula)] = {"states": st.tolist(), "ll": ll.tolist(), "accept": acc.tolist()}
    out = {"cases": cases,
           "mh": {"x": x.tolist(), "y": y.tolist(), "terms": terms, "ig": ig, "delta": 0.05,
                  "z": z.tolist(), "u": u.tolist(), "by_formula": mh}}
    path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "gp_golden.json")
    with open(path, "w") as f:
        json.dump(out, f)
    print("wrote", path, os.path.getsize(path), "bytes")


if __name__ == "__main__":
    main()
