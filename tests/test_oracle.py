"""CPU tests of the oracle (oracle/gp_oracle.c) -- the checker the GPU parity tests rely on.

The reference ships no golden vectors for this path: the oracle is pinned by
hand-checkable closed forms, an independent scipy/LAPACK evaluation, a long-double evaluation
(noise floor) and the frozen fixtures in tests/golden/gp_golden.json.
"""
import json
import math
import os

import numpy as np
import pytest
import scipy.linalg as sl

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "gp_golden.json")
SE_W = [("se", 3.0, 5.5), ("white", 0.5)]


def test_kernel_functions_match_reference_semantics(oracle):
    # KernelFunction.scala:13-15: h * exp(-d^2 / sigma^2) -- no factor 2, h not squared
    assert oracle.cov_fn([("se", 3.0, 5.5)], 2.0) == 3.0 * math.exp(-(2.0 * 2.0) / (5.5 * 5.5))
    # KernelFunction.scala:33-35: sigma (not sigma^2) at exactly zero distance, else 0
    assert oracle.cov_fn([("white", 0.5)], 0.0) == 0.5
    assert oracle.cov_fn([("white", 0.5)], 1e-300) == 0.0
    # KernelFunction.scala:47-55: left fold from 0.0 in list order
    a, b, c = 3.0 * math.exp(-(1.0 * 1.0) / (5.5 * 5.5)), 1.0 * math.exp(-(1.0 * 1.0) / (0.8 * 0.8)), 0.0
    assert oracle.cov_fn([("se", 3.0, 5.5), ("se", 1.0, 0.8), ("white", 0.5)], 1.0) == ((0.0 + a) + b) + c


def test_closed_form_n1(oracle):
    x, y = np.array([[0.3]]), np.array([1.7])
    r = oracle.loglik(x, y, SE_W, nugget=1e-3)
    K = 3.0 + 0.5 + 1e-3
    L = math.sqrt(K)
    u = 1.7 / L
    S, C = math.log(L), 0.5 * math.log(2 * math.pi)
    assert r["info"] == 0 and r["n"] == 1
    assert r["Q"] == u * u and r["U1"] == u and r["S"] == S
    assert r["ll_gpml"] == pytest.approx(-0.5 * u * u - S - C, rel=1e-15)
    # as-parsed Scala expression: (-0.5*u) dot (u - S - C)
    assert r["ll_ref"] == pytest.approx((-0.5 * u) * ((u - S) - C), rel=1e-15)


def test_closed_form_n2(oracle):
    x, y = np.array([[0.0], [1.0]]), np.array([0.4, -0.9])
    r = oracle.loglik(x, y, SE_W, nugget=1e-3)
    k0, k1 = 3.0 + 0.5 + 1e-3, 3.0 * math.exp(-1.0 / 30.25)
    l11 = math.sqrt(k0); l21 = k1 / l11; l22 = math.sqrt(k0 - l21 * l21)
    u1 = 0.4 / l11; u2 = (-0.9 - l21 * u1) / l22
    assert r["Q"] == pytest.approx(u1 * u1 + u2 * u2, rel=1e-14)
    assert r["S"] == pytest.approx(math.log(l11) + math.log(l22), rel=1e-14)
    assert r["U1"] == pytest.approx(u1 + u2, rel=1e-13)


@pytest.mark.parametrize("n,dim", [(20, 1), (300, 1), (257, 2)])
def test_against_scipy_and_long_double(oracle, n, dim):
    rng = np.random.default_rng(n)
    x = rng.uniform(-10, 10, size=(n, dim))
    K = oracle.build_cov(x, SE_W, 1e-3)
    assert np.array_equal(K, K.T)
    y = np.linalg.cholesky(K) @ rng.standard_normal(n)
    r = oracle.loglik(x, y, SE_W)
    rl = oracle.loglik(x, y, SE_W, long_double=True)
    L = np.linalg.cholesky(K)
    u = sl.solve_triangular(L, y, lower=True)
    S, C = np.log(np.diag(L)).sum(), n * 0.5 * math.log(2 * math.pi)
    for key, twin in (("ll_gpml", -0.5 * u @ u - S - C), ("ll_ref", -0.5 * u @ u + 0.5 * (S + C) * u.sum())):
        assert abs(r[key] - twin) <= 1e-11 * abs(twin)
        assert abs(r[key] - rl[key]) <= 1e-11 * abs(rl[key])  # double-precision noise floor
    Lo, info = oracle.potrf_lower(K)
    assert info == 0 and np.allclose(Lo, L, rtol=0, atol=1e-12) and np.all(np.triu(Lo, 1) == 0)


def test_white_kernel_hits_duplicate_locations_off_diagonal(oracle):
    x = np.array([[1.0], [2.0], [1.0]])
    K = oracle.build_cov(x, SE_W, 0.0)
    assert K[0, 2] == 3.0 + 0.5 and K[2, 0] == 3.0 + 0.5
    assert K[0, 1] == 3.0 * math.exp(-1.0 / 30.25)


def test_plane_mean_is_subtracted_in_loglik_only(oracle):
    rng = np.random.default_rng(5)
    x = rng.uniform(-3, 3, size=(30, 2)); y = rng.standard_normal(30)
    beta = [0.3, -0.2, 0.7]
    r1 = oracle.loglik(x, y, SE_W, beta=beta)
    r2 = oracle.loglik(x, y - (beta[0] + x @ np.array(beta[1:])), SE_W)
    assert r1["ll_gpml"] == pytest.approx(r2["ll_gpml"], rel=1e-13)


def test_non_positive_definite_reports_lapack_info(oracle):
    K = np.array([[1.0, 2.0, 0.0], [2.0, 1.0, 0.0], [0.0, 0.0, 1.0]])
    _, info = oracle.potrf_lower(K)
    assert info == 2
    # negative white noise makes K indefinite -> info > 0 and NaN likelihood, never a crash
    r = oracle.loglik(np.arange(5.0), np.ones(5), [("se", 1.0, 5.0), ("white", -2.0)])
    assert r["info"] > 0 and math.isnan(r["ll_ref"])


def test_triangular_solves(oracle):
    rng = np.random.default_rng(9)
    n = 70
    A = rng.standard_normal((n, n)); K = A @ A.T + n * np.eye(n)
    L = np.linalg.cholesky(K); b = rng.standard_normal(n); B = rng.standard_normal((n, 4))
    assert np.allclose(oracle.trsv_lower(L, b), sl.solve_triangular(L, b, lower=True), rtol=1e-12, atol=1e-13)
    assert np.allclose(oracle.trsv_upper(L.T, b), sl.solve_triangular(L.T, b, lower=False), rtol=1e-12, atol=1e-13)
    assert np.allclose(oracle.trsm_upper(L.T, B), sl.solve_triangular(L.T, B, lower=False), rtol=1e-12, atol=1e-13)


def test_predict_matches_gpml_algorithm(oracle):
    rng = np.random.default_rng(11)
    n, m = 40, 9
    x = np.sort(rng.uniform(-10, 10, n)); y = rng.standard_normal(n); xs = rng.uniform(-10, 10, m)
    mean, var, info = oracle.predict(x, y, SE_W, xs, nugget=1e-6)
    assert info == 0
    K = oracle.build_cov(x, SE_W, 1e-6)
    d = np.abs(x[:, None] - xs[None, :])
    Ks = 3.0 * np.exp(-(d * d) / 30.25) + np.where(d == 0, 0.5, 0.0)
    assert np.allclose(mean, Ks.T @ np.linalg.solve(K, y), rtol=1e-9, atol=1e-11)
    # Predict.scala:78,92: kyy includes the white-noise term (variance of y*, not f*)
    assert np.allclose(var, (3.0 + 0.5) - np.einsum("ij,ij->j", Ks, np.linalg.solve(K, Ks)), rtol=1e-9, atol=1e-11)


def test_golden_fixtures(oracle):
    g = json.load(open(GOLDEN))
    assert len(g["cases"]) >= 8
    for c in g["cases"]:
        terms = [tuple(t) for t in c["terms"]]
        r = oracle.loglik(np.array(c["x"]), np.array(c["y"]), terms, beta=c["beta"], nugget=c["nugget_ll"])
        for k in ("Q", "U1", "S", "ll_ref", "ll_gpml"):
            assert r[k] == c["ll"][k], (c["n"], k)  # same binary, same machine class: bit-exact
        mean, var, info = oracle.predict(np.array(c["x"]), np.array(c["y"]), terms, np.array(c["xstar"]),
                                         nugget=c["nugget_pred"])
        assert info == 0 and np.array_equal(mean, c["pred_mean"]) and np.array_equal(var, c["pred_var"])


def test_mh_semantics(oracle):
    g = json.load(open(GOLDEN))["mh"]
    terms = [tuple(t) for t in g["terms"]]
    x, y, z, u = (np.array(g[k]) for k in ("x", "y", "z", "u"))
    for formula in (0, 1):
        want = g["by_formula"][str(formula)]
        # Breeze re-evaluates ll(t) every step (2 evals); caching it gives identical chains
        for evals in (2, 1):
            st, ll, acc, rc = oracle.mh_run(x, y, terms, g["ig"], g["delta"], z, u, formula=formula,
                                            evals_per_iter=evals)
            assert rc == 0
            assert np.array_equal(acc, want["accept"])
            assert np.array_equal(st, want["states"]) and np.array_equal(ll, want["ll"])
        assert 0 < sum(want["accept"]) < len(want["accept"])
    # rejected steps repeat the previous state; accepted steps follow the log-normal random walk
    st, acc = np.array(g["by_formula"]["0"]["states"]), np.array(g["by_formula"]["0"]["accept"])
    prev = np.array([3.0, 5.5, 0.5])
    for i in range(len(acc)):
        if acc[i]:
            assert np.allclose(st[i], prev * np.exp(g["delta"] * z[i]), rtol=1e-15)
        else:
            assert np.array_equal(st[i], prev)
        prev = st[i]


def test_inverse_gamma_prior(oracle):
    # InverseGamma(shape, scale).logPdf, SimulatedGp.scala:132-151 priors
    a, b, xv = 3.0, 4.0, 1.7
    want = a * math.log(b) - math.lgamma(a) - (a + 1) * math.log(xv) - b / xv
    assert oracle.lib().orc_invgamma_logpdf(a, b, xv) == pytest.approx(want, rel=1e-15)
    ig = [[3.0, 4.0], [3.0, 6.0], [10.0, 6.0]]
    p = oracle.prior(SE_W, ig)
    f = oracle.lib().orc_invgamma_logpdf
    assert p == (0.0 + (f(3.0, 4.0, 3.0) + f(3.0, 6.0, 5.5))) + f(10.0, 6.0, 0.5)


def test_sample_plane_and_draw_restatements(oracle):
    """SURVEY.md 8f-1/-3 restatements: the Gibbs draw solves the stated normal equations; the
    design matrix keeps the reference's layout; both roots of `draw` reproduce K."""
    rng = np.random.default_rng(12)
    x = rng.uniform(-5, 5, size=(40, 2)); y = rng.standard_normal(40)
    terms = [("se", 3.0, 5.5), ("white", 0.5)]
    z = rng.standard_normal(3)
    beta, prec, mean = oracle.sample_plane(x, y, terms, 1.0, 2.0, z)
    X = oracle.design_matrix(x)
    assert np.array_equal(X[:, 1], x.reshape(-1)[:40])          # interleaved, as the reference builds it
    K = oracle.build_cov(x, terms, 1e-6)
    assert np.allclose(prec, X.T @ np.linalg.solve(K, X) + np.eye(3) / 2.0, rtol=1e-10)
    L = np.linalg.cholesky(prec)
    assert np.allclose(L @ (beta - mean), z, rtol=1e-9, atol=1e-12)
    x1 = np.sort(rng.uniform(-10, 10, 50))
    ze = rng.standard_normal(50)
    eig_draw, chol_draw = oracle.draw(x1, terms, ze, beta=(0.5, 0.1))
    K1 = oracle.build_cov(x1, terms, 1e-3)
    mu = 0.5 + 0.1 * x1
    Lc = np.linalg.cholesky(K1)
    assert np.allclose(chol_draw, mu + Lc @ ze, rtol=1e-10, atol=1e-12)
    w, V = np.linalg.eigh(K1)
    R = V * np.sqrt(w)
    assert np.allclose(R @ R.T, Lc @ Lc.T, rtol=1e-9, atol=1e-10)   # two roots of the same covariance
    assert np.allclose(eig_draw, mu + R @ ze, rtol=1e-8, atol=1e-10)


def test_mll_gradient_restatement(oracle):
    """SURVEY.md 8f-2.  The analytic mode is the gradient of ll_gpml in the log parameters (checked by
    central differences); the as-executed mode follows the Scala formulas, which are NOT that gradient."""
    rng = np.random.default_rng(77)
    n = 60
    x = rng.uniform(-5, 5, size=(n, 2))
    terms = [("se", 3.0, 2.5), ("se", 0.7, 0.6), ("white", 0.5)]
    y = np.linalg.cholesky(oracle.build_cov(x, terms, 1e-3)) @ rng.standard_normal(n) + 0.3 + 0.1 * x[:, 0]
    beta = (0.3, 0.1, 0.0)
    theta = [("se", math.log(3.0), math.log(2.5)), ("se", math.log(0.7), math.log(0.6)), ("white", math.log(0.5))]
    g = oracle.mll_gradient(x, y, theta, beta=beta, mode="analytic")
    flat = [math.log(3.0), math.log(2.5), math.log(0.7), math.log(0.6), math.log(0.5)]

    def ll(v):
        t = [("se", math.exp(v[0]), math.exp(v[1])), ("se", math.exp(v[2]), math.exp(v[3])), ("white", math.exp(v[4]))]
        return oracle.loglik(x, y, t, beta=beta)["ll_gpml"]

    for p in range(5):
        e = np.zeros(5); e[p] = 1e-5
        fd = (ll(np.array(flat) + e) - ll(np.array(flat) - e)) / 2e-5
        assert abs(fd - g[p]) < 1e-5 * max(1.0, abs(g[p])), (p, fd, g[p])
    # as executed: every tangent carries the "+ derivative" constant, the white one is the constant 2
    ge = oracle.mll_gradient(x, y, [("se", 1.1, 0.9), ("white", 0.4)], mode="as_executed")
    K = oracle.build_cov(x, [("se", 1.1, 0.9), ("white", 0.4)], 1e-3)
    a = np.linalg.solve(K, y)
    s1 = a.sum() ** 2 - np.linalg.solve(K, np.ones((n, n))).trace()
    assert ge.shape == (3,) and abs(ge[2] - s1) < 1e-9 * abs(s1)
