"""GPU parity tests: the CUDA path, called through the C ABI (libgpb200.so), against the CPU oracle
on the same seeded inputs, against the frozen golden fixtures, and -- at BASELINE.json's full
sizes -- through size-independent properties (L L^T = K, linear-system residuals, LAPACK twin).

Tolerances (north_star): relative 1e-10 on the log-likelihood in FP64 (we observe ~1e-14);
identical accept/reject sequences for identical RNG streams.  cond(K) is printed with the
large-n parity numbers.
"""
import json
import math
import os

import numpy as np
import pytest
import scipy.linalg as sl

pytestmark = pytest.mark.gpu

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "gp_golden.json")
SE_W = [("se", 3.0, 5.5), ("white", 0.5)]
SE_SE_W = [("se", 3.0, 5.5), ("se", 1.0, 0.8), ("white", 0.5)]
SHORT_W = [("se", 3.0, 0.2), ("white", 0.5)]
LL_RTOL = 1e-10


def gterms(terms):
    return [(0, t[1], t[2]) if t[0] == "se" else (1, t[1], 0.0) for t in terms]


def synth(oracle, n, dim, terms, seed, lo=-10.0, hi=10.0):
    rng = np.random.default_rng(seed)
    x = rng.uniform(lo, hi, size=(n, dim))
    if n >= 8:
        x[5] = x[2]  # exact duplicate: white noise lands off the diagonal
    K = oracle.build_cov(x, terms, 1e-3)
    y = np.linalg.cholesky(K) @ rng.standard_normal(n)
    return x, y, K


def rel(a, b):
    return abs(a - b) / max(abs(b), 1e-300)


@pytest.mark.parametrize("n", [1, 2, 3, 20, 63, 64, 65, 128, 300])
@pytest.mark.parametrize("dim", [1, 2])
def test_build_cov_and_distance_entrywise(ctx, oracle, n, dim):
    x, _, _ = synth(oracle, n, dim, SE_SE_W, 100 + n)
    K = ctx.build_cov(x, gterms(SE_SE_W), 1e-3)
    Ko = oracle.build_cov(x, SE_SE_W, 1e-3)
    # only exp() differs (CUDA libm vs glibc, both <= 1 ulp): allow 4 ulp of the entry
    assert np.all(np.abs(K - Ko) <= 4 * np.spacing(np.abs(Ko)))
    assert np.array_equal(K, K.T)
    D = ctx.distance_matrix(x)
    Do = np.sqrt(((x[:, None, :] - x[None, :, :]) ** 2).sum(-1))
    assert np.all(np.abs(D - Do) <= 2 * np.spacing(Do))
    assert np.all(np.diag(D) == 0)
    if n >= 8:
        assert K[5, 2] == 3.0 + 1.0 + 0.5  # SE(0) + SE(0) + white, no nugget off the diagonal


@pytest.mark.parametrize("n", [3, 64, 65, 130, 300, 1024])
@pytest.mark.parametrize("dim", [1, 2])
@pytest.mark.parametrize("terms", [SE_W, SE_SE_W, SHORT_W, [("white", 0.5), ("se", 3.0, 0.2), ("se", 1.0, 5.0)]])
def test_hot_path_build_entrywise(ctx, oracle, n, dim, terms):
    """The matrix the factorisation actually consumes (k_build's hot-path form h exp(d^2 (-1/s^2)) with
    the hand-written exp, gpb_build_cov_hot) against the oracle's exact-order matrix, entry by entry:
    |dK| <= 2 ulp of the SE amplitudes (argument rounding 0.55 ulp(h) at worst, exp <= 0.84 ulp here
    and < 1 ulp in glibc) -- including the underflow branch (length scale 0.2: arguments below -700),
    duplicate locations, ragged sizes and dim = 2."""
    x, _, _ = synth(oracle, n, dim, terms, 300 + n)
    Kh = ctx.build_cov_hot(x, gterms(terms), 1e-3)
    Ko = oracle.build_cov(x, terms, 1e-3)
    lower = np.tril(np.ones((n, n), dtype=bool))
    tol = 2.0 * sum(np.spacing(t[1]) for t in terms if t[0] == "se")
    err = np.abs(Kh - Ko)[lower]
    print(f"hot-path build n={n} dim={dim}: max |dK| = {err.max() / (tol / 2.0):.3f} ulp(h)")
    assert err.max() <= tol
    assert np.all(Kh[~lower] == 0.0)
    assert np.all(np.diag(Kh) == np.diag(Ko))          # exp(0) = 1 exactly on both sides
    if n >= 8:
        assert Kh[5, 2] == Ko[5, 2]                    # duplicate location: white lands off the diagonal


@pytest.mark.parametrize("sched", [0, 1])
@pytest.mark.parametrize("n,dim,terms", [
    (1, 1, SE_W), (2, 1, SE_W), (3, 2, SE_SE_W), (20, 1, SE_W), (63, 2, SE_W), (64, 1, SE_SE_W),
    (65, 2, SE_W), (300, 1, SE_W), (300, 2, SE_SE_W), (1024, 2, SE_W),
    # length scale 0.2 on [-10, 10]: -d^2/s^2 reaches -20000, most SE entries underflow -- the rare
    # branch of the hot-path exp (arguments below -700 go to the library routine)
    (128, 2, SHORT_W), (192, 1, SHORT_W), (256, 2, [("white", 0.5), ("se", 3.0, 0.2), ("se", 1.0, 5.0)])])
def test_loglik_parity(ctx, oracle, sched, n, dim, terms):
    ctx.set_sched(sched)
    try:
        x, y, _ = synth(oracle, n, dim, terms, 7 * n + dim)
        ds = ctx.dataset(x, y)
        from gpb200 import _lib as G
        r = ds.loglik_batch([gterms(terms)], nugget=1e-3)[0]
        ro = oracle.loglik(x, y, terms, nugget=1e-3)
        assert r.info == 0 and r.n == n
        assert rel(r.Q, ro["Q"]) < LL_RTOL and rel(r.S, ro["S"]) < LL_RTOL
        assert abs(r.U1 - ro["U1"]) < 1e-10 * max(1.0, math.sqrt(n))
        assert rel(G.ll_value(r, G.FORMULA_AS_PARSED), ro["ll_ref"]) < LL_RTOL  # what the reference executes
        assert rel(G.ll_value(r, G.FORMULA_GPML), ro["ll_gpml"]) < LL_RTOL      # what its docs intend
        ds.close()
    finally:
        ctx.set_sched(0)


def test_plane_mean_and_batch_of_proposals(ctx, oracle):
    from gpb200 import _lib as G
    n = 200
    x, y, _ = synth(oracle, n, 2, SE_W, 31)
    ds = ctx.dataset(x, y)
    rng = np.random.default_rng(1)
    props, betas = [], []
    for b in range(9):
        props.append([("se", 3.0 * math.exp(0.2 * rng.standard_normal()), 5.5 * math.exp(0.2 * rng.standard_normal())),
                      ("white", 0.5 * math.exp(0.2 * rng.standard_normal()))])
        betas.append(list(rng.normal(size=3)))
    rs = ds.loglik_batch([gterms(p) for p in props], betas, nugget=1e-3)
    for b in range(9):
        ro = oracle.loglik(x, y, props[b], beta=betas[b], nugget=1e-3)
        assert rs[b].info == 0
        assert rel(G.ll_value(rs[b], 0), ro["ll_ref"]) < LL_RTOL
        assert rel(G.ll_value(rs[b], 1), ro["ll_gpml"]) < LL_RTOL
    # a batch gives bit-identical numbers to one-at-a-time calls (the schedule does not enter the sums)
    one = ds.loglik_batch([gterms(props[4])], [betas[4]], nugget=1e-3)[0]
    assert (one.Q, one.U1, one.S) == (rs[4].Q, rs[4].U1, rs[4].S)
    ds.close()


def test_golden_fixtures_through_the_c_abi(ctx):
    from gpb200 import _lib as G
    g = json.load(open(GOLDEN))
    for c in g["cases"]:
        terms = [tuple(t) for t in c["terms"]]
        ds = ctx.dataset(np.array(c["x"]), np.array(c["y"]))
        r = ds.loglik_batch([gterms(terms)], [c["beta"]] if c["beta"] else None, nugget=c["nugget_ll"])[0]
        assert r.info == 0
        assert rel(G.ll_value(r, 0), c["ll"]["ll_ref"]) < LL_RTOL
        assert rel(G.ll_value(r, 1), c["ll"]["ll_gpml"]) < LL_RTOL
        mean, var, info = ds.predict(gterms(terms), np.array(c["xstar"]), nugget=c["nugget_pred"])
        assert info == 0
        # nugget 1e-6 => cond(K) up to ~1e7: absolute tolerance scaled to the prior variance
        assert np.allclose(mean, c["pred_mean"], rtol=1e-7, atol=1e-8)
        assert np.allclose(var, c["pred_var"], rtol=1e-7, atol=1e-8)
        ds.close()


@pytest.mark.parametrize("n,m,dim", [(20, 1, 1), (64, 64, 1), (150, 200, 2), (300, 2000, 1)])
def test_predict_parity(ctx, oracle, n, m, dim):
    x, y, _ = synth(oracle, n, dim, SE_W, 900 + n)
    rng = np.random.default_rng(m)
    xs = rng.uniform(-10, 10, size=(m, dim))
    xs[0] = x[0]
    ds = ctx.dataset(x, y)
    mean, var, info = ds.predict(gterms(SE_W), xs, nugget=1e-6)
    mo, vo, io = oracle.predict(x, y, SE_W, xs, nugget=1e-6)
    assert info == 0 and io == 0
    assert np.allclose(mean, mo, rtol=1e-7, atol=1e-8) and np.allclose(var, vo, rtol=1e-7, atol=1e-8)
    ds.close()


@pytest.mark.parametrize("n", [64, 150, 257])
def test_potrf_near_the_ends_of_the_double_range(ctx, oracle, n):
    """dpotrf works wherever its square roots and quotients do; the diagonal blocks here are eliminated
    division-free (products of two entries), so blocks outside 2^-300 .. 2^300 are rescaled by an exact even
    power of two first: the factor of 2^(2k) K must be 2^k L, bit for bit, and an odd power must agree to
    rounding."""
    x, _, _ = synth(oracle, n, 2, SE_W, 4100 + n)
    K = oracle.build_cov(x, SE_W, 1e-3)
    L, info = ctx.potrf(K)
    assert info == 0
    for k in (350, -350, 480, -500):
        Ls, info = ctx.potrf(np.ldexp(K, 2 * k))
        assert info == 0
        assert np.array_equal(Ls, np.ldexp(L, k)), k
    Lodd, info = ctx.potrf(np.ldexp(K, 701))
    assert info == 0
    assert np.allclose(np.ldexp(Lodd, -350), math.sqrt(2.0) * L, rtol=1e-11, atol=1e-12 * np.abs(L).max())
    # not positive definite stays a LAPACK info, whatever the scale
    Kbad = np.ldexp(K, 800)
    Kbad[n - 1, n - 1] = -Kbad[n - 1, n - 1]
    _, info = ctx.potrf(Kbad)
    assert info == n


@pytest.mark.parametrize("n", [1, 5, 64, 100, 257])
def test_potrf_and_triangular_solves(ctx, oracle, n):
    rng = np.random.default_rng(n)
    A = rng.standard_normal((n, n))
    K = A @ A.T + n * np.eye(n)
    L, info = ctx.potrf(K)
    Lo, io = oracle.potrf_lower(K)
    assert info == 0 and io == 0
    assert np.all(np.triu(L, 1) == 0)  # Breeze zeroes the strict upper triangle
    assert np.allclose(L, Lo, rtol=1e-12, atol=1e-12)
    b = rng.standard_normal(n); B = rng.standard_normal((n, 3))
    assert np.allclose(ctx.trsm_lower(Lo, b), oracle.trsv_lower(Lo, b), rtol=1e-11, atol=1e-12)   # Predict.forwardSolve
    assert np.allclose(ctx.trsm_upper(Lo.T, b), oracle.trsv_upper(Lo.T, b), rtol=1e-11, atol=1e-12)  # Mcmc.backSolve vec
    assert np.allclose(ctx.trsm_upper(Lo.T, B), oracle.trsm_upper(Lo.T, B), rtol=1e-11, atol=1e-12)  # Mcmc.backSolve mat


def test_not_positive_definite_is_reported_not_fatal(ctx, oracle):
    from gpb200 import _lib as G
    K = np.array([[1.0, 2.0, 0.0], [2.0, 1.0, 0.0], [0.0, 0.0, 1.0]])
    _, info = ctx.potrf(K)
    assert info == 2  # LAPACK convention, same as the oracle
    x = np.arange(70.0); y = np.ones(70)
    bad = [("se", 1.0, 5.0), ("white", -2.0)]
    ds = ctx.dataset(x, y)
    rs = ds.loglik_batch([gterms(bad), gterms(SE_W)], nugget=1e-3)
    ro = oracle.loglik(x, y, bad)
    assert rs[0].info == ro["info"] > 0 and math.isnan(G.ll_value(rs[0], 0))
    assert rs[1].info == 0  # the healthy proposal in the same batch is unaffected
    assert rel(G.ll_value(rs[1], 1), oracle.loglik(x, y, SE_W)["ll_gpml"]) < LL_RTOL
    ds.close()


def test_unsupported_inputs_fail_loudly(ctx):
    from gpb200 import _lib as G
    ds = ctx.dataset(np.arange(4.0), np.ones(4))
    with pytest.raises(G.GpbError) as e:
        ds.loglik_batch([[(2, 1.0, 1.0)]])  # Matern: out of scope
    assert e.value.status == -5
    with pytest.raises(G.GpbError):
        ctx.dataset(np.ones((4, 3)), np.ones(4))  # dim 3
    ds.close()


def test_schedulers_agree_bitwise(ctx, oracle):
    x, y, _ = synth(oracle, 700, 2, SE_SE_W, 77)
    ds = ctx.dataset(x, y)
    out = []
    for sched in (0, 1, 0):
        ctx.set_sched(sched)
        r = ds.loglik_batch([gterms(SE_SE_W)] * 3)
        out.append([(v.Q, v.U1, v.S) for v in r])
    ctx.set_sched(0)
    assert out[0] == out[1] == out[2]
    ds.close()


def test_mh_accept_reject_sequence_matches_oracle(ctx, oracle):
    """north_star: identical accept/reject sequences for identical RNG streams."""
    g = json.load(open(GOLDEN))["mh"]
    terms = [tuple(t) for t in g["terms"]]
    x, y, z, u = (np.array(g[k]) for k in ("x", "y", "z", "u"))
    ds = ctx.dataset(x, y)
    for formula in (0, 1):
        want = g["by_formula"][str(formula)]
        st, ll, acc, info = ds.mh_run([gterms(terms)], g["ig"], g["delta"], z[:, None, :], u[:, None],
                                      formula=formula)
        assert info == 0
        assert np.array_equal(acc[:, 0], want["accept"])
        assert np.allclose(st[:, 0, :], want["states"], rtol=1e-13, atol=0)
        assert np.allclose(ll[:, 0], want["ll"], rtol=1e-10, atol=0)
    ds.close()


def test_mh_many_chains_match_independent_oracle_chains(ctx, oracle):
    from gpb200 import _lib as G
    n, chains, iters = 96, 6, 60
    x, y, _ = synth(oracle, n, 2, SE_SE_W, 5)
    ig = [[3.0, 4.0], [3.0, 6.0], [3.0, 4.0], [3.0, 6.0], [10.0, 6.0]]
    z, u = G.philox_streams(11, 0, chains, iters, 5)
    rng = np.random.default_rng(3)
    inits = [[("se", 3.0 * math.exp(0.1 * rng.standard_normal()), 5.5), ("se", 1.0, 0.8 * math.exp(0.1 * rng.standard_normal())),
              ("white", 0.5)] for _ in range(chains)]
    ds = ctx.dataset(x, y)
    st, ll, acc, info = ds.mh_run([gterms(t) for t in inits], ig, 0.05, z, u, formula=G.FORMULA_GPML)
    assert info == 0
    min_margin = 1.0
    for c in range(chains):
        so, lo, ao, rc = oracle.mh_run(x, y, inits[c], ig, 0.05, z[:, c, :], u[:, c], formula=1, evals_per_iter=2)
        assert rc == 0
        assert np.array_equal(acc[:, c], ao), f"chain {c}: accept/reject sequences differ"
        assert np.allclose(st[:, c, :], so, rtol=1e-12, atol=0)
        assert np.allclose(ll[:, c], lo, rtol=1e-10, atol=0)
    assert 0 < acc.sum() < acc.size
    ds.close()


def test_reference_api_mirror_end_to_end(oracle):
    """The drop-in surface: GaussianProcess.loglikelihood / Predict.fit / KernelParameters.sample with
    the reference's signatures, as the docs use them (docs/src/main/tut/docs/prediction.md:27-58)."""
    import gpb200 as gp
    rng = np.random.default_rng(8)
    xs = np.sort(rng.uniform(-10, 10, 30))
    params = gp.GaussianProcess.Parameters(gp.MeanParameters.zero, [gp.KernelParameters.se(3.0, 5.5), gp.KernelParameters.white(0.5)])
    K = oracle.build_cov(xs, SE_W, 1e-3)
    ys = np.linalg.cholesky(K) @ rng.standard_normal(30)
    observed = [gp.GaussianProcess.Data(gp.One(float(a)), float(b)) for a, b in zip(xs, ys)]
    ll = gp.GaussianProcess.loglikelihood(observed, gp.Location.euclidean)
    assert rel(ll(params), oracle.loglik(xs, ys, SE_W)["ll_ref"]) < LL_RTOL
    assert rel(ll(params, formula=gp.FORMULA_GPML), oracle.loglik(xs, ys, SE_W)["ll_gpml"]) < LL_RTOL
    fitted = gp.Predict.fit([gp.One(-1.0), gp.One(0.5)], observed, gp.Location.euclidean, params)
    mo, vo, _ = oracle.predict(xs, ys, SE_W, np.array([-1.0, 0.5]))
    assert fitted[0][0] == gp.One(-1.0)
    assert fitted[0][1].mean == pytest.approx(mo[0], rel=1e-7, abs=1e-9)
    assert fitted[1][1].sigma == pytest.approx(math.sqrt(vo[1]), rel=1e-7)
    D = gp.GaussianProcess.distanceMatrix([d.x for d in observed], gp.Location.euclidean)
    assert D.shape == (30, 30) and D[3, 7] == abs(xs[3] - xs[7])
    Kc = gp.KernelFunction.buildCov([d.x for d in observed], gp.KernelFunction.apply(params.kernelParameters), gp.Location.euclidean)
    assert np.allclose(Kc + 1e-3 * np.eye(30), K, rtol=1e-15, atol=1e-15)
    # one MH step function, host closures for prior / proposal as in SimulatedGp.scala:115-151
    step = gp.KernelParameters.sample(observed, gp.Location.euclidean, lambda ks: 0.0,
                                      lambda ks, r: [k.map(lambda v: v * math.exp(r.normal(0, 0.05))) for k in ks])
    p = params
    for _ in range(5):
        p = step(p, rng)
    assert isinstance(p, gp.GaussianProcess.Parameters) and len(p.toList()) == 3
    with pytest.raises(gp.NotConvergedException):  # a bad proposal raises
        ll(gp.GaussianProcess.Parameters(gp.MeanParameters.zero, [gp.KernelParameters.se(1.0, 50.0), gp.KernelParameters.white(-2.0)]))
    assert np.allclose(gp.Predict.forwardSolve(np.linalg.cholesky(K), ys), sl.solve_triangular(np.linalg.cholesky(K), ys, lower=True), rtol=1e-11)
    assert np.allclose(gp.Mcmc.backSolve(np.linalg.cholesky(K).T, ys), sl.solve_triangular(np.linalg.cholesky(K).T, ys), rtol=1e-11)


# ---- BASELINE.json full sizes: properties + LAPACK twin (the oracle would take minutes) ------------
def test_full_size_c2_n4096(ctx, oracle):
    """configs[1]: single GP log-marginal-likelihood + predictive, n=4096, 2-D, SE+noise."""
    from gpb200 import _lib as G
    n = 4096
    rng = np.random.default_rng(2)
    x = rng.uniform(-10, 10, size=(n, 2))
    K = oracle.build_cov(x, SE_W, 1e-3)
    L = np.linalg.cholesky(K)
    y = L @ rng.standard_normal(n)
    ds = ctx.dataset(x, y)
    r = ds.loglik_batch([gterms(SE_W)])[0]
    u = sl.solve_triangular(L, y, lower=True)
    S, C = np.log(np.diag(L)).sum(), n * 0.5 * math.log(2 * math.pi)
    print("cond(K) ~", float(np.linalg.cond(K)))
    assert r.info == 0
    assert rel(G.ll_value(r, 1), -0.5 * u @ u - S - C) < LL_RTOL
    assert rel(G.ll_value(r, 0), -0.5 * u @ u + 0.5 * (S + C) * u.sum()) < LL_RTOL
    # factor property through the matrix-level entry point: L L^T reproduces K
    Lg, info = ctx.potrf(K)
    assert info == 0
    assert np.abs(Lg @ Lg.T - K).max() < 1e-11 * np.abs(K).max() * 64
    # predictive round trip: at the training locations the posterior mean solves K a = y
    xs = rng.uniform(-10, 10, size=(512, 2))
    mean, var, info = ds.predict(gterms(SE_W), xs, nugget=1e-6)
    K6 = K + (1e-6 - 1e-3) * np.eye(n)
    d = np.sqrt(((x[:, None, :] - xs[None, :, :]) ** 2).sum(-1))
    Ks = 3.0 * np.exp(-(d * d) / 30.25)
    cf = sl.cho_factor(K6, lower=True)
    assert np.allclose(mean, Ks.T @ sl.cho_solve(cf, y), rtol=1e-6, atol=1e-7)
    assert np.allclose(var, 3.5 - np.einsum("ij,ij->j", Ks, sl.cho_solve(cf, Ks)), rtol=1e-6, atol=1e-7)
    ds.close()


def test_full_size_c3_batch_of_chains_n1024(ctx, oracle):
    """configs[2]: 256 independent proposals x n=1024 in one pass; spot-check against the oracle and
    check every proposal against the linearity property Q(c*y) = c^2 Q(y)."""
    from gpb200 import _lib as G
    n, B = 1024, 256
    x, y, _ = synth(oracle, n, 2, SE_W, 3)
    rng = np.random.default_rng(4)
    props = [[("se", 3.0 * math.exp(0.05 * rng.standard_normal()), 5.5 * math.exp(0.05 * rng.standard_normal())),
              ("white", 0.5 * math.exp(0.05 * rng.standard_normal()))] for _ in range(B)]
    ds = ctx.dataset(x, y)
    rs = ds.loglik_batch([gterms(p) for p in props])
    ds2 = ctx.dataset(x, 2.0 * y)
    rs2 = ds2.loglik_batch([gterms(p) for p in props])
    for b in range(B):
        assert rs[b].info == 0
        assert rel(rs2[b].Q, 4.0 * rs[b].Q) < 1e-13 and rs2[b].S == rs[b].S
    for b in (0, 97, 255):
        ro = oracle.loglik(x, y, props[b])
        assert rel(G.ll_value(rs[b], 0), ro["ll_ref"]) < LL_RTOL and rel(G.ll_value(rs[b], 1), ro["ll_gpml"]) < LL_RTOL
    ds.close(); ds2.close()


def test_full_size_c5_sum_of_kernels_n8192(ctx, oracle):
    """configs[4] shape (SE + SE + white, n=8192): LAPACK twin on the host for BOTH formulas (the
    as-parsed value is the parity target), and a batch of 3 proposals bit-equal to three single calls."""
    from gpb200 import _lib as G
    n = 8192
    rng = np.random.default_rng(5)
    x = rng.uniform(-15, 15, size=(n, 2))
    d2 = ((x[:, None, :] - x[None, :, :]) ** 2).sum(-1)
    d = np.sqrt(d2)

    def cov(t):
        return (t[0][1] * np.exp(-(d * d) / (t[0][2] * t[0][2])) + t[1][1] * np.exp(-(d * d) / (t[1][2] * t[1][2]))
                + np.where(d == 0, t[2][1], 0.0) + 1e-3 * np.eye(n))

    L = np.linalg.cholesky(cov(SE_SE_W))
    y = L @ rng.standard_normal(n)
    ds = ctx.dataset(x, y)
    props = [SE_SE_W, [("se", 2.7, 5.9), ("se", 1.2, 0.7), ("white", 0.55)], [("se", 3.4, 5.1), ("se", 0.8, 0.9), ("white", 0.45)]]
    batch = ds.loglik_batch([gterms(p) for p in props])
    C = n * 0.5 * math.log(2 * math.pi)
    for b, p in enumerate(props):
        single = ds.loglik_batch([gterms(p)])[0]
        assert (single.Q, single.U1, single.S, single.info) == (batch[b].Q, batch[b].U1, batch[b].S, batch[b].info)
        if b > 0:
            L = np.linalg.cholesky(cov(p))
        u = sl.solve_triangular(L, y, lower=True)
        S = np.log(np.diag(L)).sum()
        assert batch[b].info == 0
        assert rel(G.ll_value(batch[b], 1), -0.5 * u @ u - S - C) < LL_RTOL                      # GPML
        assert rel(G.ll_value(batch[b], 0), -0.5 * u @ u + 0.5 * (S + C) * u.sum()) < LL_RTOL   # as parsed (GaussianProcess.scala:126)
    ds.close()


def test_full_size_c4_n32768_predictive_10k(ctx):
    """configs[3]: single GP n=32768 (FP64 blocked Cholesky) + 10 000-point predictive mean/variance.
    The CPU oracle would need hours at this size, so the twin is cuSOLVER/cuBLAS FP64 through torch
    on the same device (an independent implementation of the same algebra), plus the
    size-independent properties Q(2y) = 4 Q(y) and S(2y) = S(y)."""
    import torch
    from gpb200 import _lib as G
    n, m = 32768, 10000
    dev = torch.device("cuda", 0)
    gen = torch.Generator(device="cpu").manual_seed(4)
    x = (torch.rand((n, 2), generator=gen, dtype=torch.float64) * 60.0 - 30.0)
    xs = (torch.rand((m, 2), generator=gen, dtype=torch.float64) * 60.0 - 30.0)
    z = torch.randn(n, generator=gen, dtype=torch.float64)
    xd, xsd = x.to(dev), xs.to(dev)

    def cov(a, b):
        d = torch.cdist(a, b)
        return 3.0 * torch.exp(-(d * d) / (5.5 * 5.5))

    K = cov(xd, xd)
    K.diagonal().add_(0.5 + 1e-3)            # white at d == 0 (no duplicates here) + nugget 1e-3
    L = torch.linalg.cholesky(K)
    y = (L @ z.to(dev))
    u = torch.linalg.solve_triangular(L, y[:, None], upper=False)[:, 0]
    Q_t, U1_t, S_t = float(u @ u), float(u.sum()), float(torch.log(L.diagonal()).sum())
    ds = ctx.dataset(x.numpy(), y.cpu().numpy())
    r = ds.loglik_batch([gterms(SE_W)], nugget=1e-3)[0]
    assert r.info == 0 and r.n == n
    C = n * 0.5 * math.log(2 * math.pi)
    assert rel(G.ll_value(r, 1), -0.5 * Q_t - S_t - C) < LL_RTOL
    assert rel(r.Q, Q_t) < LL_RTOL and rel(r.S, S_t) < LL_RTOL and abs(r.U1 - U1_t) < 1e-8
    ds2 = ctx.dataset(x.numpy(), 2.0 * y.cpu().numpy())
    r2 = ds2.loglik_batch([gterms(SE_W)], nugget=1e-3)[0]
    assert rel(r2.Q, 4.0 * r.Q) < 1e-13 and r2.S == r.S
    ds2.close()
    # predictive (Predict.fit semantics: nugget 1e-6, raw y, var = k(0) - v.v)
    K.diagonal().add_(1e-6 - 1e-3)
    L = torch.linalg.cholesky(K)
    del K
    Ks = cov(xd, xsd)                         # n x m
    V = torch.linalg.solve_triangular(L, Ks, upper=False)
    u6 = torch.linalg.solve_triangular(L, y[:, None], upper=False)[:, 0]
    mean_t = (V.T @ u6).cpu().numpy()
    var_t = (3.5 - (V * V).sum(0)).cpu().numpy()
    del V, Ks, L
    torch.cuda.empty_cache()
    mean, var, info = ds.predict(gterms(SE_W), xs.numpy(), nugget=1e-6)
    assert info == 0
    assert np.allclose(mean, mean_t, rtol=1e-7, atol=1e-8)
    assert np.allclose(var, var_t, rtol=1e-7, atol=1e-8)
    ds.close()


@pytest.mark.parametrize("n,dim", [(40, 1), (300, 2), (1000, 2)])
def test_sample_plane_parity(ctx, oracle, n, dim):
    """MeanParameters.samplePlane through the C ABI vs the numpy restatement."""
    rng = np.random.default_rng(50 + n)
    x = rng.uniform(-10, 10, size=(n, dim))
    y = 0.7 + 0.2 * x[:, 0] + rng.standard_normal(n)
    X = oracle.design_matrix(x)
    z = rng.standard_normal(dim + 1)
    beta_o, prec_o, mean_o = oracle.sample_plane(x, y, SE_W, 1.0, 4.0, z, nugget=1e-6)
    ds = ctx.dataset(x, y)
    G, w, info = ds.plane_stats(gterms(SE_W), X, nugget=1e-6)
    assert info == 0
    assert np.allclose(G + np.eye(dim + 1) / 4.0, prec_o, rtol=1e-9)
    beta, info = ds.sample_plane(gterms(SE_W), X, 1.0, 4.0, z, nugget=1e-6)
    assert info == 0 and np.allclose(beta, beta_o, rtol=1e-8, atol=1e-10)
    ds.close()


def test_draw_uses_the_cholesky_root(ctx, oracle):
    """GaussianProcess.draw = mu + root z; the device root is the Cholesky factor."""
    rng = np.random.default_rng(61)
    n = 500
    x = np.sort(rng.uniform(-10, 10, n))
    z = rng.standard_normal(n)
    mu = 0.5 + 0.1 * x
    out, info = ctx.draw(x, gterms(SE_W), z, mu, nugget=1e-3)
    _, chol_draw = oracle.draw(x, SE_W, z, beta=(0.5, 0.1), nugget=1e-3)
    assert info == 0 and np.allclose(out, chol_draw, rtol=1e-9, atol=1e-11)


def _theta(terms):
    return [(t[0], math.log(t[1]), math.log(t[2])) if t[0] == "se" else (t[0], math.log(t[1])) for t in terms]


@pytest.mark.parametrize("n,dim,terms", [(40, 1, SE_W), (130, 2, SE_W), (300, 2, SE_SE_W), (1000, 1, SE_SE_W)])
def test_mll_gradient_parity(ctx, oracle, n, dim, terms):
    """GaussianProcess.mllGradient through the C ABI vs the numpy restatement, both the
    as-executed Scala formulas and the analytic gradient; plane mean; duplicate location."""
    x, y, _ = synth(oracle, n, dim, terms, 700 + n)
    beta = (0.2, -0.05) if dim == 1 else (0.2, -0.05, 0.03)
    ds = ctx.dataset(x, y)
    from gpb200 import _lib as G
    # analytic: log parameters
    th = _theta(terms)
    go = oracle.mll_gradient(x, y, th, beta=beta, mode="analytic")
    g, info = ds.mll_gradient([gterms(th)], [beta], mode=G.GRAD_ANALYTIC)
    assert info[0] == 0
    assert np.allclose(g[0], go, rtol=1e-8, atol=1e-8 * np.abs(go).max()), (g[0], go)
    # as executed: kxx is built from the unconstrained values themselves -> keep them positive (kxx PD)
    th2 = [("se", 1.1, 0.9), ("se", 0.4, 0.2), ("white", 0.3)][: len(terms) - 1] + [("white", 0.3)]
    go2 = oracle.mll_gradient(x, y, th2, mode="as_executed")
    g2, info2 = ds.mll_gradient([gterms(th2)], mode=G.GRAD_AS_EXECUTED)
    assert info2[0] == 0
    assert np.allclose(g2[0], go2, rtol=1e-8, atol=1e-8 * np.abs(go2).max()), (g2[0], go2)
    ds.close()


def test_mll_gradient_batch_and_not_pd(ctx, oracle):
    x, y, _ = synth(oracle, 200, 1, SE_W, 812)
    ds = ctx.dataset(x, y)
    from gpb200 import _lib as G
    ths = [[("se", math.log(3.0 + b), math.log(5.5 - b)), ("white", math.log(0.5 + 0.1 * b))] for b in range(3)]
    g, info = ds.mll_gradient([gterms(t) for t in ths], mode=G.GRAD_ANALYTIC)
    assert np.all(info == 0)
    for b in range(3):
        go = oracle.mll_gradient(x, y, ths[b], mode="analytic")
        assert np.allclose(g[b], go, rtol=1e-8, atol=1e-8 * np.abs(go).max())
    # as executed with a negative unconstrained white parameter: kxx is indefinite, the reference's LU
    # would carry on; this path reports it (include/gpb200.h)
    g, info = ds.mll_gradient([gterms([("se", 1.1, 1.7), ("white", -0.7)])], mode=G.GRAD_AS_EXECUTED)
    assert info[0] > 0 and np.all(np.isnan(g[0]))
    ds.close()


def test_full_size_mll_gradient_n4096_matches_differences_of_the_device_ll(ctx, oracle):
    """Size-independent property at C2's size: the analytic gradient is the derivative of the (parity-
    checked) device log-likelihood; central differences in the log parameters."""
    from gpb200 import _lib as G
    rng = np.random.default_rng(4096)
    n = 4096
    x = np.sort(rng.uniform(0, 400, n))
    y = np.sin(x / 7.0) * 2.0 + 0.7 * rng.standard_normal(n)
    ds = ctx.dataset(x, y)
    v0 = np.array([math.log(3.0), math.log(5.5), math.log(0.5)])
    g, info = ds.mll_gradient([[(0, v0[0], v0[1]), (1, v0[2], 0.0)]], mode=G.GRAD_ANALYTIC)
    assert info[0] == 0

    def ll(v):
        r = ds.loglik_batch([[(0, math.exp(v[0]), math.exp(v[1])), (1, math.exp(v[2]), 0.0)]])[0]
        assert r.info == 0
        return G.ll_value(r, G.FORMULA_GPML)

    for p in range(3):
        e = np.zeros(3); e[p] = 1e-4
        fd = (ll(v0 + e) - ll(v0 - e)) / 2e-4
        assert abs(fd - g[0][p]) < 2e-5 * max(1.0, abs(g[0][p])), (p, fd, g[0][p])
    ds.close()


def test_hmc_chain_follows_the_oracle_driven_chain(oracle):
    """KernelParameters.sampleHmc (KernelParameters.scala:121-156, Hmc.scala): the same host leapfrog
    driven once by the device log-likelihood + gradient and once by the oracle's, same streams."""
    import gpb200 as gp
    rng = np.random.default_rng(31)
    n = 120
    xs = np.sort(rng.uniform(-10, 10, n))
    ys = np.linalg.cholesky(oracle.build_cov(xs, SE_W, 1e-3)) @ rng.standard_normal(n)
    observed = [gp.GaussianProcess.Data(gp.One(float(a)), float(b)) for a, b in zip(xs, ys)]
    init = gp.GaussianProcess.Parameters(gp.MeanParameters.zero, [gp.KernelParameters.se(3.0, 5.5), gp.KernelParameters.white(0.5)])
    prior = [gp.GradDist.normal(1.0, 2.0), gp.GradDist.normal(1.5, 2.0), gp.GradDist.normal(-0.5, 2.0)]
    m = np.eye(3)
    old_f, old_g = gp.GaussianProcess.formula, gp.GaussianProcess.gradient_mode
    gp.GaussianProcess.formula, gp.GaussianProcess.gradient_mode = gp.FORMULA_GPML, gp.GRAD_ANALYTIC
    try:
        chain = gp.KernelParameters.sampleHmc(observed, gp.Location.euclidean, init, prior, m, 4, 0.02, np.random.default_rng(5))
        dev = [next(chain) for _ in range(8)]
    finally:
        gp.GaussianProcess.formula, gp.GaussianProcess.gradient_mode = old_f, old_g

    def terms_of(v, f):
        return [("se", f(v[0]), f(v[1])), ("white", f(v[2]))]

    def ll(v):
        return sum(pr.logPdf(a) for pr, a in zip(prior, v)) + oracle.loglik(xs, ys, terms_of(v, math.exp))["ll_gpml"]

    def grad(v):
        return oracle.mll_gradient(xs, ys, terms_of(v, float), mode="analytic") + np.array([pr.gradLogPdf(a) for pr, a in zip(prior, v)])

    r2 = np.random.default_rng(5)
    t = np.log([3.0, 5.5, 0.5])
    ref = [t]
    for _ in range(7):
        t, _ = gp.Hmc.step(0.02, 4, m, ll, grad)(t, r2)
        ref.append(t)
    assert np.allclose(np.array(dev), np.array(ref), rtol=1e-7, atol=1e-9)
    assert len({tuple(np.round(v, 12)) for v in dev}) > 2   # the chain moved


def test_ehmc_device_driven_matches_oracle_driven(oracle):
    """KernelParameters.sampleEhmc (KernelParameters.scala:171-198) in the intended mode: dual-averaging
    warm-up, empirical leapfrog counts and the chain, once on the device calls and once on the oracle."""
    import warnings
    import gpb200 as gp
    rng = np.random.default_rng(32)
    n = 100
    xs = np.sort(rng.uniform(-10, 10, n))
    ys = np.linalg.cholesky(oracle.build_cov(xs, SE_W, 1e-3)) @ rng.standard_normal(n)
    observed = [gp.GaussianProcess.Data(gp.One(float(a)), float(b)) for a, b in zip(xs, ys)]
    init = gp.GaussianProcess.Parameters(gp.MeanParameters.zero, [gp.KernelParameters.se(3.0, 5.5), gp.KernelParameters.white(0.5)])
    prior = [gp.GradDist.normal(1.0, 2.0), gp.GradDist.normal(1.5, 2.0), gp.GradDist.normal(-0.5, 2.0)]

    def terms_of(v, f):
        return [("se", f(v[0]), f(v[1])), ("white", f(v[2]))]

    def ll(v):
        return sum(pr.logPdf(a) for pr, a in zip(prior, v)) + oracle.loglik(xs, ys, terms_of(v, math.exp))["ll_gpml"]

    def grad(v):
        return oracle.mll_gradient(xs, ys, terms_of(v, float), mode="analytic") + np.array([pr.gradLogPdf(a) for pr, a in zip(prior, v)])

    old = (gp.GaussianProcess.formula, gp.GaussianProcess.gradient_mode, gp.DualAverage.as_executed, gp.Ehmc.as_executed)
    gp.GaussianProcess.formula, gp.GaussianProcess.gradient_mode = gp.FORMULA_GPML, gp.GRAD_ANALYTIC
    gp.DualAverage.as_executed, gp.Ehmc.as_executed = False, False
    try:
        with warnings.catch_warnings():
            warnings.simplefilter("ignore")
            chain = gp.KernelParameters.sampleEhmc(observed, gp.Location.euclidean, init, prior, 3, 4, rng=np.random.default_rng(9), k=5)
            dev = np.array([next(chain) for _ in range(5)])
            ref_chain = gp.Ehmc.sample(3, np.eye(3), ll, grad, 4, np.log([3.0, 5.5, 0.5]), k=5, rng=np.random.default_rng(9))
            ref = np.array([next(ref_chain) for _ in range(5)])
    finally:
        gp.GaussianProcess.formula, gp.GaussianProcess.gradient_mode, gp.DualAverage.as_executed, gp.Ehmc.as_executed = old
    assert np.all(np.isfinite(dev))
    assert np.allclose(dev, ref, rtol=1e-6, atol=1e-8)


_LIGHT_SCRIPT = r"""
import json, sys, os
import numpy as np
sys.path.insert(0, sys.argv[1]); sys.path.insert(0, os.path.join(sys.argv[1], "gaussian-processes_b200"))
from gpb200 import _lib as G
ctx = G.Context(0)
out = []
for n, dim, B, m in [(300, 1, 3, 0), (1024, 2, 5, 0), (150, 2, 1, 69), (64, 1, 1, 8)]:
    rng = np.random.default_rng(1000 + n)
    x = rng.uniform(-10, 10, size=(n, dim)); y = rng.standard_normal(n)
    ds = ctx.dataset(x, y)
    props = [[(0, 3.0 + 0.1 * b, 5.5), (1, 0.5, 0.0)] for b in range(B)]
    rs = ds.loglik_batch(props, nugget=1e-3)
    rec = [[float(r.Q).hex(), float(r.U1).hex(), float(r.S).hex(), r.info] for r in rs]
    if m:
        mean, var, info = ds.predict(props[0], rng.uniform(-10, 10, size=(m, dim)), nugget=1e-3)
        rec.append([float(v).hex() for v in mean] + [float(v).hex() for v in var] + [info])
    out.append(rec)
    ds.close()
print("RESULT" + json.dumps(out))
"""


def test_light_rhs_row_tasks_change_nothing():
    """The tile row that only holds the rhs row (<= 8 rows in use) is contracted and solved as one
    8-row DMMA fragment (Geom.light_row, gpb_chol.cu).  Same operations in the same order on the live
    entries: every output is bit-identical to the full-tile path (GPB_LIGHT_ROWS=0)."""
    import subprocess
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    outs = []
    for light in ("1", "0"):
        env = dict(os.environ, GPB_LIGHT_ROWS=light)
        p = subprocess.run([sys.executable, "-c", _LIGHT_SCRIPT, root], env=env, capture_output=True, text=True, timeout=600)
        assert p.returncode == 0, p.stderr[-2000:]
        line = [l for l in p.stdout.splitlines() if l.startswith("RESULT")][-1]
        outs.append(json.loads(line[len("RESULT"):]))
    assert outs[0] == outs[1]
    assert all(rec[0][3] == 0 for rec in outs[0])


@pytest.mark.parametrize("n,batch", [(300, 2), (1024, 1), (640, 48)])
def test_schedule_trace_build_gives_the_same_results(ctx, n, batch):
    """gpb_set_trace switches to the schedule-analysis instantiations of the factorisation kernel (one per
    launch mode: one / two CTAs per SM).  Same code apart from the timestamps: the reductions must not move
    by a bit, every task must have left a record, and its clocks must be ordered."""
    rng = np.random.default_rng(n + batch)
    x = rng.uniform(-10, 10, size=(n, 2))
    y = 1.5 * rng.standard_normal(n)
    ds = ctx.dataset(x, y)
    props = [[(0, 3.0 * math.exp(0.05 * rng.standard_normal()), 5.5), (1, 0.5, 0.0)] for _ in range(batch)]
    plain = [(r.Q, r.U1, r.S, r.info) for r in ds.loglik_batch(props)]
    ctx.set_trace(1)
    try:
        traced = [(r.Q, r.U1, r.S, r.info) for r in ds.loglik_batch(props)]
        tr = ctx.get_trace().astype(np.int64)
    finally:
        ctx.set_trace(0)
    ds.close()
    assert traced == plain
    nt = (n + 63) // 64
    assert tr.shape == (batch * (nt * (nt + 1) - nt * (nt - 1) // 2), 16)
    ran = tr[:, 6] > 0                                   # void tasks (nothing to pre-accumulate) leave no release time
    assert ran.sum() >= batch * (nt * (nt + 1) // 2 + nt - 1)
    t = tr[ran]
    assert np.all(t[:, 6] >= t[:, 1]) and np.all(t[:, 7] >= t[:, 0])
    fact = t[:, 15] > 0                                  # tasks that factored a diagonal block
    assert fact.sum() == batch * nt and np.all(t[fact, 15] > t[fact, 14])


@pytest.mark.parametrize("scale", [1e200, 1e-200])
def test_loglik_with_amplitudes_near_the_ends_of_the_double_range(ctx, oracle, scale):
    """The whole likelihood path (hot-path build, factorisation with the equilibrated diagonal blocks, bordered
    solve, reductions) at kernel amplitudes where products of two covariance entries leave the double range."""
    n = 200
    rng = np.random.default_rng(77)
    x = rng.uniform(-10, 10, size=(n, 2))
    y = rng.standard_normal(n)
    terms = [("se", 3.0 * scale, 5.5), ("white", 0.5 * scale)]
    ds = ctx.dataset(x, y)
    r = ds.loglik_batch([gterms(terms)], nugget=0.0)[0]
    ro = oracle.loglik(x, y, terms, nugget=0.0)
    ds.close()
    assert r.info == 0 and ro["info"] == 0
    for got, want in ((r.Q, ro["Q"]), (r.U1, ro["U1"]), (r.S, ro["S"])):
        assert rel(got, want) < LL_RTOL, (scale, got, want)
