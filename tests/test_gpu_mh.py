"""GPU parity tests of the Metropolis-Hastings path on BASELINE.json's own configs, through the C ABI
(gpb_mh_run / gpb_mh_run_ex / gpb_mh_run_host) against the CPU oracle on identical RNG streams.

north_star: "identical accept/reject sequences for identical RNG streams".  Contract:
the sequences are identical unless some iteration's uniform lies within TAU of its acceptance
probability (|u - a| < TAU: the 1e-10 relative agreement of the log-likelihoods cannot decide such a
draw).  Every test prints min |u - a| over the run and asserts the contract.
"""
import math
import threading
from concurrent.futures import ThreadPoolExecutor

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

SE_W = [("se", 3.0, 5.5), ("white", 0.5)]
SE_SE_W = [("se", 3.0, 5.5), ("se", 1.0, 0.8), ("white", 0.5)]
IG3 = [[3.0, 4.0], [3.0, 6.0], [10.0, 6.0]]                        # SimulatedGp.scala:132-134
IG5 = [[3.0, 4.0], [3.0, 6.0], [3.0, 4.0], [3.0, 6.0], [10.0, 6.0]]
TAU = 1e-7   # tie band of the accept/reject contract


def gterms(terms):
    return [(0, t[1], t[2]) if t[0] == "se" else (1, t[1], 0.0) for t in terms]


def simulated_gp(oracle, n=300, seed=1):
    """SimulatedGp.scala:17-45: n sorted U(-10, 10) locations in 1-D, y ~ GP(0, SE(3, 5.5) + White(0.5))
    (y = L z with the nugget 1e-3; the reference draws through eigSym)."""
    rng = np.random.default_rng(seed)
    x = np.sort(rng.uniform(-10, 10, n))
    y = np.linalg.cholesky(oracle.build_cov(x, SE_W, 1e-3)) @ rng.standard_normal(n)
    return x, y


def check_contract(acc_gpu, acc_orc, u, prob, label):
    """identical unless |u - a| < TAU; prints the margin (SURVEY.md 7 'hard parts')."""
    margin = np.abs(u - prob)
    margin = margin[np.isfinite(margin)]
    mm = float(margin.min()) if margin.size else float("inf")
    print(f"[accept/reject parity] {label}: {len(acc_orc)} iterations, {int(np.sum(acc_orc))} accepted, min|u-a| = {mm:.3e}")
    if mm >= TAU:
        assert np.array_equal(acc_gpu, acc_orc), f"{label}: accept/reject sequences differ with min|u-a| = {mm:.3e}"
    else:  # a draw inside the tie band: only the prefix before it is decided
        k = int(np.argmax(np.abs(u - prob) < TAU))
        assert np.array_equal(acc_gpu[:k], acc_orc[:k]), f"{label}: sequences differ before the first tie at {k}"


@pytest.mark.parametrize("thinned", [False, True])
@pytest.mark.parametrize("formula", [0, 1])
def test_c1_shipped_example_accept_reject(ctx, oracle, thinned, formula):
    """configs[0]: SimulatedGp.scala -- 1-D, n = 300 and the shipped thinning to n = 20 (:112), 2 chains
    (:164), log-normal random walk delta = 0.05 (:115-130), inverse-gamma priors (:132-151), 500
    iterations, both likelihood formulas; the oracle runs Breeze's 2 evaluations per iteration."""
    from gpb200 import _lib as G
    x, y = simulated_gp(oracle)
    if thinned:
        keep = [(i + 1) % 15 == 0 for i in range(len(x))]     # SimulatedGp.scala:112
        x, y = x[keep], y[keep]
        assert len(x) == 20
    chains, iters = 2, 500
    z, u = G.philox_streams(1, 0, chains, iters, 3)
    ds = ctx.dataset(x, y)
    runs = {"default": ds.mh_run_ex([gterms(SE_W)] * chains, IG3, 0.05, iters, z, u, formula=formula)}
    if thinned:   # n <= 64 takes the one-launch kernel by default: the general path must agree with it
        assert runs["default"]["path"] == 1
        runs["general"] = ds.mh_run_ex([gterms(SE_W)] * chains, IG3, 0.05, iters, z, u, formula=formula,
                                       small_n=G.SMALL_OFF)
        assert runs["general"]["path"] == 0
    for c in range(chains):
        so, lo, ao, rc, prob, done = oracle.mh_run(x, y, SE_W, IG3, 0.05, z[:, c], u[:, c], formula=formula,
                                                   evals_per_iter=2, with_prob=True)
        assert rc == 0 and done == iters
        for name, r in runs.items():
            assert r["status"] == 0 and r["info"] == 0 and np.all(r["stop"] == -1)
            check_contract(r["accept"][:, c], ao, u[:, c], prob, f"C1 n={len(x)} formula={formula} chain={c} path={name}")
            assert np.allclose(r["states"][:, c], so, rtol=1e-12, atol=0)
            assert np.allclose(r["ll"][:, c], lo, rtol=1e-10, atol=0)
    ds.close()


def test_c3_accept_reject_n1024(ctx, oracle):
    """configs[2] at full size: n = 1024, 2-D, 4 chains x 50 iterations against independent oracle
    chains (one host thread each; ll(t) cached on the oracle side: the same numbers as Breeze's
    re-evaluation, half the CPU time)."""
    from gpb200 import _lib as G
    n, chains, iters = 1024, 4, 50
    rng = np.random.default_rng(3)
    x = rng.uniform(-10, 10, size=(n, 2))
    y = np.linalg.cholesky(oracle.build_cov(x, SE_W, 1e-3)) @ rng.standard_normal(n)
    z, u = G.philox_streams(3, 0, chains, iters, 3)
    r0 = np.random.default_rng(17)
    inits = [[("se", 3.0 * math.exp(0.1 * r0.standard_normal()), 5.5 * math.exp(0.1 * r0.standard_normal())),
              ("white", 0.5 * math.exp(0.1 * r0.standard_normal()))] for _ in range(chains)]
    ds = ctx.dataset(x, y)
    for formula in (0, 1):
        r = ds.mh_run_ex([gterms(t) for t in inits], IG3, 0.05, iters, z, u, formula=formula)
        assert r["status"] == 0 and r["info"] == 0 and r["path"] == 0
        oracle.lib()
        with ThreadPoolExecutor(max_workers=chains) as ex:
            outs = list(ex.map(lambda c: oracle.mh_run(x, y, inits[c], IG3, 0.05, z[:, c], u[:, c], formula=formula,
                                                       evals_per_iter=1, with_prob=True), range(chains)))
        for c, (so, lo, ao, rc, prob, done) in enumerate(outs):
            assert rc == 0
            check_contract(r["accept"][:, c], ao, u[:, c], prob, f"C3 n=1024 formula={formula} chain={c}")
            assert np.allclose(r["states"][:, c], so, rtol=1e-12, atol=0)
            assert np.allclose(r["ll"][:, c], lo, rtol=1e-10, atol=0)
        assert 0 < r["accept"].sum() < r["accept"].size
    ds.close()


def test_nan_likelihood_rejects_like_the_jvm(ctx, oracle):
    """Breeze's metropolis on the JVM rejects when exp(new - old) is NaN (scala.math.min
    propagates NaN, `u < NaN` is false).  y ~ 1e308 makes Q and U1 overflow: the as-parsed likelihood
    is -inf + inf = NaN with info == 0, for the initial state and for every proposal."""
    from gpb200 import _lib as G
    n, iters = 40, 30
    x = np.linspace(-5, 5, n)
    y = np.full(n, 1e308)
    z, u = G.philox_streams(5, 0, 1, iters, 3)
    ds = ctx.dataset(x, y)
    for small in (G.SMALL_AUTO, G.SMALL_OFF):
        r = ds.mh_run_ex([gterms(SE_W)], IG3, 0.05, iters, z, u, formula=0, small_n=small)
        assert r["status"] == 0 and r["info"] == 0
        assert np.all(np.isnan(r["ll"])) and r["accept"].sum() == 0
        assert np.all(r["states"] == np.array([3.0, 5.5, 0.5]))
    so, lo, ao, rc = oracle.mh_run(x, y, SE_W, IG3, 0.05, z[:, 0], u[:, 0], formula=0)
    assert rc == 0 and ao.sum() == 0
    ds.close()


@pytest.mark.parametrize("n", [30, 150])
def test_not_pd_proposal_stops_that_chain_only(ctx, oracle, n):
    """include/gpb200.h, gpb_mh_run: on_notpd == 0 mirrors the reference -- `cholesky` throws on the
    chain's thread, that chain ends at the failing iteration, the other chains run on; on_notpd == 1
    rejects the proposal instead.  The proposal of chain 0 at iteration 7 is made infinite."""
    from gpb200 import _lib as G
    x, y = simulated_gp(oracle, n=n, seed=9)
    chains, iters, K = 2, 20, 7
    z, u = G.philox_streams(2, 0, chains, iters, 3)
    z[K, 0, 0] = 1e7                      # h * exp(0.05 * 1e7) = inf -> K has NaN pivots
    ds = ctx.dataset(x, y)
    r = ds.mh_run_ex([gterms(SE_W)] * chains, IG3, 0.05, iters, z, u, formula=1, raise_on_notpd=False)
    assert r["status"] == G.E_NOTPD and r["info"] > 0 and r["stopped_chains"] == 1
    assert list(r["stop"]) == [K, -1]
    so, lo, ao, rc, _, done = oracle.mh_run(x, y, SE_W, IG3, 0.05, z[:, 0], u[:, 0], formula=1, with_prob=True)
    assert rc > 0 and done == K           # the oracle (like the reference) gives up at the same iteration
    assert np.array_equal(r["accept"][:K, 0], ao[:K]) and np.allclose(r["states"][:K, 0], so[:K], rtol=1e-12)
    assert np.all(np.isnan(r["states"][K:, 0])) and np.all(np.isnan(r["ll"][K:, 0])) and r["accept"][K:, 0].sum() == 0
    s1, l1, a1, rc1 = oracle.mh_run(x, y, SE_W, IG3, 0.05, z[:, 1], u[:, 1], formula=1)
    assert rc1 == 0 and np.array_equal(r["accept"][:, 1], a1) and np.allclose(r["states"][:, 1], s1, rtol=1e-12)
    with pytest.raises(G.GpbError) as e:   # the Python layer raises like the reference throws
        ds.mh_run_ex([gterms(SE_W)] * chains, IG3, 0.05, iters, z, u, formula=1)
    assert e.value.status == G.E_NOTPD
    # the legacy entry point reports the same way
    with pytest.raises(G.GpbError):
        ds.mh_run([gterms(SE_W)] * chains, IG3, 0.05, z, u, formula=1)
    # reject mode: the chain steps over the bad proposal
    r2 = ds.mh_run_ex([gterms(SE_W)] * chains, IG3, 0.05, iters, z, u, formula=1, on_notpd=1)
    assert r2["status"] == 0 and r2["info"] > 0 and np.all(r2["stop"] == -1)
    assert r2["accept"][K, 0] == 0 and np.array_equal(r2["states"][K, 0], r2["states"][K - 1, 0])
    assert np.array_equal(r2["accept"][:K], r["accept"][:K]) and np.all(np.isfinite(r2["states"]))
    assert np.array_equal(r2["accept"][:, 1], a1)
    ds.close()


def test_mh_run_ex_equals_legacy_entry_point_and_flush_size_does_not_matter(ctx, oracle):
    from gpb200 import _lib as G
    x, y = simulated_gp(oracle, n=130, seed=4)
    chains, iters = 5, 70
    z, u = G.philox_streams(8, 0, chains, iters, 3)
    inits = [gterms([("se", 3.0 + 0.1 * c, 5.5), ("white", 0.5)]) for c in range(chains)]
    ds = ctx.dataset(x, y)
    st, ll, acc, info = ds.mh_run(inits, IG3, 0.05, z, u, formula=0, beta=(0.1, -0.02))
    b0 = np.tile([0.1, -0.02], (chains, 1))
    for T in (0, 1, 7, 32, 70, 1000):
        r = ds.mh_run_ex(inits, IG3, 0.05, iters, z, u, beta0=b0, formula=0, flush_every=T)
        assert np.array_equal(r["states"], st) and np.array_equal(r["ll"], ll) and np.array_equal(r["accept"], acc)
        assert r["flushes"] == math.ceil(iters / (T if 0 < T <= iters else (32 if T == 0 else iters)))
    # zero iterations: a valid (empty) run
    r = ds.mh_run_ex(inits, IG3, 0.05, 0, z[:0], u[:0], beta0=b0)
    assert r["status"] == 0 and r["states"].shape == (0, chains, 3)
    ds.close()


def test_device_philox_streams(ctx):
    """GPB_RNG_PHILOX: uniforms bit-identical to the host generator, normals within 2 ulp-ish (device
    log / sincos), and the loop consumes exactly what gpb_philox_streams_device returns."""
    from gpb200 import _lib as G
    zh, uh = G.philox_streams(77, 3, 6, 40, 5)
    zd, ud = ctx.philox_streams_device(77, 3, 6, 40, 5)
    assert np.array_equal(uh, ud)
    assert np.all(np.abs(zh - zd) <= 8 * np.spacing(np.maximum(np.abs(zh), 1.0)))
    assert 0.8 < zd.std() < 1.2 and abs(zd.mean()) < 0.15


@pytest.mark.parametrize("n", [20, 200])
def test_device_rng_run_equals_host_stream_run(ctx, oracle, n):
    from gpb200 import _lib as G
    x, y = simulated_gp(oracle, n=n, seed=6)
    chains, iters, chain0 = 4, 90, 10
    inits = [gterms(SE_W)] * chains
    ds = ctx.dataset(x, y)
    zd, ud = ctx.philox_streams_device(123, chain0, chains, iters, 3)
    a = ds.mh_run_ex(inits, IG3, 0.05, iters, zd, ud, formula=1)
    b = ds.mh_run_ex(inits, IG3, 0.05, iters, rng=G.RNG_PHILOX, seed=123, chain0=chain0, formula=1)
    assert a["path"] == b["path"] == (1 if n <= 64 else 0)
    assert np.array_equal(a["states"], b["states"]) and np.array_equal(a["accept"], b["accept"]) and np.array_equal(a["ll"], b["ll"])
    # and the oracle follows the same chain from the device-generated streams
    so, lo, ao, rc = oracle.mh_run(x, y, SE_W, IG3, 0.05, zd[:, 2], ud[:, 2], formula=1)
    assert rc == 0 and np.array_equal(b["accept"][:, 2], ao) and np.allclose(b["states"][:, 2], so, rtol=1e-12)
    ds.close()


@pytest.mark.parametrize("dim", [1, 2])
def test_mcmc_sample_with_plane_mean_gibbs_step(ctx, oracle, dim):
    """Mcmc.sample (Mcmc.scala:63-76): MeanParameters.sample (Gibbs) then KernelParameters.sample, every
    iteration, device resident (gpb_mh_run_ex, mean_step = 1) against the host-stepped restatement
    driven by the oracle: identical accept sequence, beta to 1e-8."""
    from gpb200 import _lib as G
    n, chains, iters = 200, 3, 50
    rng = np.random.default_rng(40 + dim)
    x = rng.uniform(-10, 10, size=(n, dim))
    true_beta = np.array([0.7, 0.2, -0.1])[: dim + 1]
    mu = true_beta[0] + x @ true_beta[1:]
    y = mu + np.linalg.cholesky(oracle.build_cov(x, SE_W, 1e-3)) @ rng.standard_normal(n)
    X = oracle.design_matrix(x)                     # makeDesignMatrix, layout quirk included
    d = dim + 1
    z, u = G.philox_streams(21, 0, chains, iters, 3 + d)
    b0 = np.zeros((chains, d))
    inits = [[("se", 3.0 + 0.2 * c, 5.5 - 0.3 * c), ("white", 0.5)] for c in range(chains)]
    ds = ctx.dataset(x, y)
    for formula in (1, 0):
        r = ds.mh_run_ex([gterms(t) for t in inits], IG3, 0.05, iters, z, u, beta0=b0, design=X, formula=formula,
                         mean_step=True, mean_prior=(0.0, 4.0), flush_every=16)
        assert r["status"] == 0 and r["info"] == 0 and r["states"].shape == (iters, chains, 3 + d)
        for c in range(chains):
            so, lo, ao = oracle.mcmc_sample(x, y, inits[c], b0[c], IG3, 0.05, z[:, c], u[:, c], 0.0, 4.0, formula=formula)
            assert np.array_equal(r["accept"][:, c], ao), f"formula {formula} chain {c}"
            assert np.allclose(r["states"][:, c, :3], so[:, :3], rtol=1e-12, atol=0)
            assert np.allclose(r["states"][:, c, 3:], so[:, 3:], rtol=1e-8, atol=1e-8)
            assert np.allclose(r["ll"][:, c], lo, rtol=1e-9, atol=1e-9)
        assert r["accept"].sum() > 0
    ds.close()


def test_host_callback_sampler_takes_arbitrary_closures(ctx, oracle):
    """KernelParameters.sample(obs, dist, prior, prop) takes closures (KernelParameters.scala:231-235):
    gpb_mh_run_host with the example's proposal / prior written as host callbacks follows the oracle."""
    from gpb200 import _lib as G
    x, y = simulated_gp(oracle, n=90, seed=12)
    chains, iters = 3, 40
    z, u = G.philox_streams(31, 0, chains, iters, 3)
    shape, scale = np.array(IG3)[:, 0], np.array(IG3)[:, 1]

    def propose(it, cur):
        return np.array([[cur[c, p] * math.exp(0.05 * z[it, c, p]) for p in range(3)] for c in range(chains)])

    def log_prior(par):
        out = []
        for c in range(par.shape[0]):
            terms = [("se", par[c, 0], par[c, 1]), ("white", par[c, 2])]
            out.append(oracle.prior(terms, IG3))   # the prior is the caller's own code here
        return np.array(out)

    ds = ctx.dataset(x, y)
    st, ll, acc, info = ds.mh_run_host([gterms(SE_W)] * chains, propose, log_prior, u, formula=1)
    assert info == 0
    for c in range(chains):
        so, lo, ao, rc = oracle.mh_run(x, y, SE_W, IG3, 0.05, z[:, c], u[:, c], formula=1)
        assert rc == 0 and np.array_equal(acc[:, c], ao)
        assert np.allclose(st[:, c], so, rtol=1e-13, atol=0) and np.allclose(ll[:, c], lo, rtol=1e-10, atol=0)
    del shape, scale
    ds.close()


# ---- the gather behind the C ABI ---------------------------------------------------
def test_gather_through_a_one_rank_communicator(ctx, oracle):
    """NCCL bound inside libgpb200.so: with a 1-rank communicator the gathered run is the plain run."""
    from gpb200 import _lib as G
    x, y = simulated_gp(oracle, n=100, seed=3)
    chains, iters = 6, 45
    inits = [gterms(SE_W)] * chains
    ds = ctx.dataset(x, y)
    plain = ds.mh_run_ex(inits, IG3, 0.05, iters, rng=G.RNG_PHILOX, seed=9, formula=1, flush_every=16)
    ctx.comm_init_rank(1, 0, G.comm_unique_id())
    try:
        assert ctx.comm_info() == (0, 1, True)
        g = ds.mh_run_ex(inits, IG3, 0.05, iters, rng=G.RNG_PHILOX, seed=9, formula=1, flush_every=16,
                         gather=True, total_chains=chains)
        assert np.array_equal(g["states"], plain["states"]) and np.array_equal(g["accept"], plain["accept"])
        assert np.array_equal(ctx.gather([1.5, -2.0]), np.array([[1.5, -2.0]]))
    finally:
        ctx.comm_destroy()
    assert ctx.comm_info()[2] is False
    with pytest.raises(G.GpbError) as e:   # no communicator: refused, not emulated
        ds.mh_run_ex(inits, IG3, 0.05, iters, rng=G.RNG_PHILOX, seed=9, gather=True, total_chains=chains)
    assert e.value.status == G.E_NOCOMM
    ds.close()


def test_two_gpu_sharded_chains_gather_device_to_device(oracle):
    """Row (e): 9 chains block-partitioned over 2 GPUs (5 + 4), one Context per device driven from two
    host threads (the JVM's threading model), NCCL all-gather of every ring segment inside
    gpb_mh_run_ex: both ranks end up with all 9 chains, equal to one GPU running all of them."""
    import torch
    from gpb200 import _lib as G
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    x, y = simulated_gp(oracle, n=300, seed=2)
    total, iters = 9, 64
    inits = [gterms([("se", 3.0 + 0.05 * c, 5.5), ("white", 0.5)]) for c in range(total)]
    uid = G.comm_unique_id()
    out, errs = [None, None], []

    def rank_main(rank):
        try:
            c = G.Context(rank)
            c.comm_init_rank(2, rank, uid)
            first, count = G.comm_shard(total, 2, rank)
            d = c.dataset(x, y)
            out[rank] = d.mh_run_ex(inits[first:first + count], IG3, 0.05, iters, rng=G.RNG_PHILOX, seed=4,
                                    chain0=first, formula=1, flush_every=16, gather=True, total_chains=total)
            out[rank]["gathered_ll"] = c.gather([float(rank), 10.0 + rank])
            d.close(); c.comm_destroy(); c.close()
        except Exception as e:  # noqa: BLE001
            errs.append((rank, e))

    th = [threading.Thread(target=rank_main, args=(r,)) for r in range(2)]
    [t.start() for t in th]
    [t.join(timeout=300) for t in th]
    assert not errs, errs
    c0 = G.Context(0)
    d0 = c0.dataset(x, y)
    ref = d0.mh_run_ex(inits, IG3, 0.05, iters, rng=G.RNG_PHILOX, seed=4, chain0=0, formula=1)
    d0.close(); c0.close()
    for r in range(2):
        assert out[r]["states"].shape == (iters, total, 3)
        assert np.array_equal(out[r]["states"], ref["states"]) and np.array_equal(out[r]["accept"], ref["accept"])
        assert np.array_equal(out[r]["ll"], ref["ll"])
        assert np.array_equal(out[r]["gathered_ll"], np.array([[0.0, 10.0], [1.0, 11.0]]))
