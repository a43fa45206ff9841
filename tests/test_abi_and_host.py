"""CPU tests of the boundary: the C-ABI library loads and exports every declared symbol, its
host-only entry points work, it refuses to run without a B200 (no CPU fallback), and the host-side
mirror of the reference API keeps the reference's parameter conventions."""
import ctypes as C
import math
import os
import re

import numpy as np
import pytest
from hypothesis import given, settings, strategies as st

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_library_exports_every_symbol_the_header_declares():
    from gpb200 import _lib as G
    L = G.load()
    hdr = open(os.path.join(ROOT, "include", "gpb200.h")).read()
    hdr = re.sub(r"/\*.*?\*/", "", hdr, flags=re.S)
    declared = sorted(set(re.findall(r"\b(gpb_[a-z0-9_]+)\s*\(", hdr)))
    assert len(declared) >= 20
    for name in declared:
        assert hasattr(L, name), f"{name} declared in include/gpb200.h but not exported"
    assert sorted(G.EXPORTS) == declared


def test_no_cpu_fallback_without_a_gpu():
    import torch
    from gpb200 import _lib as G
    if torch.cuda.is_available():
        pytest.skip("a GPU is visible: the refusal path cannot be exercised")
    with pytest.raises(G.GpbError) as e:
        G.Context(0)
    assert e.value.status == -2  # GPB_E_NODEVICE


def test_product_package_never_touches_the_oracle():
    pkg = os.path.join(ROOT, "gaussian-processes_b200")
    for d, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h", ".scala")):
                src = open(os.path.join(d, f)).read()
                for needle in ("import oracle", "from oracle", "gp_oracle", "orc_", "/oracle"):
                    assert needle not in src, (d, f, needle)


def test_ll_value_formulas():
    from gpb200 import _lib as G
    r = G.LL(10.0, -2.0, 3.5, 7, 0)
    Cc = 7 * 0.5 * math.log(2 * math.pi)
    assert G.ll_value(r, G.FORMULA_GPML) == -0.5 * 10.0 - 3.5 - Cc
    assert G.ll_value(r, G.FORMULA_AS_PARSED) == -0.5 * 10.0 + 0.5 * (3.5 + Cc) * -2.0
    assert math.isnan(G.ll_value(G.LL(1.0, 1.0, 1.0, 7, 3), 0))


def test_philox_streams_are_counter_based():
    from gpb200 import _lib as G
    z, u = G.philox_streams(42, 0, 4, 50, 3)
    z2, u2 = G.philox_streams(42, 2, 2, 50, 3)  # chains 2..3 regenerated on another rank
    assert np.array_equal(z[:, 2:], z2) and np.array_equal(u[:, 2:], u2)
    z3, _ = G.philox_streams(43, 0, 4, 50, 3)
    assert not np.array_equal(z, z3)
    assert np.all((u > 0) & (u < 1)) and abs(z.mean()) < 0.2 and 0.8 < z.std() < 1.2
    # known-answer: Philox4x32-10 zero counter / zero key (Random123 kat vectors)
    L = G.load()
    zz = np.zeros(2); uu = np.zeros(1)
    assert L.gpb_philox_streams(0, 0, 1, 1, 2, zz.ctypes.data_as(C.POINTER(C.c_double)),
                                uu.ctypes.data_as(C.POINTER(C.c_double))) == 0


def test_mh_config_layout_is_the_published_one():
    """gpb_mh_config / gpb_mh_result as a Panama / JNI binding must lay them out (INTEGRATION.md,
    scala/Gpb200.scala): natural alignment, 88 and 20 bytes."""
    from gpb200 import _lib as G
    off = {f[0]: getattr(G.MhConfig, f[0]).offset for f in G.MhConfig._fields_}
    assert off == {"struct_size": 0, "formula": 4, "on_notpd": 8, "rng": 12, "delta": 16, "nugget": 24, "seed": 32,
                   "chain0": 40, "mean_step": 44, "mean_prior_mean": 48, "mean_prior_var": 56, "mean_nugget": 64,
                   "flush_every": 72, "gather": 76, "total_chains": 80, "small_n": 84}
    assert C.sizeof(G.MhConfig) == 88 and C.sizeof(G.MhResult) == 20


def test_comm_entry_points_without_a_device():
    """Host-only parts of the multi-GPU ABI: the block partition, and loud refusals (no emulation)."""
    from gpb200 import _lib as G
    L = G.load()
    assert [G.comm_shard(256, 8, r) for r in (0, 7)] == [(0, 32), (224, 32)]
    assert [G.comm_shard(10, 4, r) for r in range(4)] == [(0, 3), (3, 3), (6, 2), (8, 2)]
    with pytest.raises(G.GpbError):
        G.comm_shard(10, 4, 4)
    assert L.gpb_comm_init_rank(None, 2, 0, None) == -1          # GPB_E_INVALID
    assert L.gpb_mh_run_ex(None, None, None, 1, 1, None, 1, None, 0, None, None, None, None, None, None, None, None, None) == -1
    assert b"communicator" in L.gpb_strerror(-8)


def test_strerror_and_version():
    from gpb200 import _lib as G
    L = G.load()
    assert L.gpb_strerror(0) == b"ok"
    assert b"no CPU fallback" in L.gpb_strerror(-2)
    assert b"sm_100a" in L.gpb_version()


# ---- host mirror of the reference API ------------------------------------------------------------
def test_location_and_dimension_mismatch():
    from gpb200 import Location, One, Two
    assert Location.euclidean(One(1.0), One(4.0)) == 3.0
    assert Location.euclidean(Two(0.0, 0.0), Two(3.0, 4.0)) == 5.0
    with pytest.raises(Exception, match="Dimension mismatch"):
        Location.euclidean(One(1.0), Two(1.0, 2.0))  # Location.scala:32


def test_cov_fn_closure_is_the_reference_scalar_function():
    from gpb200 import KernelFunction, SquaredExp, White
    k = KernelFunction.apply([SquaredExp(3.0, 5.5), White(0.5)])
    assert k(0.0) == 3.5 and k(2.0) == 3.0 * math.exp(-4.0 / 30.25)


def test_arbitrary_dist_closures_are_refused_not_emulated():
    from gpb200 import GaussianProcess, One
    with pytest.raises(NotImplementedError):
        GaussianProcess.distanceMatrix([One(0.0), One(1.0)], lambda a, b: 0.0)


kernel_term = st.one_of(
    st.tuples(st.just("se"), st.floats(2, 10), st.floats(2, 10)),
    st.tuples(st.just("white"), st.floats(2, 10)))


@settings(max_examples=100, deadline=None)
@given(st.lists(kernel_term, min_size=3, max_size=3), st.lists(st.floats(-5, 5), min_size=3, max_size=3))
def test_params_array_round_trip(terms, beta):
    """The reference's only live test (core/src/test/scala/ParameterTest.scala:57-69):
    arrayToParams(p, paramsToArray(p)) === p, flattening = mean first, then SE -> (h, sigma),
    White -> (sigma)."""
    from gpb200 import GaussianProcess, KernelParameters, Plane, SquaredExp, White
    ks = [SquaredExp(t[1], t[2]) if t[0] == "se" else White(t[1]) for t in terms]
    p = GaussianProcess.Parameters(Plane(beta), ks)
    arr = KernelParameters.paramsToArray(p)
    flat = list(beta)
    for t in terms:
        flat.extend(t[1:])
    assert list(arr) == flat
    assert KernelParameters.arrayToParams(p, arr) == p


def test_chain_files_round_trip(tmp_path):
    """`${prefix}_${chain}.csv`, header of parameter names, one row per iteration."""
    import numpy as np
    from gpb200 import chains as CH
    names = CH.param_names([(0, 1.0, 1.0), (0, 1.0, 1.0), (1, 1.0, 0.0)], nbeta=2)
    assert names == ["h", "sigma", "h_1", "sigma_1", "sigma_y", "beta_0", "beta_1"]
    st = np.random.default_rng(0).standard_normal((5, 3, 3))
    paths = CH.write_chains(str(tmp_path / "gpmcmc"), st, ["h", "sigma", "sigma_y"], chain0=4)
    assert [p.rsplit("_", 1)[1] for p in paths] == ["4.csv", "5.csv", "6.csv"]
    nm, back = CH.read_chain(paths[1])
    assert nm == ["h", "sigma", "sigma_y"] and np.array_equal(back, st[:, 1])


def test_design_matrix_mirrors_the_reference_layout():
    """MeanParameters.makeDesignMatrix flattens row-major and reads column-major (MeanParameters.scala:62)."""
    import numpy as np
    from gpb200 import api as gp
    X = gp.MeanParameters.makeDesignMatrix([gp.Two(1.0, 2.0), gp.Two(3.0, 4.0), gp.Two(5.0, 6.0)])
    assert np.array_equal(X, np.array([[1, 1, 4], [1, 2, 5], [1, 3, 6.0]]))
    X1 = gp.MeanParameters.makeDesignMatrix([gp.One(7.0), gp.One(8.0)])
    assert np.array_equal(X1, np.array([[1, 7], [1, 8.0]]))


def test_hmc_family_host_logic_on_a_toy_target():
    """Hmc / DualAverage / Ehmc (Hmc.scala, DualAveraging.scala, Ehmc.scala) are host control flow around
    two callbacks; checked here on a standard-normal target (no device): closed-form numbers, the
    as-executed field permutation of DualAverage.step, and that the intended mode tunes and samples."""
    import math
    import warnings
    import gpb200 as gp
    pos = lambda t: -0.5 * float(np.sum(np.asarray(t) ** 2))
    grad = lambda t: -np.asarray(t)
    # one leapfrog on a harmonic oscillator, by hand
    t1, p1 = gp.Hmc.leapfrog(0.1, grad, np.array([1.0]), np.array([0.0]))
    assert np.allclose(t1, [1.0 - 0.005]) and np.allclose(p1, [-0.05 - 0.05 * 0.995])
    assert gp.Hmc.kinetic([3.0, 4.0]) == 12.5
    assert gp.Hmc.logAcceptance([float("nan")], [0.0], [0.0], [0.0], pos) == -math.inf
    # updateEps, DualAveraging.scala:38-45 by hand for m = 4
    hm, le, lb = gp.DualAverage.updateEps(4, 1.0, 0.65, 0.25)(0.1, 0.0, 0.5)
    ra = 1 / 14.0
    assert hm == (1 - ra) * 0.1 + ra * 0.4 and le == 1.0 - (2.0 * hm) / 0.05
    assert lb == 4 ** -0.75 * le + (1 - 4 ** -0.75) * 0.5
    assert gp.Ehmc.isUTurn([0.0, 0.0], [1.0, 0.0], [-1.0, 5.0]) and not gp.Ehmc.isUTurn([0.0], [1.0], [1.0])
    s0 = gp.DualAverageState(1, np.array([0.3]), math.log(0.5), 0.0, 0.0)
    old = (gp.DualAverage.as_executed, gp.Ehmc.as_executed)
    try:
        with warnings.catch_warnings():
            warnings.simplefilter("ignore")
            gp.DualAverage.as_executed = True
            s1 = gp.DualAverage.step(3, 0.0, np.eye(1), pos, grad)(s0, np.random.default_rng(0))
            gp.DualAverage.as_executed = False
            s2 = gp.DualAverage.step(3, 0.0, np.eye(1), pos, grad)(s0, np.random.default_rng(0))
            assert (s1.logeps, s1.logepsbar, s1.hm) == (s2.hm, s2.logeps, s2.logepsbar)   # the positional mix-up
            gp.Ehmc.as_executed = False
            eps = gp.DualAverage.tuneStepsize(150, np.array([0.5, -0.3]), 5, np.eye(2), pos, grad, rng=np.random.default_rng(2))
            assert 0.2 < eps < 4.0
            ch = gp.Ehmc.sample(5, np.eye(2), pos, grad, 150, np.array([0.5, -0.3]), k=100, rng=np.random.default_rng(3))
            xs = np.array([next(ch) for _ in range(1500)])
            assert np.all(np.abs(xs.mean(0)) < 0.2) and np.all(np.abs(xs.std(0) - 1.0) < 0.2)
            # as executed the tuner diverges and the chain stops moving
            gp.DualAverage.as_executed, gp.Ehmc.as_executed = True, True
            ch = gp.Ehmc.sample(5, np.eye(2), pos, grad, 60, np.array([0.5, -0.3]), k=20, rng=np.random.default_rng(3))
            xs = np.array([next(ch) for _ in range(20)])
            assert np.allclose(xs, xs[0])
    finally:
        gp.DualAverage.as_executed, gp.Ehmc.as_executed = old


@pytest.mark.parametrize("n,appended,batch", [(1, 1, 1), (64, 1, 1), (65, 1, 3), (300, 1, 2), (300, 0, 1), (500, 130, 2),
                                              (1024, 1, 5), (130, 200, 1)])
def test_ticket_order_of_the_factorisation_kernel_is_topological(n, appended, batch):
    """The persistent factorisation kernel draws tickets in increasing order and a CTA only ever waits for
    results of lower tickets -- that is what makes its spin-waits deadlock-free.  Checked on the kernel's
    own decoder (gpb_schedule_task runs the same decode_task on the host): every tile is produced exactly
    once, and every input of every task is produced by a lower ticket."""
    from gpb200 import _lib as G
    PLAIN, DIAG0, PART, FUSED = 0, 1, 2, 3
    sched = G.schedule(n, appended, batch)
    nt = (n + 63) // 64
    rt = nt + (appended + 63) // 64
    assert len(sched) == batch * (nt * rt - nt * (nt - 1) // 2)
    final, partial = {}, {}          # (b, i, j) -> ticket that makes the tile final / pre-accumulates it
    for t, (b, i, j, kind) in enumerate(sched):
        b, i, j, kind = int(b), int(i), int(j), int(kind)
        assert 0 <= b < batch and 0 <= j < nt and j <= i < rt
        if kind == DIAG0:
            assert (i, j) == (0, 0)
            assert final.setdefault((b, 0, 0), t) == t
        elif kind == PART:
            assert i == j and 1 <= j < nt
            assert partial.setdefault((b, i, i), t) == t
        elif kind == FUSED:
            assert i == j + 1 and i < nt
            assert final.setdefault((b, i, j), t) == t and final.setdefault((b, i, i), t) == t
        else:
            assert i > j and not (i == j + 1 and i < nt)
            assert final.setdefault((b, i, j), t) == t
    for b in range(batch):           # every tile of the lower trapezoid exactly once
        for j in range(nt):
            for i in range(j, rt):
                assert (b, i, j) in final, (b, i, j)
            if j >= 1:
                assert (b, j, j) in partial
    assert len(final) == batch * (nt * rt - nt * (nt - 1) // 2)
    for t, (b, i, j, kind) in enumerate(sched):
        b, i, j, kind = int(b), int(i), int(j), int(kind)
        if kind == DIAG0:
            continue
        if kind == PART:             # block (i, i) over the block columns k < i - 1
            deps = [(b, i, k) for k in range(i - 1)]
        else:                        # tile (i, j): both row panels left of column j, and diag(j) for the solve
            deps = [(b, i, k) for k in range(j)] + [(b, j, k) for k in range(j + 1)]
        for d in deps:
            assert final[d] < t, (t, (b, i, j, kind), d)
        if kind == FUSED and i >= 2:  # the pre-accumulated block it finishes (block 1 has nothing to pre-accumulate)
            assert partial[(b, i, i)] < t
