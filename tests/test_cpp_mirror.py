"""The C++ host-side mirror of the reference API (include/gpb200.hpp, header-only, above the C ABI).

CPU part: the header compiles warning-free, links against libgpb200.so, and a program built on it
refuses to run without a device (no CPU fallback).  GPU part: tests/cpp/test_mirror.cpp drives every
mirrored method on the B200 and prints inputs + results; they are compared with the CPU oracle here.
"""
import json
import math
import os
import subprocess

import numpy as np
import pytest
import scipy.linalg as sl

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIBDIR = os.path.join(ROOT, "gaussian-processes_b200")
SRC = os.path.join(ROOT, "tests", "cpp", "test_mirror.cpp")
EXE = os.path.join(ROOT, "tests", "cpp", "test_mirror")
SE_W = [("se", 3.0, 5.5), ("white", 0.5)]


def build_driver():
    if not os.path.exists(os.path.join(LIBDIR, "libgpb200.so")):
        import __graft_entry__ as g
        g.build()
    cmd = ["g++", "-std=c++17", "-O1", "-Wall", "-Wextra", "-Werror", "-I", os.path.join(ROOT, "include"), SRC,
           "-L", LIBDIR, "-lgpb200", "-Wl,-rpath," + LIBDIR, "-o", EXE]
    r = subprocess.run(cmd, capture_output=True, text=True)
    assert r.returncode == 0, r.stderr
    return EXE


def test_header_compiles_links_and_has_no_cpu_fallback():
    exe = build_driver()
    import torch
    if torch.cuda.is_available():
        pytest.skip("device present: the no-device behaviour is checked on the CPU container")
    r = subprocess.run([exe], capture_output=True, text=True, timeout=120)
    assert r.returncode != 0
    assert "no CPU fallback" in r.stderr


@pytest.fixture(scope="module")
def out():
    exe = build_driver()
    r = subprocess.run([exe], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stderr[-2000:]
    return json.loads(r.stdout)


def rel(a, b):
    return abs(a - b) / max(abs(b), 1e-300)


@pytest.mark.gpu
def test_loglikelihood_and_batch(out, oracle):
    x, y = np.array(out["x1"]), np.array(out["y1"])
    o = oracle.loglik(x, y, SE_W)
    assert rel(out["ll_as_parsed"], o["ll_ref"]) < 1e-10      # north_star tolerance
    assert rel(out["ll_gpml"], o["ll_gpml"]) < 1e-10
    for b, v in enumerate(out["ll_batch_gpml"]):
        terms = [("se", 2.0 + b, 4.0 + 0.5 * b), ("white", 0.4 + 0.1 * b)]
        ob = oracle.loglik(x, y, terms, beta=(0.2 * b, -0.1 + 0.05 * b))
        assert rel(v, ob["ll_gpml"]) < 1e-10
    assert out["launches"] > 0


@pytest.mark.gpu
def test_predict_fit(out, oracle):
    x, y = np.array(out["x1"]), np.array(out["y1"])
    mo, vo, info = oracle.predict(x, y, SE_W, np.array(out["xstar1"]))
    assert info == 0
    assert np.allclose(out["fit_mean"], mo, rtol=1e-7, atol=1e-9)
    assert np.allclose(out["fit_sigma"], np.sqrt(vo), rtol=1e-7)


@pytest.mark.gpu
def test_kernel_sampler_follows_the_oracle_chain(out, oracle):
    """Same streams -> the same accept / reject sequence and the same states (Breeze's metropolis
    re-evaluates ll(current) every step: evals_per_iter=2)."""
    x, y = np.array(out["x1"]), np.array(out["y1"])
    z = np.array(out["mh_z"]).reshape(25, 3)
    u = np.array(out["mh_u"])
    ig = np.array([[3.0, 6.0], [3.0, 10.0], [3.0, 1.0]])
    states, _, acc, rc = oracle.mh_run(x, y, SE_W, ig, 0.05, z, u, evals_per_iter=2)
    assert rc == 0 and 0 < acc.sum() < 25
    got = np.array(out["mh_states"]).reshape(25, 3)
    moved = np.any(np.diff(np.vstack([[3.0, 5.5, 0.5], got]), axis=0) != 0, axis=1)
    assert np.array_equal(moved.astype(np.int32), acc)
    assert np.allclose(got, states, rtol=1e-14, atol=0)


@pytest.mark.gpu
def test_matrix_level_methods(out, oracle):
    x = np.stack([out["xa"], out["xb"]], axis=1)
    n = len(x)
    y, z = np.array(out["y2"]), np.array(out["z2"])
    D = np.array(out["D"]).reshape(n, n).T
    assert np.array_equal(D, np.sqrt((x[:, None, 0] - x[None, :, 0]) ** 2 + (x[:, None, 1] - x[None, :, 1]) ** 2))
    K = np.array(out["K"]).reshape(n, n).T
    Ko = oracle.build_cov(x, SE_W, 0.0)
    assert np.allclose(K, Ko, rtol=1e-15, atol=1e-15)
    L = np.array(out["L"]).reshape(n, n).T
    Lo, info = oracle.potrf_lower(Ko + 1e-3 * np.eye(n))
    assert info == 0 and np.allclose(L, Lo, rtol=1e-10, atol=1e-12) and np.all(np.triu(L, 1) == 0)
    fwd = sl.solve_triangular(Lo, y, lower=True)
    assert np.allclose(out["fwd"], fwd, rtol=1e-10, atol=1e-12)
    assert np.allclose(out["back"], sl.solve_triangular(Lo.T, fwd), rtol=1e-9, atol=1e-11)
    B2 = np.array(out["back2"]).reshape(2, n).T
    assert np.allclose(B2, sl.solve_triangular(Lo.T, np.stack([y, z], axis=1)), rtol=1e-9, atol=1e-11)


@pytest.mark.gpu
def test_sample_plane_and_draw(out, oracle):
    x = np.stack([out["xa"], out["xb"]], axis=1)
    n = len(x)
    X = np.array(out["design"]).reshape(3, n).T
    assert np.array_equal(X, oracle.design_matrix(x))          # the reference's interleaved layout
    beta_o, _, _ = oracle.sample_plane(x, np.array(out["y2"]), SE_W, 1.0, 4.0, np.array(out["zb"]), nugget=1e-6)
    assert np.allclose(out["plane_beta"], beta_o, rtol=1e-8, atol=1e-10)
    _, chol_draw = oracle.draw(x, SE_W, np.array(out["z2"]), beta=(0.5, 0.1, -0.2), nugget=1e-3)
    assert np.allclose(out["draw"], chol_draw, rtol=1e-9, atol=1e-11)


@pytest.mark.gpu
def test_mll_gradient(out, oracle):
    x = np.stack([out["xa"], out["xb"]], axis=1)
    y = np.array(out["y2"])
    go = oracle.mll_gradient(x, y, [("se", 1.1, 0.9), ("white", 0.3)], mode="as_executed")
    assert np.allclose(out["grad_as_executed"], go, rtol=1e-8, atol=1e-8 * np.abs(go).max())
    th = [("se", math.log(3.0), math.log(5.5)), ("white", math.log(0.5))]
    ga = oracle.mll_gradient(x, y, th, beta=(0.5, 0.1, -0.2), mode="analytic")
    assert np.allclose(out["grad_analytic"], ga, rtol=1e-8, atol=1e-8 * np.abs(ga).max())


@pytest.mark.gpu
def test_error_behaviour(out):
    assert out["err_mixed_dims"] == "Dimension mismatch"       # Location.scala:32
    assert out["err_not_pd"] == "NotConvergedException"        # breeze cholesky
    assert out["err_bad_proposal"] == "NotConvergedException"
    assert out["err_matern"] == "invalid_argument"             # out of scope, refused loudly
    assert out["err_dist"] == "invalid_argument"
    assert not math.isnan(out["ll_as_parsed"])
