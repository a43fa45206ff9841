"""gpb_loglik_noalloc (include/gpb200.h): the log-likelihood evaluation whose workspace is a static
__device__ array inside the library -- no cudaMalloc / cudaFree in the call.  It has no counterpart in
the reference; the check is against the existing entry
point gpb_loglik on the same inputs.  Both run the same kernels in the same order, so the three
reductions must agree bit for bit; the asserted bound is the suite's likelihood tolerance (relative
1e-10).  Collected last on purpose."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

TERMS = [(0, 3.0, 5.5), (1, 0.5, 0.0)]


def _problem(n, dim, seed):
    rng = np.random.default_rng(seed)
    x = rng.uniform(-10, 10, (n, dim))
    y = 1.8 * rng.standard_normal(n)
    return x, y


def _both(G, ctx, torch, n, dim, terms, beta, seed):
    x, y = _problem(n, dim, seed)
    ds = ctx.dataset(x if dim > 1 else x[:, 0], y)
    ref = ds.loglik_batch([terms], None if beta is None else [beta], nugget=1e-3)[0]
    ref = (ref.Q, ref.U1, ref.S, ref.n, ref.info)
    ds.close()
    # the caller's own device buffers (the entry point itself allocates nothing)
    dx = torch.from_numpy(np.ascontiguousarray(x.T)).to("cuda:0")
    dy = torch.from_numpy(y).to("cuda:0")
    torch.cuda.synchronize()
    st, got = G.loglik_noalloc(0, torch.cuda.current_stream().cuda_stream, n, dim, dx.data_ptr(), dy.data_ptr(),
                               terms, beta, nugget=1e-3)
    assert st == 0, st
    return ref, (got.Q, got.U1, got.S, got.n, got.info)


def test_noalloc_matches_loglik_two_calls_in_a_row(ctx):
    import torch
    from gpb200 import _lib as G
    # second call: smaller system, other kernel structure, plane mean -- it finds the first call's
    # contents in the shared scratch and must not see them
    cases = [(G.NOALLOC_MAX_N, 2, TERMS, None, 11),
             (203, 1, [(0, 1.3, 0.7), (0, 0.4, 2.5), (1, 0.2, 0.0)], [0.3, -0.1], 12),
             (G.NOALLOC_MAX_N - 1, 2, TERMS, [0.1, 0.2, -0.3], 13)]
    for n, dim, terms, beta, seed in cases:
        ref, got = _both(G, ctx, torch, n, dim, terms, beta, seed)
        assert ref[3] == got[3] == n and ref[4] == got[4] == 0
        for a, b in zip(ref[:3], got[:3]):
            assert abs(a - b) <= 1e-10 * abs(a), (n, a, b)
        assert ref[:3] == got[:3], "same kernels, same order: expected bit-identical reductions"


def test_noalloc_refuses_sizes_beyond_the_static_scratch(ctx):
    import torch
    from gpb200 import _lib as G
    n = G.NOALLOC_MAX_N + 1
    dx = torch.zeros(2 * n, dtype=torch.float64, device="cuda:0")
    dy = torch.zeros(n, dtype=torch.float64, device="cuda:0")
    st, _ = G.loglik_noalloc(0, 0, n, 2, dx.data_ptr(), dy.data_ptr(), TERMS)
    assert st == -1  # GPB_E_INVALID, message on stderr
    # not positive definite: reported in info like gpb_loglik, not as a status
    st, r = G.loglik_noalloc(0, 0, 64, 2, dx.data_ptr(), dy.data_ptr(), [(0, 1.0, 1.0), (1, -5.0, 0.0)], nugget=0.0)
    assert st == 0 and r.info > 0
