"""Parity of every kernel route IN THE REGIME THE BENCHMARKS RUN IT IN (through the C ABI, device memory).

The shape tests in test_gpu_parity.py mostly give each persistent CTA one request.  The benchmarked
configurations do not: cfg4 hands every CTA ~55 requests of 12 accumulation chunks each, cfg3's TF32 route
rotates its finaliser token ring, and some instantiations are reachable only through the LWC_SMALL_IMPL knob.
Here every route gets several requests per CTA (R >= 4 x 148 SMs) and, on the tcgen05 routes, several
accumulator hand-offs per request -- including an ODD chunk count, where the accumulator a request starts on
alternates from request to request.  Checker: the f64 oracle on every request (tally semantics
src/score/completions/client.rs:410-449; tolerances of SURVEY.md 8.0).
"""
import os

import numpy as np
import pytest

from parity_util import CONF_ATOL, CONF_RTOL
from test_gpu_parity import TF32_RTOL, compare, properties_hold

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def lwc():
    import lwc_b200

    lwc_b200.load_library()
    return lwc_b200


@pytest.fixture(scope="module")
def ctx(lwc):
    c = lwc.Context(device=0)
    yield c
    c.close()


def fast_batch(R, N, D, M, seed, positive=False):
    """Cheap seeded inputs (the clustered generator of synth is too slow at a gigabyte): gaussian rows with a
    lognormal scale, votes one-hot or a probability vector on <= 20 contiguous choices, some voters absent."""
    from lwc_b200 import synth

    rng = np.random.default_rng(seed)
    E = rng.standard_normal((R, N, D), dtype=np.float32)
    if positive:
        np.abs(E, out=E)
    E *= rng.lognormal(0.0, 0.5, (R, N, 1)).astype(np.float32)
    V = synth.votes(R, N, M, seed=seed)
    w = synth.weights(R, M, seed=seed)
    if R > 8:
        V[5] = np.nan          # all voters absent
        E[6, N - 1] = 0.0      # zero-norm candidate
        w[7, 0] = np.nan       # weight None
        E[8, 1] = E[8, 0]      # duplicate candidates
    return E, V, w


# (R, N, D, M): R >= 4 requests per CTA on 148 SMs.  D = 3072 -> 96 operand slabs = 12 chunks of 8 (even);
# D = 2304 -> 72 slabs = 9 chunks (odd: the starting accumulator alternates between requests); D = 2336 -> 73
# slabs: the last stage and the last chunk are partial.
TC_REGIME = [(600, 128, 3072, 16), (600, 128, 2304, 16), (600, 40, 3072, 16), (600, 100, 2304, 16), (620, 128, 2336, 5),
             (1200, 64, 1024, 3)]


@pytest.mark.parametrize("shape", TC_REGIME, ids=lambda s: "R%d_N%d_D%d_M%d" % s)
def test_tcgen05_fp32_grade_many_requests_per_cta(ctx, lwc, oracle, shape):
    R, N, D, M = shape
    assert ctx.kernel_for(R, N, D, M) == "tc128_tf32x3"
    E, V, w = fast_batch(R, N, D, M, seed=90 + N + D)
    got = ctx.score_batch_device(E, V, w)
    want = oracle.score_f64(E, V, w, oracle.SIM_COSINE)
    assert compare(got, want, V) == 1.0
    properties_hold(got, V)
    # a request's result may not depend on where it sits in the batch (which CTA, which accumulator parity)
    perm = np.random.default_rng(3).permutation(R)[:300]
    again = ctx.score_batch_device(E[perm], V[perm], w[perm])
    for k in got:
        assert np.array_equal(again[k], got[k][perm], equal_nan=True), k


@pytest.mark.parametrize("shape", TC_REGIME, ids=lambda s: "R%d_N%d_D%d_M%d" % s)
def test_tcgen05_tf32_many_requests_per_cta(ctx, lwc, oracle, shape):
    R, N, D, M = shape
    assert ctx.kernel_for(R, N, D, M, precision=lwc.PREC_TF32) == "tc128_tf32"
    E, V, w = fast_batch(R, N, D, M, seed=190 + N + D)
    got = ctx.score_batch_device(E, V, w, precision=lwc.PREC_TF32)
    want = oracle.score_f64(E, V, w, oracle.SIM_COSINE)
    compare(got, want, V, rtol=TF32_RTOL, atol=1e-5, gap=5e-3, min_agreement=0.95, atol_of_max=TF32_RTOL)
    properties_hold(got, V)


def test_tcgen05_fp32_grade_margin_in_the_benchmark_regime(ctx, lwc, oracle):
    """All-positive rows (worst case for the truncating tensor-core accumulator) with many requests per CTA:
    the fraction of the 1e-5 bar that is used must stay what the one-request-per-CTA test measured."""
    R, N, D, M = 600, 128, 3072, 8
    E, V, w = fast_batch(R, N, D, M, seed=5, positive=True)
    got = ctx.score_batch_device(E, V, w)
    want = oracle.score_f64(E, V, w, oracle.SIM_COSINE)
    used = np.abs(got["conf"] - want["conf"]) / (CONF_RTOL * np.abs(want["conf"]) + CONF_ATOL)
    print(f"fp32-grade tc route, {R} requests: tolerance used {used.max():.3f}")
    assert used.max() < 0.5
    assert (got["winner"] == want["winner"]).all()


# ------------------------------------------------------------------ small pools: every shipped instantiation

SMALL_REGIME = [(1200, 32, 3072, 4), (1500, 16, 1536, 4), (1100, 24, 1000, 7), (900, 9, 772, 3)]


@pytest.mark.parametrize("impl", ["default", "stream", "direct"])
@pytest.mark.parametrize("prec", ["fp32", "tf32"])
@pytest.mark.parametrize("shape", SMALL_REGIME, ids=lambda s: "R%d_N%d_D%d_M%d" % s)
def test_small_pool_routes_many_requests_per_cta(lwc, oracle, shape, prec, impl):
    """score_stream_kernel<16|32, x3|tf32> (token ring between finalisers rotates: R / 148 >= 6 requests per
    CTA) and score_small_kernel<16|32, *> (one CTA per request), selected the way a deployment would select
    them: LWC_SMALL_IMPL in the environment when the context is created."""
    R, N, D, M = shape
    E, V, w = fast_batch(R, N, D, M, seed=300 + N + D)
    old = os.environ.pop("LWC_SMALL_IMPL", None)
    try:
        if impl != "default":
            os.environ["LWC_SMALL_IMPL"] = impl
        with lwc.Context(device=0) as c:
            p = lwc.PREC_FP32 if prec == "fp32" else lwc.PREC_TF32
            got = c.score_batch_device(E, V, w, precision=p)
    finally:
        os.environ.pop("LWC_SMALL_IMPL", None)
        if old is not None:
            os.environ["LWC_SMALL_IMPL"] = old
    want = oracle.score_f64(E, V, w, oracle.SIM_COSINE)
    if prec == "fp32":
        assert compare(got, want, V) == 1.0
    else:
        compare(got, want, V, rtol=TF32_RTOL, atol=1e-5, gap=5e-3, min_agreement=0.95, atol_of_max=TF32_RTOL)
    properties_hold(got, V)


@pytest.mark.parametrize("kw,u", [(2, 2), (4, 4), (8, 2)])
def test_direct_kernel_knobs(lwc, oracle, kw, u):
    """LWC_SMALL_KW / LWC_SMALL_U select other instantiations of the one-CTA-per-request kernel."""
    R, N, D, M = 400, 32, 1536, 4
    E, V, w = fast_batch(R, N, D, M, seed=17)
    env = {"LWC_SMALL_IMPL": "direct", "LWC_SMALL_KW": str(kw), "LWC_SMALL_U": str(u)}
    old = {k: os.environ.get(k) for k in env}
    try:
        os.environ.update(env)
        with lwc.Context(device=0) as c:
            got = c.score_batch_device(E, V, w)
    finally:
        for k, v in old.items():
            if v is None:
                os.environ.pop(k, None)
            else:
                os.environ[k] = v
    want = oracle.score_f64(E, V, w, oracle.SIM_COSINE)
    assert compare(got, want, V) == 1.0


def test_generic_route_many_requests(ctx, lwc, oracle):
    """The fp32 FMA route (D % 4 != 0) at a batch size where every SM runs several CTAs in turn."""
    for R, N, D, M in ((700, 20, 50, 4), (500, 100, 97, 6)):
        E, V, w = fast_batch(R, N, D, M, seed=N)
        got = ctx.score_batch_device(E, V, w)
        want = oracle.score_f64(E, V, w, oracle.SIM_COSINE)
        assert compare(got, want, V) == 1.0
