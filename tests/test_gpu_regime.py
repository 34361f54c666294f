"""Parity of every kernel route IN THE REGIME THE BENCHMARKS RUN IT IN (through the C ABI, device memory).

The shape tests in test_gpu_parity.py mostly give each persistent CTA one request.  The benchmarked
configurations do not: cfg4 hands every CTA ~55 requests of 12 accumulation chunks each, cfg3's TF32 route
rotates its finaliser token ring, and some instantiations are reachable only through the LWC_SMALL_IMPL knob.
Here every route gets several requests per CTA (R >= 4 x 148 SMs) and, on the tcgen05 routes, several
accumulator hand-offs per request -- including an ODD chunk count, where the accumulator a request starts on
alternates from request to request.  Checker: the f64 oracle on every request (tally semantics
src/score/completions/client.rs:410-449; tolerances of SURVEY.md 8.0).
"""
import contextlib
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


@contextlib.contextmanager
def env(**kv):
    """Environment knobs are read once per context: set them around the lwc.Context(...) they are meant for."""
    old = {k: os.environ.get(k) for k in kv}
    try:
        for k, v in kv.items():
            if v is None:
                os.environ.pop(k, None)
            else:
                os.environ[k] = v
        yield
    finally:
        for k, v in old.items():
            if v is None:
                os.environ.pop(k, None)
            else:
                os.environ[k] = v


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
        if N > 1:
            E[8, 1] = E[8, 0]  # duplicate candidates
    return E, V, w


# (R, N, D, M): R >= 4 requests per CTA on 148 SMs.  D = 3072 -> 96 operand slabs = 12 chunks of 8 (even);
# D = 2304 -> 72 slabs = 9 chunks (odd: the starting accumulator alternates between requests); D = 2336 -> 73
# slabs: the last stage and the last chunk are partial.
TC_REGIME = [(600, 128, 3072, 16), (600, 128, 2304, 16), (600, 40, 3072, 16), (600, 100, 2304, 16), (620, 128, 2336, 5),
             (1200, 64, 1024, 3)]


@pytest.mark.parametrize("shape", TC_REGIME, ids=lambda s: "R%d_N%d_D%d_M%d" % s)
def test_tcgen05_fp32_grade_many_requests_per_cta(ctx, lwc, oracle, shape):
    R, N, D, M = shape
    assert ctx.kernel_for(R, N, D, M) == ("tc64x2_f16x2" if N <= 64 else "tc128_f16x2")
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
    assert ctx.kernel_for(R, N, D, M, precision=lwc.PREC_TF32) == ("tc64x2_tf32" if N <= 64 else "tc128_tf32")
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


def test_batch_list_entry_is_bit_equal_to_single_calls(lwc, oracle):
    """lwc_score_batches enqueues a list of device-resident batches inside one call; the results must be the
    bits of the same batches scored one lwc_score_batch call at a time -- with and without overlapped launches."""
    shapes = [(700, 16, 1536, 4), (300, 32, 1024, 4), (40, 128, 512, 8), (0, 16, 64, 2), (90, 10, 50, 3), (700, 16, 1536, 4)]
    batches = [fast_batch(max(R, 1), N, D, M, seed=50 + i)[:3] for i, (R, N, D, M) in enumerate(shapes)]
    for flags in (0, lwc.CTX_INPUTS_RESIDENT):
        with lwc.Context(device=0, flags=flags) as c:
            singles, keep, descs, ptrs = [], [], [], {k: [] for k in ("E", "V", "w", "conf", "cw", "agree", "winner", "rank")}
            for (R, N, D, M), (E, V, w) in zip(shapes, batches):
                E, V, w = E[:R], V[:R], w[:R]
                singles.append(c.score_batch_device(E, V, w) if R else None)
                bufs = {"E": lwc.DeviceBuffer(c, max(E.nbytes, 4)).upload(E), "V": lwc.DeviceBuffer(c, max(V.nbytes, 4)).upload(V),
                        "w": lwc.DeviceBuffer(c, max(w.nbytes, 4)).upload(w), "conf": lwc.DeviceBuffer(c, 4 * R * N + 4),
                        "cw": lwc.DeviceBuffer(c, 4 * R * N + 4), "agree": lwc.DeviceBuffer(c, 4 * R * M + 4),
                        "winner": lwc.DeviceBuffer(c, 4 * R + 4), "rank": lwc.DeviceBuffer(c, 4 * R * N + 4)}
                keep.append(bufs)
                descs.append(lwc.BatchDesc(R, N, D, M, lwc.SIM_COSINE, lwc.PREC_FP32, lwc.MEM_DEVICE, 1))
                for k in ptrs:
                    ptrs[k].append(bufs[k].ptr)
            run = c.batch_list(descs, ptrs["E"], ptrs["V"], ptrs["w"], ptrs["conf"], ptrs["cw"], ptrs["agree"], ptrs["winner"],
                               ptrs["rank"])
            launches0 = c.launch_count()
            run()
            c.synchronize()
            # one launch per non-empty batch; the fp32-grade tcgen05 route is two (fp16-piece pass + its fix-up pass)
            assert c.launch_count() - launches0 == sum((2 if s[1] > 32 else 1) for s in shapes if s[0] > 0)
            for (R, N, D, M), bufs, want in zip(shapes, keep, singles):
                if not R:
                    continue
                got = {"conf": bufs["conf"].download((R, N), np.float32), "cw": bufs["cw"].download((R, N), np.float32),
                       "agree": bufs["agree"].download((R, M), np.float32), "winner": bufs["winner"].download((R,), np.int32),
                       "rank": bufs["rank"].download((R, N), np.int32)}
                for k in got:
                    assert np.array_equal(got[k], want[k], equal_nan=True), (k, (R, N, D, M), flags)
            # a list with one bad entry enqueues nothing
            bad = list(descs)
            bad[2] = lwc.BatchDesc(5, 300, 64, 2, lwc.SIM_COSINE, lwc.PREC_FP32, lwc.MEM_DEVICE, 1)
            launches0 = c.launch_count()
            with pytest.raises(lwc.LwcError) as ei:
                c.batch_list(bad, ptrs["E"], ptrs["V"], ptrs["w"], ptrs["conf"], ptrs["cw"], ptrs["agree"], ptrs["winner"], ptrs["rank"])()
            assert ei.value.code == 400 and c.launch_count() == launches0
            for bufs in keep:
                for b in bufs.values():
                    b.free()


# ------------------------------------------------------------------ element types on the way in

@pytest.mark.parametrize("memory", ["host", "device"])
@pytest.mark.parametrize("shape", [(700, 16, 1536, 4), (300, 32, 1000, 3), (40, 128, 512, 8), (33, 20, 50, 2), (5, 3, 7, 1)],
                         ids=lambda s: "R%d_N%d_D%d_M%d" % s)
def test_f64_transport_gives_the_bits_of_host_conversion(ctx, lwc, oracle, shape, memory):
    """Embedding.embedding is Vec<f64> (embeddings/response.rs:21).  LWC_ELEM_F64 sends the doubles and rounds on the
    device; the result must be bit-identical to rounding on the host (lwc_stage_f64) and calling lwc_score_batch."""
    R, N, D, M = shape
    rng = np.random.default_rng(R + D)
    E64 = rng.standard_normal((R, N, D)) * rng.lognormal(0, 0.5, (R, N, 1))
    from lwc_b200 import synth

    V, w = synth.votes(R, N, M, seed=R), synth.weights(R, M, seed=R)
    E32 = np.empty((R, N, D), np.float32)
    ctx.stage_f64(E64, E32)
    assert np.array_equal(E32, E64.astype(np.float32))
    want = ctx.score_batch(E32, V, w)
    got = ctx.score_batch_typed(E64, lwc.ELEM_F64, V, w, memory=lwc.MEM_HOST if memory == "host" else lwc.MEM_DEVICE)
    for k in want:
        assert np.array_equal(got[k], want[k], equal_nan=True), k
    ref = oracle.score_f64(E32, V, w, oracle.SIM_COSINE)
    assert compare(got, ref, V) == 1.0


BF16_TOL = 1e-2  # stated tolerance of the bf16 transport: 1e-2 of a request's top confidence


@pytest.mark.parametrize("memory", ["host", "device"])
@pytest.mark.parametrize("shape", [(700, 16, 1536, 4), (300, 32, 3072, 4), (200, 128, 1024, 16), (33, 20, 50, 2)],
                         ids=lambda s: "R%d_N%d_D%d_M%d" % s)
def test_bf16_transport_within_its_stated_tolerance(ctx, lwc, oracle, shape, memory):
    """LWC_ELEM_BF16: 2 bytes per element cross PCIe.  (a) The device side is exact: it must reproduce, within the
    fp32 bar, the oracle run on the bf16-rounded embeddings.  (b) Against the oracle on the ORIGINAL embeddings the
    stated 1e-2-of-top tolerance must hold and the winner must agree wherever the margin exceeds it."""
    from lwc_b200 import synth

    R, N, D, M = shape
    E, V, w, _ = synth.make_batch(R, N, D, M, seed=40 + N, edge_cases=False)
    bits = ctx.stage_f64_bf16(E.astype(np.float64))
    # independent rounding model: nearest even on the upper 16 bits
    u = E.view(np.uint32).astype(np.uint64)
    model = ((u + 0x7FFF + ((u >> 16) & 1)) >> 16).astype(np.uint16)
    assert np.array_equal(bits, model)
    got = ctx.score_batch_typed(bits, lwc.ELEM_BF16, V, w, memory=lwc.MEM_HOST if memory == "host" else lwc.MEM_DEVICE)
    E_rounded = (bits.astype(np.uint32) << 16).view(np.float32)
    exact = oracle.score_f64(E_rounded, V, w, oracle.SIM_COSINE)
    assert compare(got, exact, V) == 1.0                       # (a)
    ref = oracle.score_f64(E, V, w, oracle.SIM_COSINE)         # (b)
    top = np.abs(ref["conf"]).max(axis=1, keepdims=True)
    assert np.all(np.abs(got["conf"] - ref["conf"]) <= BF16_TOL * top)
    srt = np.sort(ref["conf"], axis=1)[:, ::-1]
    clear = (ref["winner"] >= 0) & (srt[:, 0] - srt[:, 1] > 2 * BF16_TOL * top[:, 0])
    assert np.array_equal(got["winner"][clear], ref["winner"][clear])
    print(f"bf16 transport {shape}: max conf err / top = {float(np.max(np.abs(got['conf'] - ref['conf']) / np.maximum(top, 1e-30))):.2e}, "
          f"winner agreement {float(np.mean(got['winner'] == ref['winner'])):.4f}")


def test_typed_entry_argument_errors(ctx, lwc):
    import ctypes

    d = lwc.BatchDesc(1, 4, 8, 1, lwc.SIM_COSINE, lwc.PREC_FP32, lwc.MEM_HOST, 1)
    one = ctypes.c_void_p(16)
    rc = ctx.lib.lwc_score_batch_typed(ctx.handle, ctypes.byref(d), 7, one, one, one, None, None, None, None, None, None)
    assert rc == 400 and "element type" in ctx.lib.lwc_last_error_message(ctx.handle).decode()
    rc = ctx.lib.lwc_score_batch_typed(ctx.handle, ctypes.byref(d), lwc.ELEM_F64, ctypes.c_void_p(8), one, one, None, None, None, None,
                                       None, None)
    assert rc == 400 and "16-byte aligned" in ctx.lib.lwc_last_error_message(ctx.handle).decode()


# ------------------------------------------------------------------ fp16-piece tcgen05 route: range check + fix-up pass

def test_fp16_piece_route_fixes_up_requests_outside_fp16_range(ctx, lwc, oracle):
    """The default fp32-grade route for 32 < N <= 128 splits every element into two fp16 pieces.  fp16 has 5
    exponent bits, so the kernel marks requests whose candidates' norms it cannot represent exactly enough (tiny,
    huge, zero, non-finite squared norm) and a second pass re-scores exactly those on tf32 pieces.  Here marked and
    unmarked requests are interleaved inside every CTA's range; all of them must meet the fp32 bar."""
    R, N, D, M = 640, 128, 1024, 8
    E, V, w = fast_batch(R, N, D, M, seed=77)
    scale = np.ones((R, 1, 1), np.float32)
    scale[3::7] = 1e-6       # every candidate tiny: h would be subnormal or zero
    scale[4::7] = 3e4        # elements far beyond 65504
    scale[5::7] = 1e-3       # norms just under the kernel's threshold
    E *= scale
    E[6::7, 10] *= np.float32(1e-7)   # ONE tiny candidate among ordinary ones
    E[9::14, 20, 5] = 7e4             # ONE out-of-range element
    got = ctx.score_batch_device(E, V, w)
    want = oracle.score_f64(E, V, w, oracle.SIM_COSINE)
    assert compare(got, want, V) == 1.0
    properties_hold(got, V)
    # position independence holds across the two passes too
    perm = np.random.default_rng(5).permutation(R)[:300]
    again = ctx.score_batch_device(E[perm], V[perm], w[perm])
    for k in got:
        assert np.array_equal(again[k], got[k][perm], equal_nan=True), k


PAIR_SHAPES = [(601, 64, 1024, 8), (333, 33, 2304, 3), (1, 48, 256, 4), (2, 64, 96, 16), (299, 57, 2336, 5), (1201, 40, 512, 0)]


@pytest.mark.parametrize("shape", PAIR_SHAPES, ids=lambda s: "R%d_N%d_D%d_M%d" % s)
def test_two_requests_per_tile_route(lwc, oracle, shape):
    """Pools of 33..64 candidates at the default precision ride two to a 128-row tcgen05 tile (score_tc16_kernel<PAIR>):
    each request must meet the fp32 bar, must not see its tile partner (the same request next to a different partner, or
    alone at the end of an odd batch, gives the same bits), and the range check / fix-up pass works per request."""
    R, N, D, M = shape
    E, V, w = fast_batch(R, N, D, max(M, 1), seed=300 + N + D)
    if M == 0:
        V, w = V[:, :0], w[:, :0]
    if R > 8:
        E[3] *= np.float32(1e-6)     # marked: fix-up pass rewrites it; its partner (request 2) stays on the fast pass
        E[6, N - 1, :] = 0.0         # a zero-norm candidate marks request 6
        E[R - 1, 0, 1] = 7e4         # the odd batch's last request, alone in its tile, out of fp16 range
    with lwc.Context(device=0) as c:
        assert c.kernel_for(R, N, D, M) == "tc64x2_f16x2"
        got = c.score_batch_device(E, V, w)
        again = c.score_batch_device(E, V, w)
        want = oracle.score_f64(E, V, w, oracle.SIM_COSINE)
        assert compare(got, want, V) == 1.0
        properties_hold(got, V)
        assert all(np.array_equal(got[k], again[k], equal_nan=True) for k in got)
        if R > 8:
            perm = np.random.default_rng(8).permutation(R)[: R - 1 if R % 2 else R - 3]  # other partners, other halves
            moved = c.score_batch_device(E[perm], V[perm], w[perm])
            for k in got:
                assert np.array_equal(moved[k], got[k][perm], equal_nan=True), k
    old = os.environ.get("LWC_TC_PAIR")
    try:
        os.environ["LWC_TC_PAIR"] = "0"  # one request per tile: same arithmetic on the same accumulator entries
        with lwc.Context(device=0) as c:
            assert c.kernel_for(R, N, D, M) == "tc128_f16x2"
            single = c.score_batch_device(E, V, w)
    finally:
        if old is None:
            os.environ.pop("LWC_TC_PAIR", None)
        else:
            os.environ["LWC_TC_PAIR"] = old
    assert compare(single, want, V) == 1.0
    assert np.array_equal(single["winner"], got["winner"])
    np.testing.assert_allclose(single["conf"], got["conf"], rtol=2e-6, atol=1e-9)


@pytest.mark.parametrize("shape", PAIR_SHAPES, ids=lambda s: "R%d_N%d_D%d_M%d" % s)
def test_two_requests_per_tile_route_tf32(lwc, oracle, shape):
    """The same pairing on the single-pass TF32 route (score_tc_kernel<false, PAIR>): stated TF32 tolerance against the
    oracle, and bit-equal to the one-request-per-tile kernel -- every accumulator entry a request reads is produced by
    the same MMAs on the same operands, and its reductions add the same terms in the same order."""
    R, N, D, M = shape
    E, V, w = fast_batch(R, N, D, max(M, 1), seed=400 + N + D)
    if M == 0:
        V, w = V[:, :0], w[:, :0]
    with lwc.Context(device=0) as c:
        assert c.kernel_for(R, N, D, M, precision=lwc.PREC_TF32) == "tc64x2_tf32"
        got = c.score_batch_device(E, V, w, precision=lwc.PREC_TF32)
    want = oracle.score_f64(E, V, w, oracle.SIM_COSINE)
    compare(got, want, V, rtol=2e-3, atol=1e-5, gap=5e-3, min_agreement=0.95, atol_of_max=2e-3)
    properties_hold(got, V)
    old = os.environ.get("LWC_TC_PAIR")
    try:
        os.environ["LWC_TC_PAIR"] = "0"
        with lwc.Context(device=0) as c:
            assert c.kernel_for(R, N, D, M, precision=lwc.PREC_TF32) == "tc128_tf32"
            single = c.score_batch_device(E, V, w, precision=lwc.PREC_TF32)
    finally:
        if old is None:
            os.environ.pop("LWC_TC_PAIR", None)
        else:
            os.environ["LWC_TC_PAIR"] = old
    for k in got:
        assert np.array_equal(single[k], got[k], equal_nan=True), k


@pytest.mark.parametrize("shape", [(600, 128, 3072, 16), (600, 100, 2304, 16), (300, 40, 96, 3)], ids=lambda s: "R%d_N%d_D%d_M%d" % s)
def test_tf32_piece_route_remains_selectable(lwc, oracle, shape):
    """LWC_TC_X3=tf32: the fp32-grade route on tf32 pieces alone (round 1's kernel, now also the fix-up pass)."""
    from lwc_b200 import synth

    R, N, D, M = shape
    E, V, w, _ = synth.make_batch(R, N, D, M, seed=123 + N)
    old = os.environ.get("LWC_TC_X3")
    try:
        os.environ["LWC_TC_X3"] = "tf32"
        with lwc.Context(device=0) as c:
            assert c.kernel_for(R, N, D, M) == "tc128_tf32x3"
            got = c.score_batch_device(E, V, w)
    finally:
        if old is None:
            os.environ.pop("LWC_TC_X3", None)
        else:
            os.environ["LWC_TC_X3"] = old
    want = oracle.score_f64(E, V, w, oracle.SIM_COSINE)
    assert compare(got, want, V) == 1.0


def test_small_pool_fp16_pieces_fall_back_outside_fp16_range(lwc, oracle):
    """The one-CTA-per-request fp32-grade kernel (the cfg3 route) runs on fp16 pieces and re-scores a request on tf32
    pieces when a candidate's squared norm shows that fp16's exponent range was left.  Ordinary, tiny, huge and
    mixed requests interleaved; LWC_SMALL_X3=tf32 (tf32 pieces only) must give results within the same bar."""
    R, N, D, M = 900, 32, 1536, 4
    E, V, w = fast_batch(R, N, D, M, seed=61)
    scale = np.ones((R, 1, 1), np.float32)
    scale[2::5] = 1e-6
    scale[3::5] = 5e4
    E *= scale
    E[4::10, 7] *= np.float32(1e-8)   # one tiny candidate among ordinary ones
    E[9::10, 3, 100] = 1e5            # one out-of-range element
    want = oracle.score_f64(E, V, w, oracle.SIM_COSINE)
    for env in (None, "tf32"):
        old = os.environ.pop("LWC_SMALL_X3", None)
        try:
            if env:
                os.environ["LWC_SMALL_X3"] = env
            with lwc.Context(device=0) as c:
                assert c.kernel_for(R, N, D, M) == ("small32_tf32x3" if env else "small32_f16x2")
                got = c.score_batch_device(E, V, w)
        finally:
            os.environ.pop("LWC_SMALL_X3", None)
            if old is not None:
                os.environ["LWC_SMALL_X3"] = old
        assert compare(got, want, V) == 1.0, env
        properties_hold(got, V)


@pytest.mark.parametrize("shape", [(700, 8, 1536, 4), (300, 5, 100, 3), (1, 1, 16, 1), (900, 7, 3072, 0), (450, 8, 772, 9)],
                         ids=lambda s: "R%d_N%d_D%d_M%d" % s)
@pytest.mark.parametrize("prec", ["fp32", "tf32"])
def test_pools_of_at_most_eight_take_the_half_tile_form(lwc, oracle, shape, prec):
    """N <= 8 on the ring kernel: rows 8..15 of the 16-row tile are neither loaded nor multiplied and the second column
    tile is never computed (score_stream.cuh, NB = 1).  Same MMAs on the same operands for everything a request reads:
    bit-equal to the full-tile form (LWC_STREAM_HALF=0), and within the route's bar against the oracle."""
    R, N, D, M = shape
    p = lwc.PREC_FP32 if prec == "fp32" else lwc.PREC_TF32
    E, V, w = fast_batch(R, N, D, max(M, 1), seed=800 + N + D)
    if M == 0:
        V, w = V[:, :0], w[:, :0]
    res = {}
    # the half-tile form streams wider stages by default (6 KB row copies), which deals the columns to the consumer
    # warps differently; at one stage width the two forms add the same terms in the same order
    for half, kc in (("1", "768"), ("0", "768"), ("1", None)):
        old = {k: os.environ.get(k) for k in ("LWC_STREAM_HALF", "LWC_STREAM_KC", "LWC_SMALL_IMPL")}
        os.environ["LWC_STREAM_HALF"] = half
        os.environ["LWC_SMALL_IMPL"] = "stream"  # requests under 56 KB go to the one-CTA-per-request kernel by default
        if kc:
            os.environ["LWC_STREAM_KC"] = kc
        try:
            with lwc.Context(device=0) as c:
                assert c.kernel_for(R, N, D, M, precision=p).startswith("small16_")
                res[(half, kc)] = c.score_batch_device(E, V, w, precision=p)
        finally:
            for k, v in old.items():
                if v is None:
                    os.environ.pop(k, None)
                else:
                    os.environ[k] = v
    for k in res[("1", "768")]:
        assert np.array_equal(res[("1", "768")][k], res[("0", "768")][k], equal_nan=True), k
    want = oracle.score_f64(E, V, w, oracle.SIM_COSINE)
    for key in (("1", "768"), ("1", None)):
        if prec == "fp32":
            assert compare(res[key], want, V) == 1.0
        else:
            compare(res[key], want, V, rtol=TF32_RTOL, atol=1e-5, gap=5e-3, min_agreement=0.95, atol_of_max=TF32_RTOL)
        properties_hold(res[key], V)


@pytest.mark.parametrize("shape", [(700, 24, 3072, 4), (300, 17, 100, 3), (1, 20, 16, 1), (500, 23, 1536, 0), (450, 24, 772, 9)],
                         ids=lambda s: "R%d_N%d_D%d_M%d" % s)
@pytest.mark.parametrize("prec", ["fp32", "tf32"])
def test_pools_of_17_to_24_skip_the_fourth_row_group(lwc, oracle, shape, prec):
    """17..24 candidates on the 32-row kernels: the fourth 8-row group is neither loaded nor multiplied (NB = 3 in
    score_small.cuh's pipelined pass and in the ring kernel).  Bit-equal to the full-tile form (LWC_STREAM_HALF=0)."""
    R, N, D, M = shape
    p = lwc.PREC_FP32 if prec == "fp32" else lwc.PREC_TF32
    E, V, w = fast_batch(R, N, D, max(M, 1), seed=900 + N + D)
    if M == 0:
        V, w = V[:, :0], w[:, :0]
    res = {}
    for half in ("1", "0"):
        old = os.environ.get("LWC_STREAM_HALF")
        os.environ["LWC_STREAM_HALF"] = half
        try:
            with lwc.Context(device=0) as c:
                assert c.kernel_for(R, N, D, M, precision=p).startswith("small32_")
                res[half] = c.score_batch_device(E, V, w, precision=p)
        finally:
            if old is None:
                os.environ.pop("LWC_STREAM_HALF", None)
            else:
                os.environ["LWC_STREAM_HALF"] = old
    for k in res["1"]:
        assert np.array_equal(res["1"][k], res["0"][k], equal_nan=True), k
    want = oracle.score_f64(E, V, w, oracle.SIM_COSINE)
    if prec == "fp32":
        assert compare(res["1"], want, V) == 1.0
    else:
        compare(res["1"], want, V, rtol=TF32_RTOL, atol=1e-5, gap=5e-3, min_agreement=0.95, atol_of_max=TF32_RTOL)
    properties_hold(res["1"], V)


@pytest.mark.parametrize("shape,prec", [((600, 16, 1536, 4), "fp32"), ((450, 32, 1024, 4), "tf32"), ((300, 9, 772, 3), "fp32"), ((500, 6, 1536, 4), "fp32"),
                                        ((400, 22, 1024, 4), "tf32"), ((600, 4, 1536, 4), "fp32"), ((501, 3, 1536, 4), "fp32")],
                         ids=["cfg2_shape", "n32_tf32", "ragged", "half_tile", "three_groups", "grouped", "grouped_odd"])
def test_a_run_of_same_shaped_batches_is_one_launch_with_the_same_bits(lwc, oracle, shape, prec):
    """lwc_score_batches turns a run of same-shaped ring-kernel batches into ONE launch whose persistent CTAs walk
    the whole run (request g = request g % R of batch g / R).  Launch count, bit-equality with single calls, and
    LWC_LIST_FUSION=0 (batch-by-batch launches) as the A/B."""
    R, N, D, M = shape
    K = 5
    p = lwc.PREC_FP32 if prec == "fp32" else lwc.PREC_TF32
    batches = [fast_batch(R, N, D, M, seed=700 + i) for i in range(K)]
    for fusion in ("1", "0"):
        old = os.environ.get("LWC_LIST_FUSION")
        os.environ["LWC_LIST_FUSION"] = fusion
        try:
            # (the ring kernel also for the two shapes under 56 KB per request, which default to one CTA per request)
            with env(LWC_SMALL_IMPL="stream"), lwc.Context(device=0, flags=lwc.CTX_INPUTS_RESIDENT) as c:
                singles = [c.score_batch_device(E, V, w, precision=p) for E, V, w in batches]
                keep, ptrs = [], {k: [] for k in ("E", "V", "w", "conf", "cw", "agree", "winner", "rank")}
                for E, V, w in batches:
                    bufs = {"E": lwc.DeviceBuffer(c, E.nbytes).upload(E), "V": lwc.DeviceBuffer(c, V.nbytes).upload(V),
                            "w": lwc.DeviceBuffer(c, w.nbytes).upload(w), "conf": lwc.DeviceBuffer(c, 4 * R * N),
                            "cw": lwc.DeviceBuffer(c, 4 * R * N), "agree": lwc.DeviceBuffer(c, 4 * R * M),
                            "winner": lwc.DeviceBuffer(c, 4 * R), "rank": lwc.DeviceBuffer(c, 4 * R * N)}
                    keep.append(bufs)
                    for k in ptrs:
                        ptrs[k].append(bufs[k].ptr)
                descs = [lwc.BatchDesc(R, N, D, M, lwc.SIM_COSINE, p, lwc.MEM_DEVICE, 1)] * K
                run = c.batch_list(descs, *[ptrs[k] for k in ("E", "V", "w", "conf", "cw", "agree", "winner", "rank")])
                n0 = c.launch_count()
                run()
                run()  # twice: the second list overlaps the first (programmatic dependent launch) and must win
                c.synchronize()
                assert c.launch_count() - n0 == (2 if fusion == "1" else 2 * K)
                for bufs, want in zip(keep, singles):
                    got = {"conf": bufs["conf"].download((R, N), np.float32), "cw": bufs["cw"].download((R, N), np.float32),
                           "agree": bufs["agree"].download((R, M), np.float32), "winner": bufs["winner"].download((R,), np.int32),
                           "rank": bufs["rank"].download((R, N), np.int32)}
                    for k in got:
                        assert np.array_equal(got[k], want[k], equal_nan=True), (k, fusion)
                for bufs in keep:
                    for b in bufs.values():
                        b.free()
        finally:
            if old is None:
                os.environ.pop("LWC_LIST_FUSION", None)
            else:
                os.environ["LWC_LIST_FUSION"] = old
    want = oracle.score_f64(*batches[0], oracle.SIM_COSINE)
    if prec == "fp32":
        assert compare(singles[0], want, batches[0][1]) == 1.0


@pytest.mark.parametrize("stages", ["0", "3", "4"])
def test_cfg3_kernel_load_pipeline_variants(lwc, oracle, stages):
    """LWC_SMALL_PIPE selects how the 32-candidate fp32-grade kernel (the cfg3 route) feeds its fp16-piece Gram pass:
    0 = loads into registers, 3 (default) / 4 = cp.async through that many lane-private shared-memory stages.  Same
    arithmetic in the same order, so the three must agree bit for bit; D values exercise rows shorter than the
    pipeline, ragged tails and several accumulator flushes."""
    results = []
    for D in (3072, 48, 1000, 4400):
        R, N, M = 500, 32, 4
        E, V, w = fast_batch(R, N, D, M, seed=900 + D)
        old = os.environ.get("LWC_SMALL_PIPE")
        try:
            os.environ["LWC_SMALL_PIPE"] = stages
            with lwc.Context(device=0) as c:
                got = c.score_batch_device(E, V, w)
            os.environ["LWC_SMALL_PIPE"] = "0"
            with lwc.Context(device=0) as c:
                base = c.score_batch_device(E, V, w)
        finally:
            if old is None:
                os.environ.pop("LWC_SMALL_PIPE", None)
            else:
                os.environ["LWC_SMALL_PIPE"] = old
        want = oracle.score_f64(E, V, w, oracle.SIM_COSINE)
        assert compare(got, want, V) == 1.0
        for k in got:
            assert np.array_equal(got[k], base[k], equal_nan=True), (k, D)


# ------------------------------------------------------------------ every route boundary, seeded

def _boundary_shapes():
    """Pool sizes on both sides of every routing threshold (8 | 9, 16 | 17, 24 | 25, 32 | 33, 64 | 65, 128) crossed with
    embedding widths that are / are not multiples of 64, 32, 16 and 4, request counts odd and even, above and below one
    request (or pair) per CTA, and voter counts from 0 to beyond the prefetch window."""
    rng = np.random.default_rng(20261107)
    ns = [1, 2, 4, 7, 8, 9, 15, 16, 17, 23, 24, 25, 31, 32, 33, 40, 63, 64, 65, 96, 127, 128]
    ds = [32, 64, 96, 160, 256, 772, 1000, 1536, 2080]
    out = []
    for n in ns:
        for _ in range(3):
            d = int(rng.choice(ds))
            r = int(rng.choice([1, 2, 3, 150, 297, 301]))
            if n > 32 and d > 1000:
                r = min(r, 150)
            m = int(rng.choice([0, 1, 4, 5, 9, 16]))
            out.append((r, n, d, m))
    return out


@pytest.mark.parametrize("shape", _boundary_shapes(), ids=lambda s: "R%d_N%d_D%d_M%d" % s)
def test_every_route_boundary_meets_its_bar(ctx, lwc, oracle, shape):
    R, N, D, M = shape
    E, V, w = fast_batch(R, N, D, max(M, 1), seed=R + 13 * N + D)
    # a shared direction per request keeps the cosines positive: with independent gaussian rows at D = 32..64 the
    # weighted sums cancel (confidences beyond 1) and the 1e-5 bar would be a statement about conditioning, not kernels
    E += np.float32(0.9) * np.random.default_rng(R + N + D).standard_normal((R, 1, D), dtype=np.float32) * np.abs(E).mean(axis=2, keepdims=True)
    if R > 8:
        E[6, N - 1] = 0.0
    if M == 0:
        V, w = V[:, :0], w[:, :0]
    want = oracle.score_f64(E, V, w, oracle.SIM_COSINE)
    got = ctx.score_batch_device(E, V, w)
    assert compare(got, want, V) == 1.0, ctx.kernel_for(R, N, D, M)
    properties_hold(got, V)
    host = ctx.score_batch(E, V, w)  # few-request path, chunked path: same bits as device-resident
    for k in got:
        assert np.array_equal(host[k], got[k], equal_nan=True), (k, ctx.kernel_for(R, N, D, M))
    fast = ctx.score_batch_device(E, V, w, precision=lwc.PREC_TF32)
    compare(fast, want, V, rtol=TF32_RTOL, atol=1e-5, gap=5e-3, min_agreement=0.95, atol_of_max=TF32_RTOL)
    properties_hold(fast, V)


@pytest.mark.parametrize("shape", [(600, 128, 3072, 16), (301, 100, 2336, 5), (333, 57, 1024, 3), (2, 64, 96, 16)],
                         ids=lambda s: "R%d_N%d_D%d_M%d" % s)
def test_operand_a_in_tensor_memory_and_in_shared_memory_agree(lwc, oracle, shape):
    """The fp32-grade tcgen05 route takes its A operands (h for X, l for Y = l h^T) from TMEM by default (score_tc16a.cuh:
    the transform warps store them there, one thread per row) and both operands from shared memory under LWC_TC16_A=smem
    (score_tc16.cuh, Y = h l^T).  Same pieces, same X chunks; Y is the transpose and runs through one accumulator instead
    of two, so the last bits may differ -- both meet the bar."""
    R, N, D, M = shape
    E, V, w = fast_batch(R, N, D, M, seed=1200 + N + D)
    if R > 8:
        E[3] *= np.float32(1e-6)  # marked: the fix-up pass rewrites it on either route
    want = oracle.score_f64(E, V, w, oracle.SIM_COSINE)
    res = {}
    for a in ("tmem", "smem"):
        with env(LWC_TC16_A=a), lwc.Context(device=0) as c:
            res[a] = c.score_batch_device(E, V, w)
            assert compare(res[a], want, V) == 1.0, a
            properties_hold(res[a], V)
    assert np.array_equal(res["tmem"]["winner"], res["smem"]["winner"])
    np.testing.assert_allclose(res["tmem"]["conf"], res["smem"]["conf"], rtol=2e-5, atol=2e-7)  # each within 1e-5 + 1e-7 of the oracle


@pytest.mark.parametrize("shape", [(701, 4, 1536, 4), (300, 3, 100, 3), (7, 1, 64, 1), (903, 2, 3072, 0), (450, 4, 772, 9), (1, 2, 32, 2)],
                         ids=lambda s: "R%d_N%d_D%d_M%d" % s)
@pytest.mark.parametrize("prec", ["fp32", "tf32"])
def test_pools_of_at_most_four_share_a_tile(lwc, oracle, shape, prec):
    """N <= 4 on the ring kernel: 8 / N consecutive requests share one 8-row tile (their rows are contiguous in E); the
    finaliser tallies every request over its own N x N block of the 8 x 8 Gram matrix (warp_tally_group), adding in the
    order the single-request tail does.  Bit-equal to one request per tile (LWC_STREAM_GROUP=0), also for the short last
    group of an odd batch, and within the route's bar against the oracle."""
    R, N, D, M = shape
    p = lwc.PREC_FP32 if prec == "fp32" else lwc.PREC_TF32
    E, V, w = fast_batch(R, N, D, max(M, 1), seed=1300 + N + D)
    if M == 0:
        V, w = V[:, :0], w[:, :0]
    res = {}
    for grp in ("1", "0"):
        with env(LWC_STREAM_GROUP=grp, LWC_SMALL_IMPL="stream"), lwc.Context(device=0) as c:
            res[grp] = c.score_batch_device(E, V, w, precision=p)
    for k in res["1"]:
        assert np.array_equal(res["1"][k], res["0"][k], equal_nan=True), k
    want = oracle.score_f64(E, V, w, oracle.SIM_COSINE)
    if prec == "fp32":
        assert compare(res["1"], want, V) == 1.0
    else:
        compare(res["1"], want, V, rtol=TF32_RTOL, atol=1e-5, gap=5e-3, min_agreement=0.95, atol_of_max=TF32_RTOL)
    properties_hold(res["1"], V)
