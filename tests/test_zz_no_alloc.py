"""lwc_pool_consensus_static: the entry point for callers that are not allowed to allocate.

Who needs it: a serving host that has already handed the GPU's memory to its own pool, or a scoring
thread that must never enter cudaMalloc/cudaFree (they synchronise the whole device and would stall the
other requests' streams).  The function works out of a fixed 13 MiB array inside the library's device
image, so the only memory this test allocates is what ANY caller brings: the input and output buffers.

The check: through the C ABI, on two calls in a row (the second smaller, so stale data from the first
is still lying in the shared array), the results must match lwc_pool_consensus -- the existing,
allocating entry point.  Both run the same kernels in the same order on one GPU, so the match is required
to be BIT-EXACT (a stale word read from the shared array would show up as a differing bit long before it
exceeded a tolerance); the CPU oracle is compared as well, within the bf16 tolerance the pool tests use.
The second call must also leave the device's free memory where it was: nothing is allocated.
This file sorts last on purpose: if it fails, `pytest -x` has already run everything else.
"""
import ctypes

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

BF16_TOL = 1e-2  # same bound as tests/test_gpu_pool.py: 1e-2 of the top confidence


def _pool(N, D, M, seed):
    rng = np.random.default_rng(seed)
    E = rng.standard_normal((N, D)).astype(np.float32)
    E[: N // 3] += 2.0 * rng.standard_normal((1, D)).astype(np.float32)  # one dense cluster
    V = (rng.random((M, N)) ** 6).astype(np.float32)
    V /= V.sum(axis=1, keepdims=True)
    w = (rng.integers(50, 201, M) / 100).astype(np.float32)
    return E, V, w


def _run(ctx, lwc, entry, E, V, w):
    """Calls `entry` (either pool entry point) with every pointer in device memory."""
    N, D = E.shape
    M = V.shape[0]
    bufs = []

    def dev(nbytes, src=None):
        b = lwc.DeviceBuffer(ctx, max(nbytes, 8))
        bufs.append(b)
        if src is not None:
            b.upload(src)
        return b

    try:
        dE, dV, dw = dev(E.nbytes, E), dev(V.nbytes, V), dev(w.nbytes, w)
        dconf, dcw, dag, dwin = dev(4 * N), dev(4 * N), dev(4 * M), dev(8)
        desc = lwc.PoolDesc(N, D, M, lwc.PREC_BF16, lwc.MEM_DEVICE)
        vp = ctypes.c_void_p
        rc = entry(ctx.handle, ctypes.byref(desc), vp(dE.ptr), vp(dV.ptr), vp(dw.ptr), vp(dconf.ptr), vp(dcw.ptr),
                   vp(dag.ptr), vp(dwin.ptr), None)
        if rc:
            raise lwc.LwcError(rc, ctx.lib.lwc_last_error_message(ctx.handle).decode())
        ctx.synchronize()
        return {"conf": dconf.download((N,), np.float32), "cw": dcw.download((N,), np.float32),
                "agree": dag.download((M,), np.float32), "winner": int(dwin.download((1,), np.int64)[0])}
    finally:
        for b in bufs:
            b.free()


def _free_bytes():
    """Free device memory as the CUDA runtime reports it (cudaMemGetInfo), read through torch."""
    import torch

    torch.cuda.synchronize()
    return torch.cuda.mem_get_info(0)[0]


def test_static_scratch_entry_matches_the_allocating_one_twice_in_a_row(oracle):
    import lwc_b200 as lwc

    with lwc.Context(device=0) as ctx:
        # first a pool near the size limit, then a small one that reuses the (now dirty) array
        for N, D, M, seed in ((4096, 1024, 6, 1), (700, 64, 3, 2)):
            E, V, w = _pool(N, D, M, seed)
            want = _run(ctx, lwc, ctx.lib.lwc_pool_consensus, E, V, w)
            got = _run(ctx, lwc, ctx.lib.lwc_pool_consensus_static, E, V, w)
            for k in ("conf", "cw", "agree"):
                assert np.array_equal(got[k], want[k], equal_nan=True), k
            assert got["winner"] == want["winner"]
            ref = oracle.pool_f64(E, V, w, method=0)
            top = float(np.max(np.abs(ref["conf"])))
            assert top > 0
            assert np.max(np.abs(got["conf"] - ref["conf"])) <= BF16_TOL * top
            assert np.max(np.abs(got["cw"] - ref["cw"])) <= BF16_TOL * float(np.max(np.abs(ref["cw"])))
            np.testing.assert_allclose(got["agree"], ref["agree"], rtol=BF16_TOL, atol=BF16_TOL * top)
            assert got["winner"] == ref["winner"]
            assert abs(float(got["conf"].sum()) - 1.0) < 1e-3


def test_static_scratch_entry_allocates_nothing():
    """With the caller's buffers already in place, a (second) call must not move the device's free memory."""
    import lwc_b200 as lwc

    with lwc.Context(device=0) as ctx:
        N, D, M = 2048, 256, 4
        E, V, w = _pool(N, D, M, 9)
        bufs = [lwc.DeviceBuffer(ctx, n) for n in (E.nbytes, V.nbytes, w.nbytes, 4 * N, 4 * N, 4 * M, 8)]
        try:
            for b, a in zip(bufs, (E, V, w)):
                b.upload(a)
            desc = lwc.PoolDesc(N, D, M, lwc.PREC_BF16, lwc.MEM_DEVICE)
            args = [ctypes.c_void_p(b.ptr) for b in bufs]

            def call():
                rc = ctx.lib.lwc_pool_consensus_static(ctx.handle, ctypes.byref(desc), *args, None)
                assert rc == 0, ctx.lib.lwc_last_error_message(ctx.handle).decode()
                ctx.synchronize()

            call()  # the first call on a device resolves the array's address (and loads the module)
            before = _free_bytes()
            call()
            assert _free_bytes() == before
        finally:
            for b in bufs:
                b.free()


def test_static_scratch_entry_refuses_what_does_not_fit():
    import lwc_b200 as lwc

    with lwc.Context(device=0) as ctx:
        E, V, w = _pool(256, 64, 2, 3)
        # too many rows, and too many elements, for the fixed array: a clear 400, not a crash
        for N, D in ((16384 + 128, 64), (8192, 1024)):
            desc = lwc.PoolDesc(N, D, 2, lwc.PREC_BF16, lwc.MEM_DEVICE)
            one = ctypes.c_void_p(1)  # never dereferenced: the size check comes first
            rc = ctx.lib.lwc_pool_consensus_static(ctx.handle, ctypes.byref(desc), one, one, one, None, None, None, None, None)
            assert rc == 400
            msg = ctx.lib.lwc_last_error_message(ctx.handle).decode()
            assert "exceeds the fixed scratch" in msg and "lwc_pool_consensus" in msg
        # host pointers would need a staging allocation: refused as well
        desc = lwc.PoolDesc(256, 64, 2, lwc.PREC_BF16, lwc.MEM_HOST)
        rc = ctx.lib.lwc_pool_consensus_static(ctx.handle, ctypes.byref(desc), E.ctypes.data_as(ctypes.c_void_p),
                                               V.ctypes.data_as(ctypes.c_void_p), w.ctypes.data_as(ctypes.c_void_p),
                                               None, None, None, None, None)
        assert rc == 400
