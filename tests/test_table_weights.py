"""Training-table weights: CUDA path vs the f64 restatement of the same definition.

The reference slot is `unimplemented!()` (src/score/completions/weight.rs:99-117), so parity here is against
this build's own specification (include/lwc_b200.h), not against reference arithmetic: parity unpinned.
"""
import numpy as np
import pytest

SIM_ATOL = 2e-6   # fp32 cosine of unit-scale rows vs f64
W_RTOL = 2e-5


def make_table(seed, rows, D, M, nq):
    rng = np.random.default_rng(seed)
    T = rng.standard_normal((rows, D)).astype(np.float32)
    Q = rng.random((rows, M)).astype(np.float32)
    # prompts near existing rows (clear neighbours) plus fresh ones
    pick = rng.integers(0, rows, nq)
    q = (T[pick] + 0.3 * rng.standard_normal((nq, D))).astype(np.float32)
    q[::3] = rng.standard_normal((len(q[::3]), D)).astype(np.float32)
    base = (1.0 + rng.random(M)).astype(np.float32)
    minw = (base * 0.5).astype(np.float32)
    maxw = (base * 1.5).astype(np.float32)
    return T, Q, q, base, minw, maxw


def test_oracle_table_weights_properties(oracle):
    T, Q, q, base, minw, maxw = make_table(1, 300, 48, 5, 7)
    w, idx, sim = oracle.table_weights(T, Q, q, 4, base, minw, maxw)
    assert ((w >= minw - 1e-7) & (w <= maxw + 1e-7)).all()
    assert (np.diff(sim, axis=1) <= 0).all()
    Tn = T / np.linalg.norm(T, axis=1, keepdims=True)
    qn = q / np.linalg.norm(q, axis=1, keepdims=True)
    S = qn.astype(np.float64) @ Tn.astype(np.float64).T
    order = np.argsort(-S, axis=1, kind="stable")[:, :4]
    assert (order == idx).all()
    # a table whose every row rates all LLMs equally gives back the base weights
    w2, _, _ = oracle.table_weights(T, np.ones_like(Q), q, 4, base, minw, maxw)
    np.testing.assert_allclose(w2, np.broadcast_to(base, w2.shape), rtol=1e-12)
    # the query itself in the table is its own first neighbour
    w3, idx3, sim3 = oracle.table_weights(T, Q, T[17:18] * 3.0, 1, base, minw, maxw)
    assert idx3[0, 0] == 17 and abs(sim3[0, 0] - 1.0) < 1e-6
    with pytest.raises(ValueError):
        oracle.table_weights(T, Q, q, 4, base, maxw, minw)


@pytest.mark.gpu
@pytest.mark.parametrize("rows,D,M,nq,top", [
    (1, 8, 1, 1, 1), (5, 7, 3, 2, 8), (1000, 64, 8, 19, 5), (4099, 1536, 16, 9, 16), (20000, 384, 128, 8, 64),
    (3001, 3072, 4, 3, 4), (50001, 1536, 16, 8, 16),
])
@pytest.mark.parametrize("impl", ["ring", "direct"])
def test_table_weights_match_oracle(oracle, monkeypatch, rows, D, M, nq, top, impl):
    import lwc_b200
    if impl == "direct":
        monkeypatch.setenv("LWC_TABLE_IMPL", "direct")
    else:
        monkeypatch.delenv("LWC_TABLE_IMPL", raising=False)
    T, Q, q, base, minw, maxw = make_table(rows + top, rows, D, M, nq)
    with lwc_b200.Context(0) as ctx:
        ctx.table_load(T, Q)
        w, idx, sim = ctx.table_weights(q, top, base, minw, maxw)
    we, idxe, sime = oracle.table_weights(T, Q, q, top, base, minw, maxw)
    np.testing.assert_allclose(sim, sime, atol=SIM_ATOL, rtol=0)
    # neighbour order must agree wherever f64 separates neighbours by more than the fp32 cosine error
    for p in range(nq):
        for k in range(top):
            if idx[p, k] == idxe[p, k]:
                continue
            assert idxe[p, k] >= 0 and idx[p, k] >= 0
            near = np.abs(sime[p] - sime[p, k]) < 2 * SIM_ATOL
            assert near.sum() > 1 or k == top - 1, (p, k, idx[p], idxe[p])
    same = (idx == idxe).all(axis=1)
    assert same.mean() > 0.8
    np.testing.assert_allclose(w[same], we[same], rtol=W_RTOL)
    assert ((w >= minw) & (w <= maxw)).all()


@pytest.mark.gpu
def test_table_weights_errors_and_reload(oracle):
    import lwc_b200
    T, Q, q, base, minw, maxw = make_table(3, 64, 32, 4, 3)
    with lwc_b200.Context(0) as ctx:
        with pytest.raises(lwc_b200.LwcError) as e:
            ctx.table_weights(q, 2, base, minw, maxw)
        assert e.value.code == 400 and "no table" in str(e.value)
        ctx.table_load(T, Q)
        for bad in (0, 65):
            with pytest.raises(lwc_b200.LwcError):
                ctx.table_weights(q, bad, base, minw, maxw)
        with pytest.raises(lwc_b200.LwcError):
            ctx.table_weights(q, 2, base, maxw, minw)
        w1, i1, _ = ctx.table_weights(q, 2, base, minw, maxw)
        # zero prompt -> no positive neighbour -> base weights
        wz, _, sz = ctx.table_weights(np.zeros((1, 32), np.float32), 2, base, minw, maxw)
        np.testing.assert_array_equal(wz[0], base)
        assert (sz == 0).all()
        # both score-pass routes agree (their summation orders differ)
        import os
        os.environ["LWC_TABLE_IMPL"] = "direct"
        try:
            wd, idd, simd = ctx.table_weights(q, 2, base, minw, maxw)
        finally:
            del os.environ["LWC_TABLE_IMPL"]
        wr, idr, simr = ctx.table_weights(q, 2, base, minw, maxw)
        assert (idd == idr).all()
        np.testing.assert_allclose(simd, simr, atol=SIM_ATOL, rtol=0)
        np.testing.assert_allclose(wd, wr, rtol=W_RTOL)
        # reload a smaller table; results follow the new table, and the run is reproducible
        ctx.table_load(T[:10], Q[:10])
        w2, i2, _ = ctx.table_weights(q, 2, base, minw, maxw)
        assert i2.max() < 10
        w3, i3, _ = ctx.table_weights(q, 2, base, minw, maxw)
        assert (w2 == w3).all() and (i2 == i3).all()
        we, ie, _ = oracle.table_weights(T[:10], Q[:10], q, 2, base, minw, maxw)
        assert (i2 == ie).all()
        np.testing.assert_allclose(w2, we, rtol=W_RTOL)


@pytest.mark.gpu
def test_table_topk_tree_three_levels_and_ties():
    """1.1M rows with top 64 needs three levels of the top-k tree; identical rows must come out lowest row first."""
    import lwc_b200
    rows, top = 1_100_000, 64
    rng = np.random.default_rng(11)
    perm = rng.permutation(rows)
    T = np.zeros((rows, 4), np.float32)
    T[:, 0] = 1.0
    T[:, 1] = perm * 0.01          # cosine against [1, 0, 0, 0] falls strictly with perm
    Q = np.ones((rows, 2), np.float32)
    q = np.array([[1, 0, 0, 0], [2, -2, 0, 0]], np.float32)
    one = np.ones(2, np.float32)
    expect = np.argsort(perm)[:top]
    with lwc_b200.Context(0) as ctx:
        ctx.table_load(T, Q)
        _, idx, sim = ctx.table_weights(q, top, one, one, one)
        assert (idx[0] == expect).all() and (idx[1] == expect).all()
        assert (np.diff(sim, axis=1) < 0).all()
        Tt = np.ones((20000, 8), np.float32)
        Tt[::2] *= 4.0  # same direction, other length: the cosine does not change
        ctx.table_load(Tt, np.ones((20000, 2), np.float32))
        _, idx, sim = ctx.table_weights(np.ones((1, 8), np.float32), top, one, one, one)
        assert (idx[0] == np.arange(top)).all()
        # fewer rows than `top`: the tail is -1 / 0
        ctx.table_load(Tt[:5], np.ones((5, 2), np.float32))
        _, idx, sim = ctx.table_weights(np.ones((1, 8), np.float32), 8, one, one, one)
        assert (idx[0] == [0, 1, 2, 3, 4, -1, -1, -1]).all() and (sim[0, 5:] == 0).all()


@pytest.mark.gpu
def test_table_weights_device_memory_feeds_score_batch(oracle):
    """Device pointers in, device weights out, asynchronous on the context's stream: the weights go straight into a
    scoring call on that stream, and equal what the host-memory call returns bit for bit."""
    import torch
    import lwc_b200
    from lwc_b200 import synth
    R, N, D, M, top = 19, 16, 256, 4, 5
    T, Q, q, base, minw, maxw = make_table(21, 3000, D, M, R)
    E, V, _, _ = synth.make_batch(R, N, D, M)
    with lwc_b200.Context(0) as ctx:
        ctx.table_load(T, Q)
        w_host, idx_host, sim_host = ctx.table_weights(q, top, base, minw, maxw)
        st = torch.cuda.ExternalStream(ctx.stream())
        with torch.cuda.stream(st):
            dq = torch.from_numpy(q).cuda()
            dE, dV = torch.from_numpy(E).cuda(), torch.from_numpy(V).cuda()
            dw = torch.empty((R, M), dtype=torch.float32, device="cuda")
            didx = torch.empty((R, top), dtype=torch.int32, device="cuda")
            dconf = torch.empty((R, N), dtype=torch.float32, device="cuda")
            dwin = torch.empty((R,), dtype=torch.int32, device="cuda")
            torch.cuda.synchronize()
            ctx.table_weights_device(dq.data_ptr(), R, top, base, minw, maxw, dw.data_ptr(), didx.data_ptr(), None)
            desc = lwc_b200.BatchDesc(R, N, D, M, lwc_b200.SIM_COSINE, lwc_b200.PREC_FP32, lwc_b200.MEM_DEVICE, 1)
            ctx.score_raw(desc, dE.data_ptr(), dV.data_ptr(), dw.data_ptr(), dconf.data_ptr(), None, None, dwin.data_ptr(),
                          None)
            ctx.synchronize()
        assert (dw.cpu().numpy() == w_host).all() and (didx.cpu().numpy() == idx_host).all()
        ref = oracle.score_f64(E, V, w_host, oracle.SIM_COSINE)
        np.testing.assert_allclose(dconf.cpu().numpy(), ref["conf"], rtol=1e-5, atol=1e-7)
        assert ctx.lib.lwc_table_weights(ctx.handle, None, 1, 1, None, None, None, None, None, None, 7) == 400


def shard_candidates(oracle, T, Q, q, top, base, minw, maxw, bounds):
    """Per-shard neighbour lists as lwc_table_weights reports them (here from the f64 checker), plus Q rows."""
    idx, sim, qr = [], [], []
    for a, b in bounds:
        _, i, s = oracle.table_weights(T[a:b], Q[a:b], q, top, base, minw, maxw)
        rows = np.where(i >= 0, i, 0)
        idx.append(i)
        sim.append(s.astype(np.float32))
        qr.append(Q[a:b][rows])
    return np.stack(idx), np.stack(sim), np.stack(qr)


def test_table_merge_of_row_shards_matches_the_unsplit_table(oracle):
    import lwc_b200 as lwc
    """Host-only merge (no GPU): three uneven row shards, one of them smaller than `top`."""
    T, Q, q, base, minw, maxw = make_table(31, 700, 40, 6, 11)
    top = 8
    bounds = [(0, 5), (5, 400), (400, 700)]
    idx, sim, qr = shard_candidates(oracle, T, Q, q, top, base, minw, maxw, bounds)
    w, rows, sims = lwc.table_merge([a for a, _ in bounds], idx, sim, qr, base, minw, maxw)
    we, ie, se = oracle.table_weights(T, Q, q, top, base, minw, maxw)
    assert (rows == ie).all()
    np.testing.assert_allclose(sims, se, atol=1e-6)
    np.testing.assert_allclose(w, we, rtol=W_RTOL)
    # ties across shards go to the lowest global row; fewer candidates than `top` leave -1 / 0
    one = np.ones(2, np.float32)
    idx2 = np.array([[[0, -1]], [[0, -1]]], np.int32)
    sim2 = np.array([[[0.5, 0.0]], [[0.5, 0.0]]], np.float32)
    w2, rows2, sims2 = lwc.table_merge([10, 3], idx2, sim2, np.ones((2, 1, 2, 2), np.float32), one, one, one)
    assert rows2.tolist() == [[3, 10]] and sims2.tolist() == [[0.5, 0.5]]
    idx3 = np.full((2, 1, 3), -1, np.int32)
    idx3[1, 0, 0] = 4
    w3, rows3, _ = lwc.table_merge([0, 100], idx3, np.full((2, 1, 3), 0.25, np.float32), np.ones((2, 1, 3, 2), np.float32),
                                   one, one, one)
    assert rows3.tolist() == [[104, -1, -1]]
    with pytest.raises(lwc.LwcError):
        lwc.table_merge([0], np.zeros((1, 1, 1), np.int32), np.zeros((1, 1, 1), np.float32), np.ones((1, 1, 1, 2), np.float32),
                        one, 2 * one, one)


@pytest.mark.gpu
def test_table_split_over_two_contexts_gives_the_bits_of_the_unsplit_table():
    import lwc_b200
    T, Q, q, base, minw, maxw = make_table(41, 9000, 128, 5, 13)
    top = 6
    bounds = [(0, 3500), (3500, 9000)]
    with lwc_b200.Context(0) as ctx:
        ctx.table_load(T, Q)
        w_full, i_full, s_full = ctx.table_weights(q, top, base, minw, maxw)
        idx, sim, qr = [], [], []
        for a, b in bounds:
            ctx.table_load(T[a:b], Q[a:b])
            _, i, s = ctx.table_weights(q, top, base, minw, maxw)
            idx.append(i), sim.append(s), qr.append(Q[a:b][np.where(i >= 0, i, 0)])
    w, rows, sims = lwc_b200.table_merge([a for a, _ in bounds], np.stack(idx), np.stack(sim), np.stack(qr), base, minw, maxw)
    assert (rows == i_full).all() and (sims == s_full).all()
    assert (w == w_full).all()
