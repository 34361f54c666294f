"""world_size-2 `gloo` tests of the host-side multi-rank logic (no GPU): request/row sharding, the
max-over-ranks timing rule, the unique-id broadcast, and reassembly of per-rank results."""
import os

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp


def _worker(rank, world, port, out):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    import sys

    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    sys.path.insert(0, os.path.join(root, "llm-weighted-consensus_b200"))
    from lwc_b200 import shard

    res = {}
    # timing rule: every rank must end up with the slowest rank's time
    res["max"] = shard.max_over_ranks(dist, 1.0 + rank)
    # unique id travels from rank 0
    ident = bytes(range(128)) if rank == 0 else b""
    res["id"] = shard.broadcast_bytes(dist, ident if rank == 0 else bytes(128), src=0)
    # request sharding covers [0, R) exactly once, in rank order
    res["block"] = shard.request_block(4097, rank, world)
    # pool sharding + reassembly: each rank "scores" its rows with a stand-in (row sums) and the slices are gathered
    rng = np.random.default_rng(0)
    E = rng.standard_normal((64, 8)).astype(np.float32)
    V = rng.random((3, 64)).astype(np.float32)
    V[1] = np.nan
    Eo, Vo = shard.pool_shard(E, V, rank, world)
    mine = torch.from_numpy(Eo.sum(axis=1))
    parts = [torch.empty_like(mine) for _ in range(world)]
    dist.all_gather(parts, mine)
    res["gathered_ok"] = bool(np.allclose(torch.cat(parts).numpy(), E.sum(axis=1)))
    res["nan_rows_everywhere"] = bool(np.all(np.isnan(Vo[1])) and not np.any(np.isnan(Vo[0])))
    out[rank] = res
    dist.destroy_process_group()


def test_two_rank_host_logic():
    world = 2
    mgr = mp.Manager()
    out = mgr.dict()
    port = 29500 + (os.getpid() % 2000)
    mp.spawn(_worker, args=(world, port, out), nprocs=world, join=True)
    assert out[0]["max"] == out[1]["max"] == 2.0
    assert out[0]["id"] == out[1]["id"] == bytes(range(128))
    assert out[0]["block"] == (0, 2048) and out[1]["block"] == (2048, 4097)
    assert all(out[r]["gathered_ok"] and out[r]["nan_rows_everywhere"] for r in range(world))


def test_shard_helpers_edge_cases():
    import sys

    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    sys.path.insert(0, os.path.join(root, "llm-weighted-consensus_b200"))
    from lwc_b200 import shard

    covered = []
    for r in range(8):
        b, e = shard.request_block(65536 + 3, r, 8)
        covered += list(range(b, e))
    assert covered == list(range(65536 + 3))
    assert shard.request_block(3, 5, 8) == (1, 2)  # more ranks than requests: some blocks are empty
    assert shard.request_block(3, 0, 8) == (0, 0)
    with pytest.raises(ValueError):
        shard.pool_rows(10, 0, 4)
    assert shard.max_over_ranks(None, 3.5) == 3.5


def _table_worker(rank, world, port, out):
    """Each rank owns a row shard of the training table; the neighbour lists (here from the f64 checker, standing in
    for lwc_table_weights on that rank's GPU) are all-gathered and merged on every rank with lwc_table_merge."""
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    import sys

    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    sys.path.insert(0, os.path.join(root, "llm-weighted-consensus_b200"))
    sys.path.insert(0, root)
    import lwc_b200
    import oracle
    from lwc_b200 import shard

    rng = np.random.default_rng(77)
    rows, D, M, nq, top = 501, 24, 4, 5, 7
    T = rng.standard_normal((rows, D)).astype(np.float32)
    Q = rng.random((rows, M)).astype(np.float32)
    q = rng.standard_normal((nq, D)).astype(np.float32)
    base = np.ones(M, np.float32)
    a, b = shard.request_block(rows, rank, world)  # contiguous row block of this rank
    _, idx, sim = oracle.table_weights(T[a:b], Q[a:b], q, top, base, base * 0.5, base * 2)
    qr = Q[a:b][np.where(idx >= 0, idx, 0)]
    g_idx = [torch.empty((nq, top), dtype=torch.int32) for _ in range(world)]
    g_sim = [torch.empty((nq, top), dtype=torch.float32) for _ in range(world)]
    g_q = [torch.empty((nq, top, M), dtype=torch.float32) for _ in range(world)]
    g_off = [torch.empty(1, dtype=torch.int64) for _ in range(world)]
    dist.all_gather(g_idx, torch.from_numpy(idx.copy()))
    dist.all_gather(g_sim, torch.from_numpy(sim.astype(np.float32)))
    dist.all_gather(g_q, torch.from_numpy(qr.copy()))
    dist.all_gather(g_off, torch.tensor([a], dtype=torch.int64))
    w, grows, _ = lwc_b200.table_merge([int(o.item()) for o in g_off], torch.stack(g_idx).numpy(), torch.stack(g_sim).numpy(),
                                       torch.stack(g_q).numpy(), base, base * 0.5, base * 2)
    we, ie, _ = oracle.table_weights(T, Q, q, top, base, base * 0.5, base * 2)
    out[rank] = {"rows_ok": bool((grows == ie).all()), "w_ok": bool(np.allclose(w, we, rtol=2e-5)), "w": w.tobytes()}
    dist.destroy_process_group()


def test_two_rank_table_shards_merge_to_the_unsplit_table():
    world = 2
    mgr = mp.Manager()
    out = mgr.dict()
    port = 31500 + (os.getpid() % 2000)
    mp.spawn(_table_worker, args=(world, port, out), nprocs=world, join=True)
    assert all(out[r]["rows_ok"] and out[r]["w_ok"] for r in range(world))
    assert out[0]["w"] == out[1]["w"]  # every rank derives the same weights
