"""Run under torchrun with N >= 2 ranks, one GPU each:
    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29611 tests/mgpu_pool_check.py
Row-sharded lwc_pool_consensus (NCCL all-gather of the normalised rows) against the single-GPU result
of the same pool and against the CPU oracle (centroid identity).  Prints MGPU_POOL_OK on rank 0."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "llm-weighted-consensus_b200"), os.path.join(ROOT, "tests")):
    sys.path.insert(0, p)


def main():
    import torch
    import torch.distributed as dist

    import lwc_b200
    from lwc_b200 import shard

    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("gloo")  # plumbing only: the data path uses the library's own NCCL communicator
    ident = lwc_b200.comm_unique_id() if rank == 0 else bytes(128)
    ident = shard.broadcast_bytes(dist, ident, src=0)
    ctx = lwc_b200.Context(device=local)
    ctx.comm_init(world, rank, ident)

    N, D, M = 8192 * world, 256, 8
    rng = np.random.default_rng(3)
    E = rng.standard_normal((N, D)).astype(np.float32)
    E[: N // 4] += 1.5 * rng.standard_normal((1, D)).astype(np.float32)
    V = (rng.random((M, N)) ** 6).astype(np.float32)
    V /= V.sum(axis=1, keepdims=True)
    V[2] = np.nan
    w = (rng.integers(50, 201, M) / 100).astype(np.float32)
    Eo, Vo = shard.pool_shard(E, V, rank, world)
    got_nccl = ctx.pool_consensus(Eo, Vo, w)  # exchange step = NCCL all-gather before the tensor pass
    # now the fused route: peers' shards are pulled over NVLink inside the tensor kernel
    mine = ctx.pool_p2p_export(Eo.shape[0], D)
    handles = [None] * world
    dist.all_gather_object(handles, mine)
    ctx.pool_p2p_attach(b"".join(handles))
    got = ctx.pool_consensus(Eo, Vo, w)
    got2 = ctx.pool_consensus(Eo, Vo, w)      # twice: arrival counters and buffers are reused
    top_local = float(np.max(np.abs(got_nccl["conf"]))) + 1e-30
    assert np.max(np.abs(got["conf"] - got_nccl["conf"])) <= 1e-4 * top_local, "fused and NCCL routes disagree"
    assert got["winner"] == got_nccl["winner"]
    for k in ("conf", "cw", "agree"):
        assert np.array_equal(got[k], got2[k], equal_nan=True), k

    # Growing the exported buffers frees what the peers have mapped: a larger pool must end the fused route on
    # every rank (NCCL all-gather from then on), and the first shape must still score correctly afterwards.
    Eb = np.concatenate([Eo, Eo[::-1]])
    Vb = np.concatenate([Vo, Vo[:, ::-1]], axis=1)
    big = ctx.pool_consensus(Eb, Vb, w)
    sums = [None] * world
    dist.all_gather_object(sums, float(big["conf"].sum()))
    assert abs(sum(sums) - 1.0) < 1e-3, sums
    after = ctx.pool_consensus(Eo, Vo, w)
    for k in ("conf", "cw", "agree"):
        assert np.array_equal(after[k], got_nccl[k], equal_nan=True), ("after growth: not the NCCL route's bits", k)
    # export + attach again: the fused route is back and gives its earlier bits
    mine = ctx.pool_p2p_export(Eo.shape[0], D)
    dist.all_gather_object(handles, mine)
    ctx.pool_p2p_attach(b"".join(handles))
    again = ctx.pool_consensus(Eo, Vo, w)
    for k in ("conf", "cw", "agree"):
        assert np.array_equal(again[k], got[k], equal_nan=True), ("re-attached fused route", k)

    # every rank must hold identical winner/agree; conf slices reassemble to the whole
    conf_parts = [None] * world
    dist.all_gather_object(conf_parts, got["conf"])
    meta = [None] * world
    dist.all_gather_object(meta, (got["winner"], got["agree"].tolist()))
    ok = True
    if rank == 0:
        import oracle

        conf = np.concatenate(conf_parts)
        assert all(m[0] == meta[0][0] for m in meta), meta
        assert all(np.array_equal(np.array(m[1]), np.array(meta[0][1]), equal_nan=True) for m in meta)
        want = oracle.pool_f64(E, V, w, method=1)
        top = float(want["conf"].max())
        err = float(np.max(np.abs(conf - want["conf"])) / top)
        print(f"sharded pool over {world} GPUs: N={N} max conf err / top = {err:.2e}, winner {got['winner']} vs {want['winner']}")
        assert err <= 1e-2
        srt = np.sort(want["conf"])[::-1]
        if srt[0] - srt[1] > 2e-2 * top:
            assert got["winner"] == want["winner"]
        assert abs(conf.sum() - 1.0) < 1e-3
        # the same pool on one GPU (no communicator): results must agree within the bf16 tolerance
        solo = lwc_b200.Context(device=local)
        one = solo.pool_consensus(E, V, w)
        solo.close()
        assert np.max(np.abs(one["conf"] - conf)) <= 1e-2 * top
        assert one["winner"] == got["winner"]
        print("MGPU_POOL_OK")
    ctx.comm_destroy()
    ctx.close()
    dist.barrier()
    dist.destroy_process_group()
    return 0 if ok else 1


if __name__ == "__main__":
    sys.exit(main())
