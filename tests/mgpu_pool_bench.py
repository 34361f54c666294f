"""BASELINE.json config 5 on a whole node: one pool of 262,144 candidates x 1024 dims, 16 voters, rows
sharded over the ranks, NCCL all-gather of the normalised rows over NVLink.  Run under torchrun:
    python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29631 tests/mgpu_pool_bench.py
Times the whole lwc_pool_consensus call on the device (max over ranks) and prints one JSON line."""
import ctypes
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "llm-weighted-consensus_b200")):
    sys.path.insert(0, p)


def main():
    import torch
    import torch.distributed as dist

    import lwc_b200
    from lwc_b200 import shard

    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    n_total = int(os.environ.get("LWC_POOL_N", 262144))
    D, M = 1024, 16
    steps, warmup = int(os.environ.get("LWC_POOL_STEPS", 5)), 2
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    ident = lwc_b200.comm_unique_id() if rank == 0 else bytes(128)
    ident = shard.broadcast_bytes(dist, ident, src=0, device="cuda")
    ctx = lwc_b200.Context(device=local)
    ctx.comm_init(world, rank, ident)
    b, e = shard.pool_rows(n_total, rank, world)
    n = e - b
    fused = os.environ.get("LWC_POOL_P2P", "0") == "1"
    if fused:
        mine = ctx.pool_p2p_export(n, D)
        handles = [None] * world
        dist.all_gather_object(handles, mine)
        ctx.pool_p2p_attach(b"".join(handles))
    g = torch.Generator(device="cuda").manual_seed(100 + rank)
    E = torch.randn((n, D), dtype=torch.float32, device="cuda", generator=g)
    V = torch.rand((M, n), dtype=torch.float32, device="cuda", generator=g) ** 6
    V /= V.sum(dim=1, keepdim=True) * world
    w = torch.linspace(0.5, 2.0, M, device="cuda")
    conf = torch.empty((n,), dtype=torch.float32, device="cuda")
    cw = torch.empty((n,), dtype=torch.float32, device="cuda")
    agree = torch.empty((M,), dtype=torch.float32, device="cuda")
    winner = torch.zeros((1,), dtype=torch.int64, device="cuda")
    desc = lwc_b200.PoolDesc(n, D, M, lwc_b200.PREC_BF16, lwc_b200.MEM_DEVICE)
    vp = ctypes.c_void_p

    def step():
        rc = ctx.lib.lwc_pool_consensus(ctx.handle, ctypes.byref(desc), vp(E.data_ptr()), vp(V.data_ptr()), vp(w.data_ptr()),
                                        vp(conf.data_ptr()), vp(cw.data_ptr()), vp(agree.data_ptr()), vp(winner.data_ptr()), None)
        if rc:
            raise lwc_b200.LwcError(rc, ctx.lib.lwc_last_error_message(ctx.handle).decode())

    for _ in range(warmup):
        step()
    dist.barrier()
    torch.cuda.synchronize()
    ext = torch.cuda.ExternalStream(ctx.stream())
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(ext)
    for _ in range(steps):
        step()
    e1.record(ext)
    torch.cuda.synchronize()
    dist.barrier()
    ms = shard.max_over_ranks(dist, e0.elapsed_time(e1) / steps, device="cuda")
    total = torch.tensor([float(conf.sum())], device="cuda")
    dist.all_reduce(total)
    winners = [None] * world
    dist.all_gather_object(winners, int(winner.item()))
    if rank == 0:
        flops = 2.0 * n_total * n_total * D
        print(json.dumps({"metric": "pool_all_pairs_per_sec", "value": 1e3 / ms, "unit": "pools/s", "n_gpus": world,
                          "ms_per_step": ms, "steps": steps, "scaling": "strong",
                          "config": {"workload": f"cfg5: one pool of {n_total} x {D}, {M} voters, rows sharded x{world}, " + ("gather fused into the tensor kernel (P2P over NVLink)" if fused else "NCCL allgather (bf16 rows)")},
                          "aggregate_tflops": flops / (ms * 1e-3) / 1e12, "per_gpu_tflops": flops / (ms * 1e-3) / 1e12 / world,
                          "sum_conf": float(total.item()), "winner_same_on_all_ranks": len(set(winners)) == 1}), flush=True)
    ctx.comm_destroy()
    ctx.close()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
