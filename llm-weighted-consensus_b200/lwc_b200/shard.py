"""Sharding bookkeeping for the multi-GPU paths (no device code here).

* Batched scoring shards by REQUEST: requests are independent (client.rs:385-456 touches only its own
  aggregate), rank g takes one contiguous block, results are concatenated in rank order; no collective.
* The single giant pool shards by ROW: rank g holds rows [g*N/G, (g+1)*N/G) of E and the matching
  columns of the vote matrix; lwc_pool_consensus all-gathers the normalised rows (NCCL) and every rank
  ends with identical winner/agree and its own slice of conf/cw.
"""
import numpy as np


def request_block(R: int, rank: int, world: int):
    """Contiguous [begin, end) of the R requests owned by `rank` (sizes differ by at most one)."""
    return (rank * R) // world, ((rank + 1) * R) // world


def pool_rows(n_total: int, rank: int, world: int):
    """Rows of the pool owned by `rank`; the pool path needs equal shards."""
    if n_total % world:
        raise ValueError(f"pool of {n_total} rows does not split evenly over {world} ranks")
    n = n_total // world
    return rank * n, (rank + 1) * n


def pool_shard(E, V, rank: int, world: int):
    """(E_own [n,D], V_own [M,n]) for `rank`; V's NaN rows (absent voters) stay NaN on every rank."""
    b, e = pool_rows(E.shape[0], rank, world)
    return np.ascontiguousarray(E[b:e]), np.ascontiguousarray(V[:, b:e])


def max_over_ranks(dist, value: float, device="cpu") -> float:
    """Timing rule of the bench: a multi-GPU step takes as long as its slowest rank."""
    if dist is None or not dist.is_initialized() or dist.get_world_size() == 1:
        return float(value)
    import torch

    t = torch.tensor([float(value)], dtype=torch.float64, device=device)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def broadcast_bytes(dist, payload: bytes, src: int = 0, device="cpu") -> bytes:
    """Ships the 128-byte NCCL unique id (lwc_comm_unique_id) from rank `src` to every rank."""
    import torch

    buf = torch.zeros(len(payload) if payload else 128, dtype=torch.uint8, device=device)
    if dist.get_rank() == src:
        buf.copy_(torch.frombuffer(bytearray(payload), dtype=torch.uint8))
    dist.broadcast(buf, src=src)
    return bytes(buf.cpu().numpy().tobytes())
