"""Multi-GPU host logic: one process per GPU, query rows sharded, references replicated, one
all-gather of the per-rank slabs.  Backend-agnostic: NCCL on GPUs, gloo in the CPU tests."""
from __future__ import annotations

import numpy as np

TILE = 32


def shard_range(n: int, rank: int, world: int, align: int = TILE) -> tuple[int, int]:
    """Contiguous block of query rows of `rank`: blocks are multiples of `align` rows (whole CTAs of the
    select kernels) except the last, and differ by at most `align` rows."""
    tiles = (n + align - 1) // align
    base, extra = divmod(tiles, world)
    t0 = rank * base + min(rank, extra)
    t1 = t0 + base + (1 if rank < extra else 0)
    return min(t0 * align, n), min(t1 * align, n)


def all_gather_rows(local, n: int, world: int, group=None):
    """all-gather row blocks of unequal length (shard_range) into an [n, k] tensor on every rank."""
    import torch
    import torch.distributed as dist
    sizes = [shard_range(n, r, world)[1] - shard_range(n, r, world)[0] for r in range(world)]
    mx = max(sizes)
    pad = torch.zeros((mx,) + tuple(local.shape[1:]), dtype=local.dtype, device=local.device)
    pad[: local.shape[0]] = local
    out = torch.empty((world * mx,) + tuple(local.shape[1:]), dtype=local.dtype, device=local.device)
    dist.all_gather_into_tensor(out, pad, group=group)
    return torch.cat([out[r * mx: r * mx + sizes[r]] for r in range(world)], dim=0)


def build_graph_distributed(ctx, which: int, k: int, n: int, rank: int, world: int):
    """Search this rank's shard on its GPU (tables already loaded with ctx.knn_load) and all-gather.
    -> (idx int32 [n,k], dist float32 [n,k]) torch tensors on the rank's device."""
    import torch
    qb, qe = shard_range(n, rank, world)
    dev = torch.device("cuda", ctx.device)
    idx = torch.empty((qe - qb, k), dtype=torch.int32, device=dev)
    dst = torch.empty((qe - qb, k), dtype=torch.float32, device=dev)
    if qe > qb:
        ctx.knn_search_dev(which, k, qb, qe, idx.data_ptr(), dst.data_ptr())
    ext = torch.cuda.ExternalStream(ctx.stream, device=ctx.device)
    with torch.cuda.stream(ext):
        if world > 1:
            idx = all_gather_rows(idx, n, world)
            dst = all_gather_rows(dst, n, world)
    ctx.sync()
    return idx, dst


def concat_host(parts: list[np.ndarray]) -> np.ndarray:
    return np.concatenate(parts, axis=0)
