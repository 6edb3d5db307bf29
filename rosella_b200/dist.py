"""One process per GPU (torchrun-style launchers) on top of the C ABI.

Product path: `init_comm` + `build_graph` — the exchange itself (ncclSend/Recv all-to-all of partial lists, merge, f64
finish, in-place all-gather) runs INSIDE librosella_gpu.so (rg_knn_build_dist, csrc/knn_dist.cu); torch.distributed only
carries the 128-byte communicator id and the bench's barriers.

The functions below `# ---- host-side model` restate the same exchange with torch collectives.  They are what the
world-size-2 gloo tests run on CPU (layout of the all-to-all, the merge input, uneven shards) and are not on the
product path."""
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


def record_cuts(offsets, world: int) -> list[int]:
    """The record ranges of rg_multi_tnf_count_ascii / rg_multi_sketch_contigs (csrc/rg_multi.cu) restated: rank r owns
    records [cuts[r], cuts[r + 1]) — contiguous, in order, about total / world bytes each (the first record whose start
    reaches r * (total / world)); no record is split, so per-record results only need concatenating."""
    offsets = np.asarray(offsets, np.uint64)
    n = len(offsets) - 1
    cuts = [0] * (world + 1)
    if n:
        base0, total = int(offsets[0]), int(offsets[n]) - int(offsets[0])
        for r in range(1, world):
            target = base0 + total // world * r
            cuts[r] = min(max(cuts[r - 1], int(np.searchsorted(offsets, np.uint64(target), side="left"))), n)
    cuts[world] = n
    return cuts


def init_comm(ctx, rank: int, world: int, group=None):
    """rg_comm_init on every rank: rank 0's ncclUniqueId travels through torch.distributed (any backend)."""
    import torch
    import torch.distributed as dist
    from . import gpu
    if world == 1:
        return
    on_gpu = dist.get_backend(group) == "nccl"
    uid = torch.from_numpy(gpu.comm_unique_id() if rank == 0 else np.zeros(128, np.uint8))
    if on_gpu:
        uid = uid.cuda()
    dist.broadcast(uid, src=0, group=group)
    ctx.comm_init(rank, world, uid.cpu().numpy())


def build_graph(ctx, which: int, k: int, n: int, gather: bool = True):
    """rg_knn_build_dist_dev: the whole graph `which` over the communicator of ctx (collective).
    -> (idx int32 [n,k], dist float32 [n,k]) torch tensors on the rank's device; gather=False leaves other ranks'
    rows unwritten."""
    import torch
    dev = torch.device("cuda", ctx.device)
    idx = torch.empty((n, k), dtype=torch.int32, device=dev)
    dst = torch.empty((n, k), dtype=torch.float32, device=dev)
    torch.cuda.synchronize(dev)          # the buffers come from torch's stream, the library works on ctx.stream
    ctx.knn_build_dist_dev(which, k, gather, idx.data_ptr(), dst.data_ptr())
    ctx.sync()
    return idx, dst


# ---- host-side model of the exchange (CPU tests)
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
    torch.cuda.synchronize(dev)      # the buffers above come from torch's stream; the work below runs on ctx.stream
    if qe > qb:
        ctx.knn_search_dev(which, k, qb, qe, idx.data_ptr(), dst.data_ptr())
    ext = torch.cuda.ExternalStream(ctx.stream, device=ctx.device)
    with torch.cuda.stream(ext):
        if world > 1:
            idx = all_gather_rows(idx, n, world)
            dst = all_gather_rows(dst, n, world)
    ctx.sync()
    return idx, dst


def exchange_partial_lists(part_d, part_i, n: int, rank: int, world: int, group=None):
    """The exchange step of the symmetric whole-graph coverage build.  Every rank holds partial candidate lists for
    ALL rows ([n, K'] distances and indices); the owner of a row shard needs every rank's lists of its rows.  One
    all-to-all per array: rank r sends rows shard_range(n, q, world) to rank q.
    -> ([world, rows_of_this_shard, K'] distances, same for indices), the layout rg_knn_depth_merge_dev reads."""
    import torch
    import torch.distributed as dist
    if world == 1:
        return part_d.unsqueeze(0), part_i.unsqueeze(0)
    sizes = [shard_range(n, r, world)[1] - shard_range(n, r, world)[0] for r in range(world)]
    mine = sizes[rank]
    out = []
    for t in (part_d, part_i):
        recv = torch.empty((world * mine,) + tuple(t.shape[1:]), dtype=t.dtype, device=t.device)
        dist.all_to_all_single(recv, t.contiguous(), output_split_sizes=[mine] * world, input_split_sizes=sizes, group=group)
        out.append(recv.view((world, mine) + tuple(t.shape[1:])))
    return out[0], out[1]


def build_depth_graph_symmetric(ctx, k: int, n: int, rank: int, world: int):
    """Whole coverage graph on `world` GPUs with each unordered pair evaluated once: this rank's share
    of the row blocks (rg_knn_depth_partial_dev) -> all-to-all of the partial lists -> merge + f64 distances of this
    rank's row shard (rg_knn_depth_merge_dev) -> all-gather.  -> (idx int32 [n,k], dist float32 [n,k]) on the device."""
    import torch
    from . import gpu
    dev = torch.device("cuda", ctx.device)
    kp = ctx.knn_kprime(gpu.GRAPH_DEPTH, k)
    qb, qe = shard_range(n, rank, world)
    part_d = torch.empty((n, kp), dtype=torch.float32, device=dev)
    part_i = torch.empty((n, kp), dtype=torch.int32, device=dev)
    idx = torch.empty((qe - qb, k), dtype=torch.int32, device=dev)
    dst = torch.empty((qe - qb, k), dtype=torch.float32, device=dev)
    torch.cuda.synchronize(dev)      # the buffers above come from torch's stream; the work below runs on ctx.stream
    ctx.knn_depth_partial_dev(k, rank, world, part_d.data_ptr(), part_i.data_ptr())
    ext = torch.cuda.ExternalStream(ctx.stream, device=ctx.device)
    with torch.cuda.stream(ext):
        all_d, all_i = exchange_partial_lists(part_d, part_i, n, rank, world)
        if qe > qb:
            ctx.knn_depth_merge_dev(k, world, all_d.data_ptr(), all_i.data_ptr(), qb, qe, idx.data_ptr(), dst.data_ptr())
        if world > 1:
            idx = all_gather_rows(idx, n, world)
            dst = all_gather_rows(dst, n, world)
    ctx.sync()
    return idx, dst

