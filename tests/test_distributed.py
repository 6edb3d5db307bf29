"""CPU coverage of the N>1 host logic: world-size-2 gloo run of the row sharding and the
all-gather of per-rank slabs.  The per-rank lists come from the oracle here (no GPU in this test);
on GPUs the same code path gathers what rg_knn_search_dev wrote."""
import os
import socket

import numpy as np
import pytest

from rosella_b200 import dist as rdist


def test_shard_range_partitions_rows():
    for n in (1, 31, 32, 33, 1000, 1_000_000):
        for world in (1, 2, 3, 8):
            blocks = [rdist.shard_range(n, r, world) for r in range(world)]
            assert blocks[0][0] == 0 and blocks[-1][1] == n
            for (a0, a1), (b0, b1) in zip(blocks, blocks[1:]):
                assert a1 == b0 and a0 <= a1
            sizes = [b - a for a, b in blocks]
            assert max(sizes) - min(sizes) <= 32 + 31
            assert all(a % 32 == 0 for a, _ in blocks if a < n)


def _worker(rank, world, port, n, k, out):
    import torch
    import torch.distributed as dist

    import oracle
    from rosella_b200 import synth
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    tnf, cov, _ = synth.knn_tables(n, 3, seed=5)
    qb, qe = rdist.shard_range(n, rank, world)
    idx, dst, _ = oracle.knn(cov, oracle.DEPTH, k, qrows=np.arange(qb, qe, dtype=np.uint64), threads=2)
    gi = rdist.all_gather_rows(torch.from_numpy(idx.astype(np.int32)), n, world)
    gd = rdist.all_gather_rows(torch.from_numpy(dst), n, world)
    if rank == 0:
        np.savez(out, idx=gi.numpy().astype(np.uint32), dist=gd.numpy())
    dist.barrier()
    dist.destroy_process_group()


def test_gloo_world2_all_gather_matches_single_process(tmp_path):
    import torch.multiprocessing as mp

    import oracle
    from rosella_b200 import synth
    n, k, world = 150, 7, 2
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    out = str(tmp_path / "gathered.npz")
    mp.spawn(_worker, args=(world, port, n, k, out), nprocs=world, join=True)
    got = np.load(out)
    _, cov, _ = synth.knn_tables(n, 3, seed=5)
    idx, dst, _ = oracle.knn(cov, oracle.DEPTH, k)
    assert np.array_equal(got["idx"], idx) and np.array_equal(got["dist"], dst)


@pytest.mark.gpu
def test_single_gpu_distributed_path_equals_direct(ctx):
    """build_graph_distributed at world size 1 (no process group needed) equals rg_knn_search."""
    from rosella_b200 import gpu, synth
    tnf, cov, _ = synth.knn_tables(700, 4, seed=8)
    ctx.knn_load(tnf, cov)
    i1, d1, _ = ctx.knn_search(gpu.GRAPH_DEPTH, 9)
    i2, d2 = rdist.build_graph_distributed(ctx, gpu.GRAPH_DEPTH, 9, 700, 0, 1)
    assert np.array_equal(i2.cpu().numpy().astype(np.uint32), i1) and np.array_equal(d2.cpu().numpy(), d1)
