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


def test_record_cuts_are_contiguous_and_balanced():
    rng = np.random.default_rng(8)
    for n in (0, 1, 2, 7, 500):
        lens = rng.integers(0, 5000, n)
        offs = np.zeros(n + 1, np.uint64); offs[1:] = np.cumsum(lens)
        offs += np.uint64(17)                                          # a table that does not start at byte 0
        for world in (1, 2, 3, 8):
            cuts = rdist.record_cuts(offs, world)
            assert cuts[0] == 0 and cuts[-1] == n and all(a <= b for a, b in zip(cuts, cuts[1:]))
            if n >= 200:
                share = [int(offs[b]) - int(offs[a]) for a, b in zip(cuts, cuts[1:])]
                assert max(share) - min(share) <= 2 * 5000


def _worker_records(rank, world, port, total_bp, out):
    """Every rank counts TNF and sketches ITS record range (the oracle stands in for the device) — no exchange on the
    data path; rank 0 only concatenates."""
    import torch.distributed as dist

    import oracle
    from rosella_b200 import synth
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    seq, offs = synth.contigs(total_bp, threads=1, seed=3)
    cuts = rdist.record_cuts(offs, world)
    a, b = cuts[rank], cuts[rank + 1]
    local_seq = seq[int(offs[a]):int(offs[b])]
    local_offs = offs[a:b + 1] - offs[a]
    counts, nk, _ = oracle.tnf_count(local_seq, local_offs, fast=True)
    h, so, snk, _ = oracle.sketch_contigs(local_seq, local_offs)
    parts = [None] * world
    dist.all_gather_object(parts, (counts, nk, h, so, snk))
    if rank == 0:
        sk_off = [np.zeros(1, np.uint64)]
        for p in parts:
            sk_off.append(p[3][1:] + sk_off[-1][-1])
        np.savez(out, counts=np.concatenate([p[0] for p in parts]), nk=np.concatenate([p[1] for p in parts]),
                 hashes=np.concatenate([p[2] for p in parts]), sk_off=np.concatenate(sk_off), snk=np.concatenate([p[4] for p in parts]))
    dist.barrier()
    dist.destroy_process_group()


def test_gloo_world2_record_ranges_need_no_exchange(tmp_path):
    """TNF counting and MinHash sketching on N ranks: contiguous record ranges, results concatenated."""
    import torch.multiprocessing as mp

    import oracle
    from rosella_b200 import synth
    world, total_bp = 2, 300_000
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    out = str(tmp_path / "records.npz")
    mp.spawn(_worker_records, args=(world, port, total_bp, out), nprocs=world, join=True)
    got = np.load(out)
    seq, offs = synth.contigs(total_bp, threads=1, seed=3)
    counts, nk, _ = oracle.tnf_count(seq, offs, fast=True)
    h, so, snk, _ = oracle.sketch_contigs(seq, offs)
    assert np.array_equal(got["counts"], counts) and np.array_equal(got["nk"], nk)
    assert np.array_equal(got["hashes"], h) and np.array_equal(got["sk_off"], so) and np.array_equal(got["snk"], snk)


def _worker_sym(rank, world, port, n, k, kp, out):
    """Partial lists of a pair partition (numpy stand-in for rg_knn_depth_partial_dev), exchanged with
    exchange_partial_lists, merged by (distance, index) (stand-in for rg_knn_depth_merge_dev), all-gathered."""
    import torch
    import torch.distributed as dist

    import oracle
    from rosella_b200 import synth
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    _, cov, _ = synth.knn_tables(n, 3, seed=5)
    full_i, full_d, _ = oracle.knn(cov, oracle.DEPTH, n - 1)             # every pair, sorted per row
    pd = np.full((n, kp), np.inf, np.float32)
    pi = np.full((n, kp), -1, np.int32)
    for r in range(n):
        nb = full_i[r].astype(np.int64)
        mine = (np.minimum(r, nb) + np.maximum(r, nb)) % world == rank   # each unordered pair belongs to one rank
        sel = np.where(mine)[0][:kp]
        pd[r, :len(sel)] = full_d[r, sel]
        pi[r, :len(sel)] = full_i[r, sel].astype(np.int32)
    all_d, all_i = rdist.exchange_partial_lists(torch.from_numpy(pd), torch.from_numpy(pi), n, rank, world)
    qb, qe = rdist.shard_range(n, rank, world)
    assert tuple(all_d.shape) == (world, qe - qb, kp)
    md, mi = all_d.numpy().transpose(1, 0, 2).reshape(qe - qb, -1), all_i.numpy().transpose(1, 0, 2).reshape(qe - qb, -1)
    idx = np.empty((qe - qb, k), np.int32)
    dst = np.empty((qe - qb, k), np.float32)
    for r in range(qe - qb):
        order = np.lexsort((mi[r].astype(np.uint32), md[r]))[:k]
        idx[r], dst[r] = mi[r][order], md[r][order]
    gi = rdist.all_gather_rows(torch.from_numpy(idx), n, world)
    gd = rdist.all_gather_rows(torch.from_numpy(dst), n, world)
    if rank == 0:
        np.savez(out, idx=gi.numpy().astype(np.uint32), dist=gd.numpy())
    dist.barrier()
    dist.destroy_process_group()


def test_gloo_world2_partial_list_exchange(tmp_path):
    """Host logic of the symmetric multi-GPU coverage build: all-to-all of partial lists, merge, all-gather."""
    import torch.multiprocessing as mp

    import oracle
    from rosella_b200 import synth
    n, k, kp, world = 150, 7, 16, 2
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    out = str(tmp_path / "sym.npz")
    mp.spawn(_worker_sym, args=(world, port, n, k, kp, out), nprocs=world, join=True)
    got = np.load(out)
    _, cov, _ = synth.knn_tables(n, 3, seed=5)
    idx, dst, _ = oracle.knn(cov, oracle.DEPTH, k)
    assert np.array_equal(got["idx"], idx) and np.array_equal(got["dist"], dst)


@pytest.mark.gpu
@pytest.mark.parametrize("variant", [0, 1])
def test_partial_build_and_merge_on_one_gpu(ctx, variant):
    """rg_knn_depth_partial_dev for parts 0..2 of 3 run one after the other on one GPU, stacked the way the all-to-all
    delivers them, merged by rg_knn_depth_merge_dev: the lists equal rg_knn_search's bit for bit."""
    import torch
    from rosella_b200 import gpu, synth
    n, k, parts = 2500, 30, 3
    _, cov, _ = synth.knn_tables(n, 5, seed=31, flight=variant == 1)
    cov[2000, 1::2] = cov[17, 1::2]
    cov[1999] = cov[18]
    ctx.knn_load(None, cov, variant)
    ctx.set_option("depth_build", 1)
    i1, d1, _ = ctx.knn_search(gpu.GRAPH_DEPTH, k)
    ctx.set_option("depth_build", 0)
    ctx.set_option("sym_block", 224)
    ctx.set_option("sym_chunk", 512)
    kp = ctx.knn_kprime(gpu.GRAPH_DEPTH, k)
    assert kp >= k
    dev = torch.device("cuda", ctx.device)
    pd = torch.empty((parts, n, kp), dtype=torch.float32, device=dev)
    pi = torch.empty((parts, n, kp), dtype=torch.int32, device=dev)
    torch.cuda.synchronize()
    for p in range(parts):
        ctx.knn_depth_partial_dev(k, p, parts, pd[p].data_ptr(), pi[p].data_ptr())
    ctx.sync()
    filled = (pi != -1).sum(dim=2)                                       # every pair is held by exactly one part
    assert int(filled.sum(dim=0).min()) >= min(kp, n - 1)
    for qb, qe in ((0, n), (37, 1201)):
        idx = torch.empty((qe - qb, k), dtype=torch.int32, device=dev)
        dst = torch.empty((qe - qb, k), dtype=torch.float32, device=dev)
        sd, si = pd[:, qb:qe].contiguous(), pi[:, qb:qe].contiguous()
        torch.cuda.synchronize()
        ctx.knn_depth_merge_dev(k, parts, sd.data_ptr(), si.data_ptr(), qb, qe, idx.data_ptr(), dst.data_ptr())
        ctx.sync()
        assert np.array_equal(idx.cpu().numpy().astype(np.uint32), i1[qb:qe]) and np.array_equal(dst.cpu().numpy(), d1[qb:qe])
    i2, d2 = rdist.build_depth_graph_symmetric(ctx, k, n, 0, 1)          # world size 1 through the same host code
    ctx.set_option("sym_block", 0)
    ctx.set_option("sym_chunk", 0)
    assert np.array_equal(i2.cpu().numpy().astype(np.uint32), i1) and np.array_equal(d2.cpu().numpy(), d1)


@pytest.mark.gpu
def test_single_gpu_distributed_path_equals_direct(ctx):
    """build_graph_distributed at world size 1 (no process group needed) equals rg_knn_search."""
    from rosella_b200 import gpu, synth
    tnf, cov, _ = synth.knn_tables(700, 4, seed=8)
    ctx.knn_load(tnf, cov)
    i1, d1, _ = ctx.knn_search(gpu.GRAPH_DEPTH, 9)
    i2, d2 = rdist.build_graph_distributed(ctx, gpu.GRAPH_DEPTH, 9, 700, 0, 1)
    assert np.array_equal(i2.cpu().numpy().astype(np.uint32), i1) and np.array_equal(d2.cpu().numpy(), d1)
