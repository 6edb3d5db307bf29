"""More than one GPU behind the C ABI (rg_multi_*: one caller thread; rg_comm_* + *_dist: one handle per rank), plus the
list cache (rg_knn_refilter) and the row-subset build (rg_knn_build_subset).  Every result must equal the one-GPU
result bit for bit.  The 2-GPU cases skip on a 1-GPU box (`gpurun --gpus 2`, benchmarks/gpu_multi2.sh run them)."""
import threading

import numpy as np
import pytest

import oracle
from rosella_b200 import dist as rdist
from rosella_b200 import gpu, synth


def n_gpus():
    import torch
    return torch.cuda.device_count()


def test_comm_shard_matches_the_python_model():
    for n in (1, 31, 32, 33, 1000, 100_000, 1_000_003):
        for world in (1, 2, 3, 8):
            cover = []
            for r in range(world):
                assert gpu.comm_shard(n, r, world) == rdist.shard_range(n, r, world)
                cover.append(gpu.comm_shard(n, r, world))
            assert cover[0][0] == 0 and cover[-1][1] == n and all(a[1] == b[0] for a, b in zip(cover, cover[1:]))


@pytest.mark.gpu
def test_multi_with_one_device_equals_the_single_handle(ctx):
    tnf, cov, _ = synth.knn_tables(2600, 5, seed=71)
    (ti, td), (di, dd) = ctx.knn_build(tnf, cov, 30)
    seq, offs = synth.contigs(400_000, threads=2)
    c, nk, f = ctx.tnf_count_ascii(seq, offs)
    with gpu.Multi([0]) as m:
        (mti, mtd), (mdi, mdd) = m.knn_build(tnf, cov, 30)
        assert np.array_equal(mti, ti) and np.array_equal(mtd, td) and np.array_equal(mdi, di) and np.array_equal(mdd, dd)
        off, nbr, w, disc = m.knn_build_mutual(tnf, cov, 30, 4)
        o2 = ctx.knn_build_mutual(tnf, cov, 30, 4)
        for a, b in zip((off, nbr, w, disc), o2):
            assert np.array_equal(a, b)
        mc, mnk, mf = m.tnf_count_ascii(seq, offs)
        assert np.array_equal(mc, c) and np.array_equal(mnk, nk) and np.array_equal(mf, f, equal_nan=True)
        for a, b in zip(m.sketch_contigs(seq, offs), ctx.sketch_contigs(seq, offs)):
            assert np.array_equal(a, b)


@pytest.mark.gpu
def test_refilter_reuses_the_kept_lists(ctx):
    """prefilter_nodes / the embed retry loop (recover_engine.rs:438-482) change keep_n_edges only."""
    tnf, cov, _ = synth.knn_tables(1500, 3, seed=13)
    k = 12
    first = ctx.knn_build_mutual(tnf, cov, k, 0)
    launches0 = ctx.timings()["kernel_launches"]
    for keep in (3, 12, 0):
        again = ctx.knn_refilter(keep, k)
        want = ctx.knn_build_mutual(tnf, cov, k, keep) if keep else first
        for a, b in zip(again, want):
            assert np.array_equal(a, b)
    ctx.knn_build_mutual(tnf, cov, k, 0)
    l0 = ctx.timings()["kernel_launches"]
    ctx.knn_refilter(5, k)
    assert ctx.timings()["kernel_launches"] - l0 <= 4 < l0 - launches0      # filter kernels only, no search
    ctx.knn_load(tnf, cov)                                                   # a new load drops the kept lists
    with pytest.raises(gpu.RosellaGpuError, match="no kNN lists"):
        ctx.knn_refilter(5, k)


@pytest.mark.gpu
def test_build_subset_equals_a_build_of_the_gathered_rows(ctx):
    """evaluate_subset (recover_engine.rs:410-435): node numbers are positions in row_ids."""
    tnf, cov, _ = synth.knn_tables(3000, 4, seed=17)
    rng = np.random.default_rng(3)
    ids = rng.permutation(3000)[:700].astype(np.uint64)
    (ti, td), (di, dd) = ctx.knn_build_subset(tnf, cov, ids, 15)
    (wti, wtd), (wdi, wdd) = ctx.knn_build(tnf[ids], cov[ids], 15)
    assert np.array_equal(ti, wti) and np.array_equal(td, wtd) and np.array_equal(di, wdi) and np.array_equal(dd, wdd)
    oi, od, _ = oracle.knn(cov[ids], oracle.DEPTH, 15)
    assert (di == oi).mean() > 0.99
    with pytest.raises(gpu.RosellaGpuError, match="outside"):
        ctx.knn_build_subset(tnf, cov, np.array([1, 3000], np.uint64), 5)


@pytest.mark.gpu
@pytest.mark.skipif("n_gpus() < 2")
@pytest.mark.parametrize("n,s,k", [(6000, 6, 30), (20_011, 3, 50)])
def test_multi_two_devices_equal_one(ctx, n, s, k):
    tnf, cov, _ = synth.knn_tables(n, s, seed=5)
    cov[n - 7] = cov[11]                           # duplicates and equal variances across ranks
    cov[n - 9, 1::2] = cov[40, 1::2]
    (ti, td), (di, dd) = ctx.knn_build(tnf, cov, k)
    with gpu.Multi([0, 1]) as m:
        for h in (m.handle(0), m.handle(1)):       # small blocks: many block pairs, both ranks own several
            h.set_option("sym_block", 512); h.set_option("sym_chunk", 1024)
        (mti, mtd), (mdi, mdd) = m.knn_build(tnf, cov, k)
        assert np.array_equal(mdi, di) and np.array_equal(mdd, dd)
        assert np.array_equal(mti, ti) and np.array_equal(mtd, td)
        t0, t1 = m.handle(0).dist_timings(), m.handle(1).dist_timings()
        assert t0["load_bcast_ms"] > 0 and t1["load_h2d_ms"] < t0["load_h2d_ms"] + 1e-3
        hashes, offs = synth.sketches(n, seed=4, mean_size=30, share_frac=0.2)
        ctx.knn_load(None, cov)
        ctx.knn_load_sketches(hashes, offs, 0.1)
        si, sd, _ = ctx.knn_search(gpu.GRAPH_DEPTH, k)
        (_, _), (msi, msd) = m.knn_build_sketched(tnf, cov, hashes, offs, k)
        assert np.array_equal(msi, si) and np.array_equal(msd, sd) and not np.array_equal(msi, di)
        got = m.knn_build_mutual(tnf, cov, k, 6)
        want = ctx.knn_build_mutual(tnf, cov, k, 6)
        for a, b in zip(got, want):
            assert np.array_equal(a, b)
        seq, offs = synth.contigs(3_000_000, threads=2)
        c, nk, f = ctx.tnf_count_ascii(seq, offs)
        mc, mnk, mf = m.tnf_count_ascii(seq, offs)
        assert np.array_equal(mc, c) and np.array_equal(mnk, nk) and np.array_equal(mf, f, equal_nan=True)
        for a, b in zip(m.sketch_contigs(seq, offs), ctx.sketch_contigs(seq, offs)):      # per-device ranges, same sketches
            assert np.array_equal(a, b)
    # one failing rank must fail the call on every rank instead of hanging the others
    bad = cov.copy(); bad[5, 1] = np.nan
    with gpu.Multi([0, 1]) as m:
        with pytest.raises(gpu.RosellaGpuError, match="non-finite"):
            m.knn_build(tnf, bad, k)


@pytest.mark.gpu
@pytest.mark.skipif("n_gpus() < 2")
@pytest.mark.parametrize("n", [20, 50, 700])
def test_multi_two_devices_tiny_tables(ctx, n):
    """Fewer rows than one 32-row tile per rank (rank 1 owns nothing at n = 20), fewer rows than k."""
    tnf, cov, _ = synth.knn_tables(n, 3, seed=n)
    (ti, td), (di, dd) = ctx.knn_build(tnf, cov, 30)
    with gpu.Multi([0, 1]) as m:
        (mti, mtd), (mdi, mdd) = m.knn_build(tnf, cov, 30)
        assert np.array_equal(mti, ti) and np.array_equal(mtd, td) and np.array_equal(mdi, di) and np.array_equal(mdd, dd)
        got = m.knn_build_mutual(tnf, cov, 30, 3)
        for a, b in zip(got, ctx.knn_build_mutual(tnf, cov, 30, 3)):
            assert np.array_equal(a, b)


@pytest.mark.gpu
@pytest.mark.skipif("n_gpus() < 2")
def test_comm_two_ranks_in_two_threads(ctx):
    """The one-process-per-GPU flavour (rg_comm_init + rg_knn_load_dist + rg_knn_build_dist), driven here by two
    threads with one handle each: rank 1 never sees the host tables."""
    n, s, k = 9000, 4, 30
    tnf, cov, _ = synth.knn_tables(n, s, seed=29)
    ctx.knn_load(tnf, cov)
    want = {w: ctx.knn_search(w, k)[:2] for w in (gpu.GRAPH_DEPTH, gpu.GRAPH_TNF, gpu.GRAPH_TNF_EUCLID)}
    uid = gpu.comm_unique_id()
    out, err = {}, []

    def worker(rank):
        try:
            c = gpu.Context(rank)
            c.comm_init(rank, 2, uid)
            c.knn_load_dist(tnf if rank == 0 else None, cov if rank == 0 else None, n, 136, s, gpu.VARIANT_RUST, root=0)
            for w in want:
                out[(rank, w, True)] = c.knn_build_dist(w, k, gather=True)
            out[(rank, "slab")] = c.knn_build_dist(gpu.GRAPH_DEPTH, k, gather=False)
            c.close()
        except Exception as e:  # noqa: BLE001
            err.append((rank, e))

    th = [threading.Thread(target=worker, args=(r,)) for r in range(2)]
    [t.start() for t in th]
    [t.join(timeout=300) for t in th]
    assert not err, err
    for rank in range(2):
        for w in want:
            assert np.array_equal(out[(rank, w, True)][0], want[w][0]) and np.array_equal(out[(rank, w, True)][1], want[w][1])
        b, e = gpu.comm_shard(n, rank, 2)
        assert np.array_equal(out[(rank, "slab")][0][b:e], want[gpu.GRAPH_DEPTH][0][b:e])
