"""GPU parity: rg_sketch_contigs (sketch_contigs, contig_sketcher.rs:47-92) through the C ABI vs the oracle's restatement
of finch's scaled MinHash — integer work, bit-exact: same offsets, same hashes, same k-mer counts."""
import numpy as np
import pytest

import oracle
from rosella_b200 import gpu, synth

pytestmark = pytest.mark.gpu


def _cat(recs):
    offs = np.zeros(len(recs) + 1, np.uint64)
    offs[1:] = np.cumsum([len(r) for r in recs])
    return np.frombuffer(b"".join(recs), np.uint8), offs


def _check(ctx, seq, offs, k=31, scale=0.1, seed=42):
    want_h, want_o, want_nk, _ = oracle.sketch_contigs(seq, offs, k, scale, seed)
    h, o, nk = ctx.sketch_contigs(seq, offs, k, scale, seed)
    assert np.array_equal(o, want_o), np.flatnonzero(o != want_o)[:5]
    assert np.array_equal(nk, want_nk)
    assert np.array_equal(h, want_h)
    return h, o


def test_synthetic_contigs_and_wrapped_fasta(ctx):
    seq, offs = synth.contigs(3_000_000, threads=4)
    h, o = _check(ctx, seq, offs)
    assert 0.08 < len(h) / len(seq) < 0.12                     # scale 0.1
    wrapped, woffs = synth.fasta_wrap(seq[: int(offs[20])], offs[:21])
    hw, ow = _check(ctx, wrapped, woffs)
    assert np.array_equal(hw, h[: int(o[20])])                 # line breaks are removed by normalize(false)


def test_many_slices_and_records_longer_than_a_slice(ctx):
    """rg_sketch_contigs works through slices of records (256 MiB by default) with the upload of the next slice running
    beside the hashing of the current one on a copy stream, two alternating device buffers.  With 100 kB slices the
    3 Mbp set takes about 40 of them, several records are longer than a slice (a slice of their own), and the result
    must not change."""
    seq, offs = synth.contigs(3_000_000, threads=4, seed=12)
    assert (np.diff(offs.astype(np.int64)) > 100_000).sum() >= 3
    want = ctx.sketch_contigs(seq, offs)
    try:
        for slice_bytes in (100_000, 37_777, 1_000_000):
            ctx.set_option("sketch_slice_bytes", slice_bytes)
            for a, b in zip(ctx.sketch_contigs(seq, offs), want):
                assert np.array_equal(a, b), slice_bytes
    finally:
        ctx.set_option("sketch_slice_bytes", 0)
    wh, wo, wnk, _ = oracle.sketch_contigs(seq, offs)
    assert np.array_equal(want[0], wh) and np.array_equal(want[1], wo) and np.array_equal(want[2], wnk)


def test_edge_cases_and_the_redo_path(ctx):
    rng = np.random.default_rng(31)
    alpha = np.frombuffer(b"ACGT", np.uint8)
    rnd = lambda n: bytes(alpha[rng.integers(0, 4, n)])
    mixed = bytearray(rnd(20000))
    for p in rng.integers(0, 19000, 30):
        mixed[p:p + int(rng.integers(1, 60))] = b"N" * 60
    lower = bytearray(rnd(5000))
    for p in rng.integers(0, 5000, 400):
        lower[p] |= 0x20
    recs = [b"", b"A", rnd(30), rnd(31), rnd(32), rnd(63), rnd(64), rnd(65), rnd(95), rnd(1500),
            b"N" * 3000,                                        # no k-mer at all
            b"ACGT" * 5000,                                     # 2 distinct canonical k-mers: candidates cannot fill the sketch
            b"A" * 40000,                                       # 1 distinct k-mer
            rnd(3000) + b"N" * 9000 + rnd(3000),                # kmers_to_sketch counts the N's: padded past max_hash -> redo
            bytes(mixed), bytes(lower), b"ACGU" * 100 + rnd(200), b"\n".join([rnd(60)] * 50), rnd(100) + b"-~.RYK" + rnd(100),
            bytes(rng.choice(np.frombuffer(b"ACGTNacgtRY-\n ", np.uint8), 30001))]
    seq, offs = _cat(recs)
    h, o = _check(ctx, seq, offs)
    assert o[11] == o[10] and o[12] - o[11] == 2 and o[13] - o[12] == 1
    # the same records behind odd prefixes: other alignments of the 64-byte thread segments
    for pad in (1, 33):
        seq2 = np.concatenate([np.full(pad, ord("C"), np.uint8), seq])
        offs2 = offs + np.uint64(pad); offs2[0] = 0
        _check(ctx, seq2, offs2)
    # zero records
    h0, o0, nk0 = ctx.sketch_contigs(np.zeros(0, np.uint8), np.zeros(1, np.uint64))
    assert len(h0) == 0 and list(o0) == [0] and len(nk0) == 0


@pytest.mark.parametrize("k,scale,seed", [(31, 0.1, 42), (21, 0.1, 42), (32, 0.05, 7), (16, 0.25, 1), (15, 0.5, 42), (8, 1.0, 3), (4, 0.1, 0)])
def test_other_kmer_lengths_scales_and_seeds(ctx, k, scale, seed):
    seq, offs = synth.contigs(400_000, threads=4, seed=77)
    _check(ctx, seq, offs, k, scale, seed)


def test_errors(ctx):
    seq, offs = _cat([b"ACGT" * 100])
    with pytest.raises(gpu.RosellaGpuError, match="kmer_length"):
        ctx.sketch_contigs(seq, offs, 33)
    with pytest.raises(gpu.RosellaGpuError, match="scale"):
        ctx.sketch_contigs(seq, offs, 31, 0.0)
    with pytest.raises(gpu.RosellaGpuError, match="decrease"):
        ctx.sketch_contigs(seq, np.array([0, 200, 100], np.uint64))


def test_gpu_sketches_feed_the_depth_graph(ctx):
    """sketch_contigs -> rg_knn_load_sketches -> DepthDistance graph (embedder.rs:133-171), all on the GPU, against the
    oracle's graph built from the oracle's sketches.  Some contigs share sequence so that the factor matters."""
    seq, offs = synth.contigs(600_000, threads=4, seed=5, min_len=1500, max_len=8000)
    n = len(offs) - 1
    seq = seq.copy()
    rng = np.random.default_rng(2)
    for _ in range(40):                                          # copy a stretch of one contig into another
        a, b = rng.integers(0, n, 2)
        la, lb = int(offs[a + 1] - offs[a]), int(offs[b + 1] - offs[b])
        ln = min(la, lb, 1200)
        seq[int(offs[b]):int(offs[b]) + ln] = seq[int(offs[a]):int(offs[a]) + ln]
    _, cov, _ = synth.knn_tables(n, 3, seed=4)
    h, o, _ = ctx.sketch_contigs(seq, offs)
    want_h, want_o, _, _ = oracle.sketch_contigs(seq, offs)
    assert np.array_equal(o, want_o) and np.array_equal(h, want_h)
    ctx.knn_load(None, cov)
    ctx.knn_load_sketches(h, o, 0.1)
    idx, dist, _ = ctx.knn_search(gpu.GRAPH_DEPTH, 10)
    oi, od = oracle.knn_depth_sketch(cov, want_h, want_o, 10, oracle.RUST, 0.1)
    assert np.all(np.abs(dist - od) <= 2.4e-7 * np.abs(od) + 1e-15)
    assert (idx == oi).mean() > 0.999                            # ties at equal f32 distance may swap
    # the same sketches taken straight from the device (hashes = None): no host round trip, same graph
    ctx.knn_load(None, cov)
    ctx.knn_load_sketches(None, o, 0.1)
    idx2, dist2, _ = ctx.knn_search(gpu.GRAPH_DEPTH, 10)
    assert np.array_equal(idx2, idx) and np.array_equal(dist2, dist)
    with pytest.raises(gpu.RosellaGpuError, match="holds no sketches"):
        ctx.knn_load_sketches(None, np.concatenate([o[:-1], o[-1:] + np.uint64(1)]), 0.1)
    plain_i, _, _ = (ctx.knn_load(None, cov), ctx.knn_search(gpu.GRAPH_DEPTH, 10))[1]
    assert (plain_i != idx).any(axis=1).sum() >= 5               # the planted shared stretches do change the graph
