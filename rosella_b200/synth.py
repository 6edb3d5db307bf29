"""Seeded synthetic inputs of the BASELINE.json shapes.  numpy only."""
from __future__ import annotations

from concurrent.futures import ThreadPoolExecutor

import numpy as np

SEED_KNN = 20261104
SEED_TNF = 20261105
NBINS = 136


def knn_tables(n: int, s: int, seed: int = SEED_KNN, flight: bool = False):
    """-> (tnf f64[n,136] CLR-transformed as the Rust path hands it to TNFDistance,
           cov f64[n,2s] interleaved mean,var, genome id per contig)

    G = max(50, n/2000) latent genomes; per genome a TNF composition ~ Dirichlet(8) and per
    (genome, sample) an abundance ~ LogNormal(1.5, 1.2), absent with probability 0.3.
    Contig length ~ exp(U(ln 2000, ln 500000)); counts ~ Multinomial(L-3, p_genome) with empty bins
    set to 1 (the Rust CLR has no pseudocount: ln 0 would flag the row, kmer_counting.rs:351).
    """
    rng = np.random.default_rng(seed)
    g = max(50, n // 2000)
    comp = rng.dirichlet(8.0 * np.ones(NBINS), g)
    abund = rng.lognormal(1.5, 1.2, (g, s))
    abund[rng.random((g, s)) < 0.3] = 0.0
    genome = rng.integers(0, g, n)
    length = np.rint(np.exp(rng.uniform(np.log(2000.0), np.log(500000.0), n))).astype(np.int64)

    tnf = np.empty((n, NBINS), np.float64)
    step = 65536
    for b in range(0, n, step):
        e = min(n, b + step)
        counts = rng.multinomial(length[b:e] - 3, comp[genome[b:e]]).astype(np.float64)
        if flight:
            x = np.log(counts + 1.0)                      # clr(x + 1), flight/rosella/binning.py:192
            tnf[b:e] = x - x.mean(axis=1, keepdims=True)
        else:
            counts[counts == 0] = 1.0
            freq = counts / counts.sum(axis=1, keepdims=True)
            tnf[b:e] = np.log(freq / freq.mean(axis=1, keepdims=True))   # kmer_counting.rs:340-359
    cov = np.empty((n, 2 * s), np.float64)
    for b in range(0, n, step):
        e = min(n, b + step)
        mean = abund[genome[b:e]] * rng.lognormal(0.0, 0.1, (e - b, s))
        var = mean * (1.0 + 3.0 * rng.random((e - b, s)))
        cov[b:e, 0::2] = mean
        cov[b:e, 1::2] = var
    return tnf, cov, genome


_ASCII = np.frombuffer(b"ACGT", np.uint8)


def _one_contig(args):
    seed, i, length, gc = args
    rng = np.random.default_rng([seed, i])
    r = rng.integers(0, 1 << 16, length, dtype=np.uint16)
    at = (1.0 - gc) / 2.0
    edges = np.array([at, at + gc / 2.0, at + gc], np.float64) * 65536.0
    seq = _ASCII[np.searchsorted(edges, r, side="right")]          # A, C, G, T with P = at, gc/2, gc/2, at
    lower = rng.random(length) < 0.01                              # 1 % lower case
    seq = np.where(lower, seq | 0x20, seq).astype(np.uint8)
    if rng.random() < 0.001 and length > 400:                      # 0.1 % of contigs: one N run of 1..200
        ln = int(rng.integers(1, 201))
        st = int(rng.integers(0, length - ln))
        seq[st:st + ln] = ord("N")
    return seq


def contigs(total_bp: int, seed: int = SEED_TNF, threads: int = 8, min_len: int = 1000, max_len: int = 500000):
    """-> (seq u8[total], offsets u64[n+1]).  Lengths ~ exp(U(ln min, ln max)) until the sum reaches
    total_bp; bases i.i.d. with per-contig GC ~ U(0.3, 0.7)."""
    rng = np.random.default_rng(seed)
    lens = []
    tot = 0
    while tot < total_bp:
        ln = int(round(float(np.exp(rng.uniform(np.log(min_len), np.log(max_len))))))
        ln = min(ln, max(total_bp - tot, min_len))
        lens.append(ln)
        tot += ln
    gcs = rng.uniform(0.3, 0.7, len(lens))
    offsets = np.zeros(len(lens) + 1, np.uint64)
    offsets[1:] = np.cumsum(lens)
    seq = np.empty(int(offsets[-1]), np.uint8)
    jobs = [(seed, i, lens[i], gcs[i]) for i in range(len(lens))]

    def run(job):
        i = job[1]
        seq[int(offsets[i]):int(offsets[i + 1])] = _one_contig(job)

    if threads > 1:
        with ThreadPoolExecutor(threads) as ex:
            list(ex.map(run, jobs, chunksize=16))
    else:
        for j in jobs:
            run(j)
    return seq, offsets


def fasta_wrap(seq: np.ndarray, offsets: np.ndarray, width: int = 60):
    """Insert a newline every `width` bases per record (what needletail hands to normalize for
    multi-line FASTA) -> (bytes u8, offsets u64)."""
    out, offs = [], [0]
    for i in range(len(offsets) - 1):
        rec = seq[int(offsets[i]):int(offsets[i + 1])]
        n = len(rec)
        rows = (n + width - 1) // width
        buf = np.full(n + rows, ord("\n"), np.uint8)
        pos = np.arange(n) + np.arange(n) // width
        buf[pos] = rec
        out.append(buf)
        offs.append(offs[-1] + len(buf))
    return (np.concatenate(out) if out else np.zeros(0, np.uint8)), np.array(offs, np.uint64)


MAX_HASH_SCALE_0_1 = (2 ** 64 - 1) // 10      # finch's max_hash for scale 0.1: u64::MAX / (1/scale) as u64


def sketches(n: int, seed: int = SEED_KNN + 7, mean_size: int = 60, share_frac: float = 0.05, empty_frac: float = 0.002):
    """Synthetic scaled MinHash sketches (what finch hands to SketchDistance, src/sketch/contig_sketcher.rs:47-92):
    per contig a sorted list of distinct u64 hashes, most of them below max_hash(scale = 0.1), a few above (finch pads
    a scaled sketch up to kmers_to_sketch).  A fraction of the contigs copy part of another contig's hashes (shared
    31-mers: repeats, overlapping contig ends), a few sketches are empty.  -> (hashes u64[...], offsets u64[n+1])"""
    rng = np.random.default_rng(seed)
    sizes = rng.poisson(mean_size, n).astype(np.int64)
    sizes[rng.random(n) < empty_frac] = 0
    lists = []
    for i in range(n):
        h = rng.integers(0, MAX_HASH_SCALE_0_1, sizes[i], dtype=np.uint64)
        if sizes[i] and rng.random() < 0.1:                          # padded beyond max_hash
            h = np.concatenate([h, rng.integers(MAX_HASH_SCALE_0_1, 2 ** 63, 3, dtype=np.uint64)])
        lists.append(h)
    donors = rng.integers(0, n, n)
    for i in np.where(rng.random(n) < share_frac)[0]:
        d = lists[donors[i]]
        if len(d) and donors[i] != i:
            take = d[rng.random(len(d)) < rng.uniform(0.05, 0.9)]
            lists[i] = np.concatenate([lists[i], take])
    lists = [np.unique(h) for h in lists]
    offsets = np.zeros(n + 1, np.uint64)
    offsets[1:] = np.cumsum([len(h) for h in lists])
    hashes = np.concatenate(lists) if n else np.zeros(0, np.uint64)
    return hashes.astype(np.uint64), offsets


def knn_tables_parallel(n: int, s: int, seed: int = SEED_KNN, threads: int = 8, flight: bool = False):
    """The generator of knn_tables with one random stream per block of 65536 rows (seeded by [seed, block]) so that the
    blocks can be drawn by several threads: same distributions, a different (still deterministic, thread-count
    independent) table.  Used for the 5M-row shape, where the single-stream generator costs minutes."""
    master = np.random.default_rng(seed)
    g = max(50, n // 2000)
    comp = master.dirichlet(8.0 * np.ones(NBINS), g)
    abund = master.lognormal(1.5, 1.2, (g, s))
    abund[master.random((g, s)) < 0.3] = 0.0
    tnf = np.empty((n, NBINS), np.float64)
    cov = np.empty((n, 2 * s), np.float64)
    genome = np.empty(n, np.int64)
    step = 65536

    def block(bi):
        b, e = bi * step, min(n, (bi + 1) * step)
        rng = np.random.default_rng([seed, bi])
        gen = rng.integers(0, g, e - b)
        genome[b:e] = gen
        length = np.rint(np.exp(rng.uniform(np.log(2000.0), np.log(500000.0), e - b))).astype(np.int64)
        counts = rng.multinomial(length - 3, comp[gen]).astype(np.float64)
        if flight:
            x = np.log(counts + 1.0)
            tnf[b:e] = x - x.mean(axis=1, keepdims=True)
        else:
            counts[counts == 0] = 1.0
            freq = counts / counts.sum(axis=1, keepdims=True)
            tnf[b:e] = np.log(freq / freq.mean(axis=1, keepdims=True))
        mean = abund[gen] * rng.lognormal(0.0, 0.1, (e - b, s))
        var = mean * (1.0 + 3.0 * rng.random((e - b, s)))
        cov[b:e, 0::2] = mean
        cov[b:e, 1::2] = var

    nb = (n + step - 1) // step
    if threads > 1:
        with ThreadPoolExecutor(threads) as ex:
            list(ex.map(block, range(nb)))
    else:
        for bi in range(nb):
            block(bi)
    return tnf, cov, genome
