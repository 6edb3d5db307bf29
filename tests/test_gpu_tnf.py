"""GPU parity: TNF counting through the C ABI vs the oracle — bit-exact (integer work)."""
import os

import numpy as np
import pytest

import oracle
from rosella_b200 import gpu, synth

pytestmark = pytest.mark.gpu


def _check(ctx, seq, offsets, expect=None):
    want = expect if expect is not None else oracle.tnf_count(seq, offsets, fast=True)
    c, nk, f = ctx.tnf_count_ascii(seq, offsets)
    assert np.array_equal(c, want[0]) and np.array_equal(nk, want[1])
    assert np.array_equal(f, want[2], equal_nan=True)          # IEEE division of equal integers: bit-exact
    codes, valid, woff = gpu.pack_host(seq, offsets)
    c2, nk2, f2 = ctx.tnf_count_packed(codes, valid, woff)
    assert np.array_equal(c2, want[0]) and np.array_equal(nk2, want[1])
    assert np.array_equal(f2, want[2], equal_nan=True)


def test_reference_fixtures(ctx, golden_dir):
    g = np.load(os.path.join(golden_dir, "tnf_fixtures.npz"))
    c, nk, f = ctx.tnf_count_ascii(g["seq"], g["offsets"])
    assert np.array_equal(c, g["counts"]) and np.array_equal(nk, g["n_kmers"])
    assert list(c[0, :6]) == [88, 87, 65, 71, 89, 104] and nk[0] == 9997     # SURVEY §4 known answers
    assert np.isnan(f[nk == 0]).all()
    _check(ctx, g["seq"], g["offsets"], oracle.tnf_count(g["seq"], g["offsets"]))


def test_synthetic_contigs_and_fasta_wrapping(ctx):
    seq, offs = synth.contigs(3_000_000, threads=4)
    _check(ctx, seq, offs)
    wrapped, woffs = synth.fasta_wrap(seq[: int(offs[20])], offs[:21])
    _check(ctx, wrapped, woffs)


def test_edge_cases(ctx):
    rng = np.random.default_rng(3)
    recs = [b"", b"A", b"ACG", b"ACGT", b"ACGTN", b"NNNNNNNN", b"\n\n\n", b"acgu" * 50,
            b"ACGT" * 8 + b"N" + b"ACGT" * 8,                  # N exactly at a 32-base word boundary
            b"A" * 31 + b"C" * 33, b"ACGT" * 2048,             # exactly one 8192-base chunk
            b"ACGT" * 2048 + b"A", b"T" * 20000,
            bytes(rng.choice(np.frombuffer(b"ACGTNacgtRY-\n ", np.uint8), 70001))]
    offs = np.zeros(len(recs) + 1, np.uint64)
    offs[1:] = np.cumsum([len(r) for r in recs])
    seq = np.frombuffer(b"".join(recs), np.uint8)
    _check(ctx, seq, offs, oracle.tnf_count(seq, offs))
    # zero records
    c, nk, f = ctx.tnf_count_ascii(np.zeros(0, np.uint8), np.zeros(1, np.uint64))
    assert c.shape == (0, 136) and nk.shape == (0,)


def test_large_properties(ctx):
    """Size-independent checks at a size the faithful oracle would not finish quickly."""
    seq, offs = synth.contigs(200_000_000, threads=8)
    ctx.tnf_load_ascii(seq, offs)
    ctx.tnf_count_resident()
    c, nk, f = ctx.tnf_fetch()
    assert np.array_equal(c.sum(axis=1, dtype=np.uint64), nk)
    lens = (offs[1:] - offs[:-1]).astype(np.int64)
    assert (nk.astype(np.int64) <= np.maximum(lens - 3, 0)).all()
    clean = nk.astype(np.int64) == lens - 3                       # contigs without an N run
    assert clean.mean() > 0.99
    # reverse-complementing a contig leaves canonical counts unchanged
    comp = np.zeros(256, np.uint8); comp[list(b"ACGTNacgt")] = list(b"TGCANtgca")
    i = int(np.argmax(lens))
    rc = comp[seq[int(offs[i]):int(offs[i + 1])]][::-1].copy()
    c2, nk2, _ = ctx.tnf_count_ascii(rc, np.array([0, len(rc)], np.uint64))
    assert np.array_equal(c2[0], c[i]) and nk2[0] == nk[i]
    # packed path agrees with the ASCII path, and both with the fast oracle
    codes, valid, woff = gpu.pack_host(seq, offs)
    c3, nk3, _ = ctx.tnf_count_packed(codes, valid, woff)
    assert np.array_equal(c3, c) and np.array_equal(nk3, nk)
    want = oracle.tnf_count(seq, offs, fast=True)
    assert np.array_equal(c, want[0]) and np.array_equal(nk, want[1])


def test_chunk_boundaries_and_many_short_records(ctx):
    """Second-generation kernel (tnf_count2.cu): a chunk is 896 units of 32 bases (28 672 bases) per CTA step, a
    thread's byte counters take at most 224 windows between flushes, and a CTA walks several records in a row."""
    rng = np.random.default_rng(11)
    alpha = np.frombuffer(b"ACGT", np.uint8)
    recs = [bytes(alpha[rng.integers(0, 4, n)]) for n in (28671, 28672, 28673, 28675, 57344, 57347, 3 * 28672 + 5)]
    recs += [b"T" * 28672 * 2, b"G" * 229, b"AC" * 14336 + b"N" + b"GT" * 7]       # saturating byte counters
    recs += [bytes(alpha[rng.integers(0, 4, int(n))]) for n in rng.integers(0, 300, 3000)]   # record switches inside a CTA
    long = alpha[rng.integers(0, 4, 400_000)].copy()
    for p in rng.integers(0, 400_000 - 300, 40):
        long[p:p + int(rng.integers(1, 200))] = ord("N")
    long[rng.integers(0, 400_000, 4000)] |= 0x20                                      # lower case
    recs.append(bytes(long))
    offs = np.zeros(len(recs) + 1, np.uint64)
    offs[1:] = np.cumsum([len(r) for r in recs])
    seq = np.frombuffer(b"".join(recs), np.uint8)
    _check(ctx, seq, offs)
    # the same records behind an odd prefix: unaligned starts for the 32-byte segments of the ASCII kernel
    for pad in (1, 17, 31):
        seq2 = np.concatenate([np.full(pad, ord("A"), np.uint8), seq])
        offs2 = offs + np.uint64(pad)
        want = oracle.tnf_count(seq, offs, fast=True)
        c, nk, _ = ctx.tnf_count_ascii(seq2, offs2)
        assert np.array_equal(c, want[0]) and np.array_equal(nk, want[1])


def test_every_byte_value_against_the_register_classifier(ctx):
    """The ASCII kernel's fast path decides "all 36 bytes are bases" from `byte & 7` and two PRMT tables
    (tnf_count2.cu, classify4): every one of the 256 byte values, at every position of a 32-byte unit and in the 4
    bytes before it, inside otherwise clean sequence — against the faithful oracle.  Near misses of the test
    ('@' 'B' 'F' '`' 'b' 'f' 'U' 'u', high-bit twins of the bases, control bytes with the same low bits) are all here."""
    rng = np.random.default_rng(23)
    alpha = np.frombuffer(b"ACGTacgtUu", np.uint8)
    recs = []
    for b in range(256):
        r = alpha[rng.integers(0, len(alpha), 32 * 40 + int(rng.integers(0, 32)))].copy()
        r[37 + 33 * np.arange(36)] = b                         # positions 37, 70, 103, ...: every offset mod 32
        recs.append(bytes(r))
    recs.append(bytes(alpha[rng.integers(0, len(alpha), 5000)]))   # clean: all units on the fast path
    offs = np.zeros(len(recs) + 1, np.uint64)
    offs[1:] = np.cumsum([len(r) for r in recs])
    seq = np.frombuffer(b"".join(recs), np.uint8)
    want = oracle.tnf_count(seq, offs)                          # the byte-by-byte restatement
    _check(ctx, seq, offs, want)
    for pad in (3, 29):                                         # other alignments of the records against the units
        seq2 = np.concatenate([np.full(pad, ord("G"), np.uint8), seq])
        offs2 = offs + np.uint64(pad); offs2[0] = 0
        c, nk, _ = ctx.tnf_count_ascii(seq2, offs2)
        w2 = oracle.tnf_count(seq2, offs2)
        assert np.array_equal(c, w2[0]) and np.array_equal(nk, w2[1])
