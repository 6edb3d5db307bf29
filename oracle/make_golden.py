#!/usr/bin/env python
"""Generate tests/golden/*.npz from the reference itself.  Run in the build container only
(needs /root/reference; the GPU box never runs this).

  flight_metrics.npz  outputs of the reference's own numba functions
                      (flight/flight/metrics.py: metabat_distance_nn :445-538, rho :601-645,
                      tnf_euclidean :64-77, coverage_distance :429-443) on seeded inputs.
                      This is what PINS the oracle's flight variant.
  tnf_fixtures.npz    the reference's bundled FASTA inputs (tests/data/ben/random10000.fna and
                      the first records of 1.fna.gz) as bytes + offsets, with counts produced by
                      a literal pure-Python model of kmer_counting.rs:64-109 written here
                      (dict lookup per k-mer, string reverse complement) — an implementation
                      independent of oracle_tnf.c.  The reference ships no expected outputs
                      (SURVEY F3), so for the Rust path this is a cross-check, not a pin.
"""
import gzip
import importlib.util
import os
import sys

import numpy as np

REF = "/root/reference"
OUT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "tests", "golden")


def load_flight_metrics():
    sys.dont_write_bytecode = True
    spec = importlib.util.spec_from_file_location("flight_metrics", f"{REF}/flight/flight/metrics.py")
    m = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(m)
    return m


def coverage_rows(rng, n, s):
    """Rows of interleaved (mean, var) with the edge cases the metric branches on."""
    mean = rng.lognormal(1.5, 1.2, (n, s))
    mean[rng.random((n, s)) < 0.3] = 0.0
    var = mean * (1 + 3 * rng.random((n, s)))
    low = rng.random((n, s)) < 0.15          # low-coverage samples: var clamps to 1 -> equal-variance branch
    mean[low] *= 0.05
    var[low] = mean[low] * 0.5
    cov = np.empty((n, 2 * s))
    cov[:, 0::2] = mean
    cov[:, 1::2] = var
    cov[1] = cov[0]                           # identical rows: every sample skipped (a_mean == b_mean)
    cov[2, 0::2] = cov[3, 0::2]               # equal means, different variances
    cov[4, 1::2] = cov[5, 1::2] + 5e-5        # |dv| < 1e-4 with var > 1 on some samples
    cov[6] = 0.0                              # absent everywhere
    return cov


def make_flight(m):
    rng = np.random.default_rng(20261104)
    out = {}
    for s in (1, 3, 10):
        cov = coverage_rows(rng, 48, s)
        n = len(cov)
        md = np.zeros((n, n)); cd = np.zeros((n, n))
        for i in range(n):
            for j in range(n):
                md[i, j] = m.metabat_distance_nn(cov[i], cov[j])
                cd[i, j] = m.coverage_distance(cov[i], cov[j])
        out[f"cov_s{s}"] = cov
        out[f"metabat_nn_s{s}"] = md
        out[f"coverage_distance_s{s}"] = cd
    tnf = rng.dirichlet(8 * np.ones(136), 40)
    counts = np.stack([rng.multinomial(3000, p) for p in tnf]).astype(np.float64)
    x = np.log(counts + 1.0)
    x -= x.mean(axis=1, keepdims=True)        # clr(x + 1) as flight/rosella/binning.py:192
    n = len(x)
    rho = np.zeros((n, n)); eu = np.zeros((n, n))
    for i in range(n):
        for j in range(n):
            rho[i, j] = m.rho(x[i], x[j])
            eu[i, j] = m.tnf_euclidean(x[i], x[j])
    out["tnf_clr"] = x
    out["rho"] = rho
    out["tnf_euclidean"] = eu
    np.savez_compressed(os.path.join(OUT, "flight_metrics.npz"), **out)
    print("flight_metrics.npz:", {k: v.shape for k, v in out.items()})


def read_fasta(fh):
    names, seqs, cur = [], [], None
    for line in fh:
        if line.startswith(b">"):
            names.append(line[1:].split()[0].decode())
            cur = []
            seqs.append(cur)
        elif cur is not None:
            cur.append(line)                  # keep the newline: needletail hands raw lines to normalize
    return names, [b"".join(s) for s in seqs]


COMP = {ord("A"): "T", ord("C"): "G", ord("G"): "C", ord("T"): "A"}


def model_count(raw: bytes):
    """Literal model of kmer_counting.rs:64-109 + needletail normalize/canonical_kmers."""
    import itertools
    canon = set()
    for t in itertools.product("ACGT", repeat=4):
        k = "".join(t)
        rc = k[::-1].translate(COMP)
        canon.add(min(k, rc))
    index = {k: i for i, k in enumerate(sorted(canon))}
    norm = []
    for ch in raw.decode("latin-1"):
        if ch in "ACGTN-": norm.append(ch)
        elif ch in "acg": norm.append(ch.upper())
        elif ch in "tuU": norm.append("T")
        elif ch in ".~": norm.append("-")
        elif ch in " \t\r\n": continue
        else: norm.append("N")
    norm = "".join(norm)
    rc = norm[::-1].translate(COMP)
    counts = np.zeros(136, np.uint32)
    nk = 0
    L = len(norm)
    for pos in range(L - 3):
        w = norm[pos:pos + 4]
        if any(c not in "ACGT" for c in w):
            continue
        r = rc[L - pos - 4:L - pos]
        k = w if w < r else r
        counts[index[k]] += 1
        nk += 1
    return counts, nk, sorted(canon)


def make_tnf():
    with open(f"{REF}/tests/data/ben/random10000.fna", "rb") as fh:
        names, seqs = read_fasta(fh)
    with gzip.open(f"{REF}/tests/data/ben/1.fna.gz", "rb") as fh:
        n2, s2 = read_fasta(fh)
    names += n2[:200]; seqs += s2[:200]
    # adversarial records built from the fixture's first contig (lower case, IUPAC, N runs, U, gaps, CRLF)
    base = seqs[0][:600]
    adv = [base.lower(), base[:100] + b"NNNNN" + base[100:200] + b"RYKM" + base[200:300],
           base[:50].replace(b"T", b"U") + b"-..~" + base[50:120], base.replace(b"\n", b"\r\n"),
           b"ACG", b"", b"ACGT", b"NNNN", b"acgtn ACGT\tACGT"]
    names += [f"adversarial_{i}" for i in range(len(adv))]; seqs += adv
    counts = np.zeros((len(seqs), 136), np.uint32); nk = np.zeros(len(seqs), np.uint64)
    canon = None
    for i, s in enumerate(seqs):
        counts[i], nk[i], canon = model_count(s)
    offsets = np.zeros(len(seqs) + 1, np.uint64)
    offsets[1:] = np.cumsum([len(s) for s in seqs])
    np.savez_compressed(os.path.join(OUT, "tnf_fixtures.npz"), seq=np.frombuffer(b"".join(seqs), np.uint8),
                        offsets=offsets, names=np.array(names), counts=counts, n_kmers=nk,
                        canonical=np.array(canon))
    print("tnf_fixtures.npz:", len(seqs), "records,", int(offsets[-1]), "bytes; first bins", counts[0, :6], nk[:2])


if __name__ == "__main__":
    os.makedirs(OUT, exist_ok=True)
    make_tnf()
    make_flight(load_flight_metrics())
