"""CPU oracle for Rosella's hot path — TEST INFRASTRUCTURE ONLY.

Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s ``cpu_baseline`` /
``--impl reference`` legs may import this package.  ``rosella_b200`` never does.
See ``oracle/oracle.h`` for the parity status of each function.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB_PATH = os.path.join(_HERE, "liboracle.so")
NBINS = 136
RUST, FLIGHT = 0, 1
TNF, DEPTH, TNF_EUCLID = 0, 1, 2


def build(force: bool = False) -> str:
    srcs = [os.path.join(_HERE, f) for f in ("oracle_tnf.c", "oracle_metrics.c", "oracle_knn.c", "oracle_sketch.c", "oracle_minhash.c", "oracle.h")]
    if force or not os.path.exists(_LIB_PATH) or any(
        os.path.getmtime(s) > os.path.getmtime(_LIB_PATH) for s in srcs
    ):
        subprocess.check_call(["make", "-C", _HERE, "-s", "liboracle.so"])
    return _LIB_PATH


_lib = None


def lib() -> C.CDLL:
    global _lib
    if _lib is None:
        if not os.path.exists(_LIB_PATH):
            build()
        _lib = C.CDLL(_LIB_PATH)
        _lib.orc_normalize.restype = C.c_size_t
        _lib.orc_sketch_max_hash.restype = C.c_uint64
        _lib.orc_sketch_record.restype = C.c_uint64
        for f in ("orc_kmer_correlation", "orc_rho_flight", "orc_tnf_euclidean", "orc_metabat", "orc_min_jaccard", "orc_silhouette"):
            getattr(_lib, f).restype = C.c_double
        for f in ("orc_tnf_distance_f32", "orc_depth_distance_f32", "orc_sketch_distance_f32", "orc_depth_distance_sketch_f32"):
            getattr(_lib, f).restype = C.c_float
    return _lib


def _p(a, t):
    return a.ctypes.data_as(C.POINTER(t))


def canonical_kmers() -> list[str]:
    buf = np.zeros(NBINS * 4, np.uint8)
    lib().orc_canonical_kmers(_p(buf, C.c_uint8))
    return [bytes(buf[4 * i:4 * i + 4]).decode() for i in range(NBINS)]


def code_to_bin() -> np.ndarray:
    lut = np.zeros(256, np.uint8)
    lib().orc_code_to_bin(_p(lut, C.c_uint8))
    return lut


def normalize(seq: bytes) -> bytes:
    a = np.frombuffer(seq, np.uint8)
    out = np.zeros(max(len(a), 1), np.uint8)
    n = lib().orc_normalize(_p(a, C.c_uint8), C.c_size_t(len(a)), _p(out, C.c_uint8))
    return bytes(out[:n])


def tnf_count(seq: np.ndarray, offsets: np.ndarray, fast: bool = False, threads: int = 0):
    """-> (counts u32[n,136], n_kmers u64[n], freqs f64[n,136])"""
    seq = np.ascontiguousarray(seq, np.uint8)
    offsets = np.ascontiguousarray(offsets, np.uint64)
    n = len(offsets) - 1
    counts = np.zeros((n, NBINS), np.uint32)
    nk = np.zeros(n, np.uint64)
    freqs = np.zeros((n, NBINS), np.float64)
    args = [_p(seq, C.c_uint8), _p(offsets, C.c_uint64), C.c_uint64(n),
            _p(counts, C.c_uint32), _p(nk, C.c_uint64), _p(freqs, C.c_double)]
    if fast:
        lib().orc_tnf_count_fast(*args, C.c_int(threads))
    else:
        lib().orc_tnf_count(*args)
    return counts, nk, freqs


def clr(x: np.ndarray, variant: int = RUST) -> np.ndarray:
    x = np.ascontiguousarray(x, np.float64)
    out = np.empty_like(x)
    f = lib().orc_clr_flight if variant == FLIGHT else lib().orc_clr
    f(_p(x, C.c_double), _p(out, C.c_double), C.c_uint64(x.shape[0]), C.c_uint32(x.shape[1]))
    return out


def kmer_correlation(x, y) -> float:
    x = np.ascontiguousarray(x, np.float64); y = np.ascontiguousarray(y, np.float64)
    return lib().orc_kmer_correlation(_p(x, C.c_double), _p(y, C.c_double), C.c_uint32(len(x)))


def rho_flight(x, y) -> float:
    x = np.ascontiguousarray(x, np.float64); y = np.ascontiguousarray(y, np.float64)
    return lib().orc_rho_flight(_p(x, C.c_double), _p(y, C.c_double), C.c_uint32(len(x)))


def tnf_euclidean(x, y) -> float:
    x = np.ascontiguousarray(x, np.float64); y = np.ascontiguousarray(y, np.float64)
    return lib().orc_tnf_euclidean(_p(x, C.c_double), _p(y, C.c_double), C.c_uint32(len(x)))


def metabat(a, b, variant: int = RUST) -> float:
    a = np.ascontiguousarray(a, np.float64); b = np.ascontiguousarray(b, np.float64)
    return lib().orc_metabat(_p(a, C.c_double), _p(b, C.c_double), C.c_uint32(len(a) // 2), C.c_int(variant))


def knn(data: np.ndarray, which: int, k: int, variant: int = RUST, qrows=None, threads: int = 0):
    """Exhaustive exact top-k.  -> (idx u32[nq,k], dist f32[nq,k], dist64 f64[nq,k])"""
    data = np.ascontiguousarray(data, np.float64)
    n, w = data.shape
    if qrows is None:
        nq, qp = n, None
    else:
        qrows = np.ascontiguousarray(qrows, np.uint64)
        nq, qp = len(qrows), _p(qrows, C.c_uint64)
    idx = np.zeros((nq, k), np.uint32)
    dist = np.zeros((nq, k), np.float32)
    d64 = np.zeros((nq, k), np.float64)
    lib().orc_knn(_p(data, C.c_double), C.c_uint64(n), C.c_uint32(w), C.c_int(which), C.c_int(variant),
                  qp, C.c_uint64(nq), C.c_uint32(k), _p(idx, C.c_uint32), _p(dist, C.c_float),
                  _p(d64, C.c_double), C.c_int(threads))
    return idx, dist, d64


def mutual_knn(idx: np.ndarray, keep_n_edges: int) -> np.ndarray:
    idx = np.ascontiguousarray(idx, np.uint32)
    keep = np.zeros(idx.shape, np.uint8)
    lib().orc_mutual_knn(_p(idx, C.c_uint32), C.c_uint64(idx.shape[0]), C.c_uint32(idx.shape[1]),
                         C.c_uint32(keep_n_edges), _p(keep, C.c_uint8))
    return keep


def intersect(base: np.ndarray, other: np.ndarray, keep_n_edges: int) -> np.ndarray:
    base = np.ascontiguousarray(base, np.uint32); other = np.ascontiguousarray(other, np.uint32)
    keep = np.zeros(base.shape, np.uint8)
    lib().orc_intersect(_p(base, C.c_uint32), _p(other, C.c_uint32), C.c_uint64(base.shape[0]),
                        C.c_uint32(base.shape[1]), C.c_uint32(keep_n_edges), _p(keep, C.c_uint8))
    return keep


def build_mutual_kgraph(depth_idx, depth_dist, tnf_idx, keep_n_edges: int):
    """EmbedderEngine::build_mutual_kgraph's post-processing (embedder.rs:68-84) with the filters applied:
    intersect(depth, tnf, max(keep_n_edges, 5)) then mutual_knn(., keep_n_edges) on the filtered lists.
    -> (offsets u64[n+1], neighbours u32, weights f32, disconnected u32)"""
    PAD = 0xFFFFFFFF
    depth_idx = np.ascontiguousarray(depth_idx, np.uint32)
    n, k = depth_idx.shape
    keep1 = intersect(depth_idx, tnf_idx, max(keep_n_edges, 5)).astype(bool)
    f_idx = np.full((n, k), PAD, np.uint32)                 # the filtered lists, compacted to the left
    f_dist = np.zeros((n, k), np.float32)
    for i in range(n):
        m = keep1[i]
        c = int(m.sum())
        f_idx[i, :c] = depth_idx[i][m]
        f_dist[i, :c] = depth_dist[i][m]
    keep2 = mutual_knn(f_idx, keep_n_edges).astype(bool)
    counts = keep2.sum(axis=1).astype(np.uint64)
    offsets = np.zeros(n + 1, np.uint64)
    offsets[1:] = np.cumsum(counts)
    return offsets, f_idx[keep2], f_dist[keep2], np.where(counts == 0)[0].astype(np.uint32)


def murmur3_x64_128(key: bytes, seed: int = 0) -> tuple[int, int]:
    """MurmurHash3_x64_128 restated (oracle_minhash.c); finch's hash is the first half."""
    buf = np.frombuffer(bytes(key), np.uint8) if len(key) else np.zeros(1, np.uint8)
    out = np.zeros(2, np.uint64)
    lib().orc_murmur3_x64_128(_p(buf, C.c_uint8), C.c_uint64(len(key)), C.c_uint64(seed), _p(out, C.c_uint64))
    return int(out[0]), int(out[1])


def sketch_max_hash(scale: float = 0.1) -> int:
    return int(lib().orc_sketch_max_hash(C.c_double(scale)))


def sketch_contigs(seq, offsets, k: int = 31, scale: float = 0.1, seed: int = 42):
    """sketch_contigs (contig_sketcher.rs:47-92) -> (hashes u64[...], sk_offsets u64[n+1], n_valid_kmers u64[n], norm_len u64[n])."""
    seq = np.ascontiguousarray(seq, np.uint8); offsets = np.ascontiguousarray(offsets, np.uint64)
    n = len(offsets) - 1
    sk_off = np.zeros(n + 1, np.uint64); nk = np.zeros(n, np.uint64); nl = np.zeros(n, np.uint64)
    sp = _p(seq, C.c_uint8) if len(seq) else _p(np.zeros(1, np.uint8), C.c_uint8)
    args = (sp, _p(offsets, C.c_uint64), C.c_uint64(n), C.c_uint32(k), C.c_double(scale), C.c_uint64(seed))
    lib().orc_sketch_contigs(*args, _p(sk_off, C.c_uint64), None, _p(nk, C.c_uint64), _p(nl, C.c_uint64))
    hashes = np.zeros(max(int(sk_off[-1]), 1), np.uint64)
    lib().orc_sketch_contigs(*args, _p(sk_off, C.c_uint64), _p(hashes, C.c_uint64), None, None)
    return hashes[: int(sk_off[-1])], sk_off, nk, nl


def min_jaccard(q, r, scale: float = 0.1) -> float:
    """raw_distance(...).2 of sketch_distances.rs:61-126 for two sorted u64 hash lists."""
    q = np.ascontiguousarray(q, np.uint64); r = np.ascontiguousarray(r, np.uint64)
    return lib().orc_min_jaccard(_p(q, C.c_uint64), C.c_uint64(len(q)), _p(r, C.c_uint64), C.c_uint64(len(r)), C.c_double(scale))


def depth_distance_sketch(a, b, qa, qb, variant: int = RUST, scale: float = 0.1) -> float:
    """DepthDistance::eval (embedder.rs:133-171): metabat x (1 - min_jaccard), f32."""
    a = np.ascontiguousarray(a, np.float64); b = np.ascontiguousarray(b, np.float64)
    qa = np.ascontiguousarray(qa, np.uint64); qb = np.ascontiguousarray(qb, np.uint64)
    return lib().orc_depth_distance_sketch_f32(_p(a, C.c_double), _p(b, C.c_double), C.c_uint32(len(a) // 2), C.c_int(variant),
                                               _p(qa, C.c_uint64), C.c_uint64(len(qa)), _p(qb, C.c_uint64), C.c_uint64(len(qb)),
                                               C.c_double(scale))


def knn_depth_sketch(cov, hashes, offsets, k: int, variant: int = RUST, scale: float = 0.1, qrows=None, threads: int = 0):
    """Exhaustive exact top-k of the DepthDistance graph with the MinHash factor.  -> (idx u32[nq,k], dist f32[nq,k])"""
    cov = np.ascontiguousarray(cov, np.float64)
    hashes = np.ascontiguousarray(hashes, np.uint64); offsets = np.ascontiguousarray(offsets, np.uint64)
    n, w = cov.shape
    if qrows is None:
        nq, qp = n, None
    else:
        qrows = np.ascontiguousarray(qrows, np.uint64)
        nq, qp = len(qrows), _p(qrows, C.c_uint64)
    idx = np.zeros((nq, k), np.uint32)
    dist = np.zeros((nq, k), np.float32)
    lib().orc_knn_depth_sketch(_p(cov, C.c_double), C.c_uint64(n), C.c_uint32(w), C.c_int(variant), _p(hashes, C.c_uint64),
                               _p(offsets, C.c_uint64), C.c_double(scale), qp, C.c_uint64(nq), C.c_uint32(k),
                               _p(idx, C.c_uint32), _p(dist, C.c_float), C.c_int(threads))
    return idx, dist


def condensed_euclidean(data) -> np.ndarray:
    """condensed_pairwise_distance, src/clustering/cluster_utils.rs:7-34."""
    data = np.ascontiguousarray(data, np.float64)
    n, d = data.shape
    out = np.zeros(n * (n - 1) // 2)
    lib().orc_condensed_euclidean(_p(data, C.c_double), C.c_uint64(n), C.c_uint32(d), _p(out, C.c_double))
    return out


def silhouette(dist, labels, n_clusters: int):
    """silhouette_score of cluster_utils.rs:38-120 on a condensed distance vector (a = mean over all other members).
    -> (score, per-point values)"""
    dist = np.ascontiguousarray(dist, np.float64)
    labels = np.ascontiguousarray(labels, np.int32)
    n = len(labels)
    pp = np.zeros(n)
    score = lib().orc_silhouette(_p(dist, C.c_double), C.c_uint64(n), _p(labels, C.c_int32), C.c_uint32(n_clusters), _p(pp, C.c_double))
    return score, pp


def max_threads() -> int:
    return lib().orc_max_threads()
