"""Thin Python harness over the C ABI (tests, bench, smoke).  One `Context` = one rg_handle = one GPU.
All compute happens in librosella_gpu.so; nothing here falls back to the CPU."""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import _lib
from ._lib import (GRAPH_DEPTH, GRAPH_TNF, GRAPH_TNF_EUCLID, NBINS, VARIANT_FLIGHT, VARIANT_RUST, DistTimings, RosellaGpuError,
                   Timings)

__all__ = ["Context", "Multi", "comm_unique_id", "comm_shard", "canonical_kmers", "pack_host", "GRAPH_TNF", "GRAPH_DEPTH", "GRAPH_TNF_EUCLID", "VARIANT_RUST", "VARIANT_FLIGHT"]


def canonical_kmers() -> list[str]:
    buf = np.zeros(NBINS * 4, np.uint8)
    rc = _lib.load().rg_tnf_canonical_kmers(_lib.ptr(buf, C.c_uint8))
    if rc:
        raise RosellaGpuError(rc, "rg_tnf_canonical_kmers")
    return [bytes(buf[4 * i:4 * i + 4]).decode() for i in range(NBINS)]


def pack_host(seq: np.ndarray, offsets: np.ndarray, threads: int = 0):
    """needletail normalize(false) + 2-bit/validity packing on the host -> (codes u64, valid u32, word_offsets u64)."""
    lib = _lib.load()
    seq = np.ascontiguousarray(seq, np.uint8)
    offsets = np.ascontiguousarray(offsets, np.uint64)
    n = len(offsets) - 1
    words = lib.rg_tnf_packed_words(_lib.ptr(seq, C.c_uint8), _lib.ptr(offsets, C.c_uint64), n)
    codes = np.zeros(max(words, 1), np.uint64)
    valid = np.zeros(max(words, 1), np.uint32)
    woff = np.zeros(n + 1, np.uint64)
    rc = lib.rg_tnf_pack_host(_lib.ptr(seq, C.c_uint8), _lib.ptr(offsets, C.c_uint64), n, _lib.ptr(codes, C.c_uint64),
                              _lib.ptr(valid, C.c_uint32), _lib.ptr(woff, C.c_uint64), threads)
    if rc:
        raise RosellaGpuError(rc, "rg_tnf_pack_host")
    return codes[:words], valid[:words], woff


def comm_unique_id() -> np.ndarray:
    """rg_comm_unique_id: 128 bytes that rank 0 hands to the other ranks (file, MPI, torch.distributed ...)."""
    buf = np.zeros(128, np.uint8)
    rc = _lib.load().rg_comm_unique_id(_lib.ptr(buf, C.c_uint8))
    if rc:
        raise RosellaGpuError(rc, (_lib.load().rg_last_error(None) or b"").decode())
    return buf


def comm_shard(n: int, rank: int, world: int) -> tuple[int, int]:
    b, e = C.c_uint64(0), C.c_uint64(0)
    rc = _lib.load().rg_comm_shard(n, rank, world, C.byref(b), C.byref(e))
    if rc:
        raise RosellaGpuError(rc, "rg_comm_shard")
    return int(b.value), int(e.value)


class Context:
    def __init__(self, device: int = 0, _borrowed=None):
        self.lib = _lib.load()
        self._owned = _borrowed is None
        if _borrowed is not None:        # a handle that belongs to an rg_multi (options / timings only)
            self.h = C.c_void_p(_borrowed)
            self.device = device
            return
        h = C.c_void_p()
        rc = self.lib.rg_create(device, C.byref(h))
        if rc:
            raise RosellaGpuError(rc, (self.lib.rg_last_error(None) or b"").decode())
        self.h = h
        self.device = device

    def close(self):
        if getattr(self, "h", None) and self._owned:
            self.lib.rg_destroy(self.h)
        self.h = None

    __del__ = close

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    def _chk(self, rc: int):
        if rc:
            raise RosellaGpuError(rc, (self.lib.rg_last_error(self.h) or b"").decode())

    def set_option(self, key: str, value: int):
        """rg_set_option: depth_build, sym_block, sym_chunk, tnf_select_impl, certify, kprime_depth, kprime_tnf, sketch_slice_bytes."""
        self._chk(self.lib.rg_set_option(self.h, key.encode(), int(value)))

    def sync(self):
        self._chk(self.lib.rg_sync(self.h))

    @property
    def stream(self) -> int:
        return self.lib.rg_stream(self.h) or 0

    def timings(self) -> dict:
        t = Timings()
        self._chk(self.lib.rg_get_timings(self.h, C.byref(t)))
        return t.as_dict()

    # ------------------------------------------------------------------ TNF
    def _tnf_out(self, n):
        return np.zeros((n, NBINS), np.uint32), np.zeros(n, np.uint64), np.zeros((n, NBINS), np.float64)

    def tnf_count_ascii(self, seq, offsets, out=None):
        """out: optional (counts u32[n,136], n_kmers u64[n], freqs f64[n,136]) host arrays to fill (e.g. pinned, reused)."""
        seq = np.ascontiguousarray(seq, np.uint8)
        offsets = np.ascontiguousarray(offsets, np.uint64)
        n = len(offsets) - 1
        counts, nk, freqs = out if out is not None else self._tnf_out(n)
        self._chk(self.lib.rg_tnf_count_ascii(self.h, _lib.ptr(seq, C.c_uint8), _lib.ptr(offsets, C.c_uint64), n,
                                              _lib.ptr(counts, C.c_uint32), _lib.ptr(nk, C.c_uint64),
                                              _lib.ptr(freqs, C.c_double)))
        return counts, nk, freqs

    def tnf_count_packed(self, codes, valid, word_offsets):
        codes = np.ascontiguousarray(codes, np.uint64)
        valid = np.ascontiguousarray(valid, np.uint32)
        word_offsets = np.ascontiguousarray(word_offsets, np.uint64)
        n = len(word_offsets) - 1
        counts, nk, freqs = self._tnf_out(n)
        self._chk(self.lib.rg_tnf_count_packed(self.h, _lib.ptr(codes, C.c_uint64), _lib.ptr(valid, C.c_uint32),
                                               _lib.ptr(word_offsets, C.c_uint64), n, _lib.ptr(counts, C.c_uint32),
                                               _lib.ptr(nk, C.c_uint64), _lib.ptr(freqs, C.c_double)))
        return counts, nk, freqs

    def tnf_load_packed(self, codes, valid, word_offsets):
        codes = np.ascontiguousarray(codes, np.uint64)
        valid = np.ascontiguousarray(valid, np.uint32)
        word_offsets = np.ascontiguousarray(word_offsets, np.uint64)
        self._tnf_n = len(word_offsets) - 1
        self._chk(self.lib.rg_tnf_load_packed(self.h, _lib.ptr(codes, C.c_uint64), _lib.ptr(valid, C.c_uint32),
                                              _lib.ptr(word_offsets, C.c_uint64), self._tnf_n))

    def tnf_load_ascii(self, seq, offsets):
        seq = np.ascontiguousarray(seq, np.uint8)
        offsets = np.ascontiguousarray(offsets, np.uint64)
        self._tnf_n = len(offsets) - 1
        self._chk(self.lib.rg_tnf_load_ascii(self.h, _lib.ptr(seq, C.c_uint8), _lib.ptr(offsets, C.c_uint64), self._tnf_n))

    def tnf_count_resident(self):
        self._chk(self.lib.rg_tnf_count_resident(self.h))

    def tnf_fetch(self):
        counts, nk, freqs = self._tnf_out(self._tnf_n)
        self._chk(self.lib.rg_tnf_fetch(self.h, _lib.ptr(counts, C.c_uint32), _lib.ptr(nk, C.c_uint64),
                                        _lib.ptr(freqs, C.c_double)))
        return counts, nk, freqs

    # ------------------------------------------------------------------ kNN
    def knn_load(self, tnf, cov, variant: int = VARIANT_RUST):
        tnf, cov = _lib.as_f64(tnf), _lib.as_f64(cov)
        n = (tnf if tnf is not None else cov).shape[0]
        d = tnf.shape[1] if tnf is not None else NBINS
        s = cov.shape[1] // 2 if cov is not None else 1
        self._n = n
        self._chk(self.lib.rg_knn_load(self.h, _lib.ptr(tnf, C.c_double), _lib.ptr(cov, C.c_double), n, d, s, variant))

    def knn_load_sketches(self, hashes, offsets, scale: float = 0.1):
        """MinHash sketches of the loaded rows (sorted u64 hash lists) -> RG_GRAPH_DEPTH becomes DepthDistance."""
        offsets = np.ascontiguousarray(offsets, np.uint64)
        hp = None                                    # None: the sketches the last sketch_contigs left on the device
        if hashes is not None:
            hashes = np.ascontiguousarray(hashes, np.uint64)
            hp = _lib.ptr(hashes, C.c_uint64)
        self._chk(self.lib.rg_knn_load_sketches(self.h, hp, _lib.ptr(offsets, C.c_uint64), len(offsets) - 1, scale))

    def sketch_contigs(self, seq, offsets, kmer_length: int = 31, scale: float = 0.1, seed: int = 42):
        """rg_sketch_contigs + rg_sketch_fetch (contig_sketcher.rs:47-92): finch's scaled MinHash per record
        -> (hashes u64[...], sketch_offsets u64[n+1], n_valid_kmers u64[n])."""
        seq = np.ascontiguousarray(seq, np.uint8)
        offsets = np.ascontiguousarray(offsets, np.uint64)
        n = len(offsets) - 1
        sk_off = np.zeros(n + 1, np.uint64)
        nk = np.zeros(max(n, 1), np.uint64)
        total = C.c_uint64(0)
        sp = _lib.ptr(seq, C.c_uint8) if len(seq) else None
        self._chk(self.lib.rg_sketch_contigs(self.h, sp, _lib.ptr(offsets, C.c_uint64), n, kmer_length, scale, seed,
                                             _lib.ptr(sk_off, C.c_uint64), _lib.ptr(nk, C.c_uint64), C.byref(total)))
        hashes = np.zeros(max(total.value, 1), np.uint64)
        self._chk(self.lib.rg_sketch_fetch(self.h, _lib.ptr(hashes, C.c_uint64), len(hashes)))
        return hashes[: total.value], sk_off, nk[:n]

    def knn_search(self, which: int, k: int, q_begin: int = 0, q_end: int | None = None):
        q_end = self._n if q_end is None else q_end
        nq = max(q_end - q_begin, 0)
        idx = np.zeros((nq, k), np.uint32)
        dist = np.zeros((nq, k), np.float32)
        nv = np.zeros(nq, np.uint32)
        self._chk(self.lib.rg_knn_search(self.h, which, k, q_begin, q_end, _lib.ptr(idx, C.c_uint32),
                                         _lib.ptr(dist, C.c_float), _lib.ptr(nv, C.c_uint32)))
        return idx, dist, nv

    def knn_search_dev(self, which: int, k: int, q_begin: int, q_end: int, idx_ptr: int, dist_ptr: int,
                       nvalid_ptr: int = 0):
        """Outputs are raw device addresses (e.g. torch tensor .data_ptr()); asynchronous on self.stream."""
        self._chk(self.lib.rg_knn_search_dev(self.h, which, k, q_begin, q_end, idx_ptr, dist_ptr, nvalid_ptr or None))

    def knn_build_mutual(self, tnf, cov, k: int, keep_n_edges: int, variant: int = VARIANT_RUST):
        """build_mutual_kgraph with the filters applied -> (offsets u64[n+1], neighbours u32[E], weights f32[E],
        disconnected u32[...])"""
        tnf, cov = _lib.as_f64(tnf), _lib.as_f64(cov)
        n, d, s = tnf.shape[0], tnf.shape[1], cov.shape[1] // 2
        self._n = n
        off = np.zeros(n + 1, np.uint64)
        nbr, w = np.zeros(n * k, np.uint32), np.zeros(n * k, np.float32)
        disc, nd = np.zeros(n, np.uint32), C.c_uint64(0)
        self._chk(self.lib.rg_knn_build_mutual(self.h, _lib.ptr(tnf, C.c_double), _lib.ptr(cov, C.c_double), n, d, s, k,
                                               variant, keep_n_edges, _lib.ptr(off, C.c_uint64), _lib.ptr(nbr, C.c_uint32),
                                               _lib.ptr(w, C.c_float), C.byref(nd), _lib.ptr(disc, C.c_uint32)))
        e = int(off[-1])
        return off, nbr[:e].copy(), w[:e].copy(), disc[:nd.value].copy()

    def knn_refilter(self, keep_n_edges: int, k: int):
        """intersect + mutual_knn re-applied to the lists of the last knn_build_mutual (no load, no search)."""
        n = self._n
        off = np.zeros(n + 1, np.uint64)
        nbr, w = np.zeros(n * k, np.uint32), np.zeros(n * k, np.float32)
        disc, nd = np.zeros(n, np.uint32), C.c_uint64(0)
        self._chk(self.lib.rg_knn_refilter(self.h, keep_n_edges, _lib.ptr(off, C.c_uint64), _lib.ptr(nbr, C.c_uint32),
                                           _lib.ptr(w, C.c_float), C.byref(nd), _lib.ptr(disc, C.c_uint32)))
        e = int(off[-1])
        return off, nbr[:e].copy(), w[:e].copy(), disc[:nd.value].copy()

    def knn_build_subset(self, tnf, cov, row_ids, k: int, variant: int = VARIANT_RUST, which_mask: int = 3):
        tnf, cov = _lib.as_f64(tnf), _lib.as_f64(cov)
        row_ids = np.ascontiguousarray(row_ids, np.uint64)
        n = (tnf if tnf is not None else cov).shape[0]
        d = tnf.shape[1] if tnf is not None else NBINS
        s = cov.shape[1] // 2 if cov is not None else 1
        m = len(row_ids)
        ti, td = np.zeros((m, k), np.uint32), np.zeros((m, k), np.float32)
        di, dd = np.zeros((m, k), np.uint32), np.zeros((m, k), np.float32)
        self._chk(self.lib.rg_knn_build_subset(self.h, _lib.ptr(tnf, C.c_double), _lib.ptr(cov, C.c_double), n, d, s,
                                               _lib.ptr(row_ids, C.c_uint64), m, k, variant, which_mask,
                                               _lib.ptr(ti, C.c_uint32), _lib.ptr(td, C.c_float),
                                               _lib.ptr(di, C.c_uint32), _lib.ptr(dd, C.c_float)))
        self._n = m
        return (ti, td), (di, dd)

    # ------------------------------------------------------------------ one process per GPU (NCCL inside the library)
    def comm_init(self, rank: int, world: int, uid: np.ndarray):
        uid = np.ascontiguousarray(uid, np.uint8)
        assert uid.size == 128
        self._chk(self.lib.rg_comm_init(self.h, rank, world, _lib.ptr(uid, C.c_uint8)))
        self.rank, self.world = rank, world

    def comm_destroy(self):
        self._chk(self.lib.rg_comm_destroy(self.h))

    def knn_load_dist(self, tnf, cov, n: int, d: int, s: int, variant: int = VARIANT_RUST, root: int = 0):
        """Collective.  Only `root` needs the tables (others may pass None)."""
        tnf, cov = _lib.as_f64(tnf), _lib.as_f64(cov)
        self._n = n
        self._chk(self.lib.rg_knn_load_dist(self.h, _lib.ptr(tnf, C.c_double), _lib.ptr(cov, C.c_double), n, d, s, variant, root))

    def knn_build_dist_dev(self, which: int, k: int, gather: bool, idx_ptr: int, dist_ptr: int):
        """Collective, asynchronous on self.stream; idx / dist are device arrays [n][k]."""
        self._chk(self.lib.rg_knn_build_dist_dev(self.h, which, k, int(gather), idx_ptr, dist_ptr))

    def knn_build_dist(self, which: int, k: int, gather: bool = True, out=None):
        """Collective -> (idx u32 [n,k], dist f32 [n,k]) host arrays; gather=False fills only this rank's rows."""
        if out is not None:
            idx, dist = out
            assert idx.shape == (self._n, k) and idx.dtype == np.uint32 and dist.dtype == np.float32
        else:
            idx = np.zeros((self._n, k), np.uint32)
            dist = np.zeros((self._n, k), np.float32)
        self._chk(self.lib.rg_knn_build_dist(self.h, which, k, int(gather), _lib.ptr(idx, C.c_uint32), _lib.ptr(dist, C.c_float)))
        return idx, dist

    def dist_timings(self, which: int | None = None) -> dict:
        """Phase times of the last distributed build (of graph `which` when given)."""
        t = DistTimings()
        if which is None:
            self._chk(self.lib.rg_get_dist_timings(self.h, C.byref(t)))
        else:
            self._chk(self.lib.rg_get_dist_timings_of(self.h, which, C.byref(t)))
        return t.as_dict()

    def knn_kprime(self, which: int, k: int) -> int:
        return int(self.lib.rg_knn_kprime(self.h, which, k))

    def knn_depth_partial_dev(self, k: int, part: int, n_parts: int, cand_dist_ptr: int, cand_idx_ptr: int):
        """This part's share of the symmetric whole-graph coverage build: partial K' lists for every row, device
        arrays [n][K'] (K' = knn_kprime(GRAPH_DEPTH, k)); asynchronous on self.stream."""
        self._chk(self.lib.rg_knn_depth_partial_dev(self.h, k, part, n_parts, cand_dist_ptr, cand_idx_ptr))

    def knn_depth_merge_dev(self, k: int, n_parts: int, parts_dist_ptr: int, parts_idx_ptr: int, q_begin: int, q_end: int,
                            idx_ptr: int, dist_ptr: int, nvalid_ptr: int = 0):
        """Merge [n_parts][q_end - q_begin][K'] partial lists and finish them like knn_search_dev."""
        self._chk(self.lib.rg_knn_depth_merge_dev(self.h, k, n_parts, parts_dist_ptr, parts_idx_ptr, q_begin, q_end,
                                                  idx_ptr, dist_ptr, nvalid_ptr or None))

    def knn_build(self, tnf, cov, k: int, variant: int = VARIANT_RUST, which_mask: int = 3):
        tnf, cov = _lib.as_f64(tnf), _lib.as_f64(cov)
        n = (tnf if tnf is not None else cov).shape[0]
        d = tnf.shape[1] if tnf is not None else NBINS
        s = cov.shape[1] // 2 if cov is not None else 1
        ti, td = np.zeros((n, k), np.uint32), np.zeros((n, k), np.float32)
        di, dd = np.zeros((n, k), np.uint32), np.zeros((n, k), np.float32)
        self._chk(self.lib.rg_knn_build(self.h, _lib.ptr(tnf, C.c_double), _lib.ptr(cov, C.c_double), n, d, s, k, variant,
                                        which_mask, _lib.ptr(ti, C.c_uint32), _lib.ptr(td, C.c_float),
                                        _lib.ptr(di, C.c_uint32), _lib.ptr(dd, C.c_float)))
        self._n = n
        return (ti, td), (di, dd)

    def knn_flagged_rows(self, cap: int = 1 << 20):
        """Rows of the last search that went through the exhaustive pass -> (rows, forced, kth_dist, select_eps)."""
        rows, dk, eps = np.zeros(cap, np.uint32), np.zeros(cap, np.float32), np.zeros(cap, np.float32)
        cnt = C.c_uint64(0)
        self._chk(self.lib.rg_knn_flagged_rows(self.h, _lib.ptr(rows, C.c_uint32), _lib.ptr(dk, C.c_float),
                                               _lib.ptr(eps, C.c_float), cap, C.byref(cnt)))
        m = min(cnt.value, cap)
        return rows[:m] & 0x7FFFFFFF, (rows[:m] >> 31).astype(bool), dk[:m], eps[:m]

    def rank_stat(self, tnf, cov, lengths, k: int = 0, condensed: bool = True):
        """flight's makeRankStat on the device -> (stat condensed f64 or None, idx u32 [n,k] or None, stat_knn f64 [n,k] or None)"""
        tnf, cov = _lib.as_f64(tnf), _lib.as_f64(cov)
        lengths = np.ascontiguousarray(lengths, np.float64)
        n, d, s = tnf.shape[0], tnf.shape[1], cov.shape[1] // 2
        stat = np.zeros(n * (n - 1) // 2, np.float64) if condensed else None
        idx = np.zeros((n, k), np.uint32) if k else None
        sk = np.zeros((n, k), np.float64) if k else None
        self._chk(self.lib.rg_rank_stat(self.h, _lib.ptr(tnf, C.c_double), _lib.ptr(cov, C.c_double), _lib.ptr(lengths, C.c_double),
                                        n, d, s, k, _lib.ptr(stat, C.c_double), _lib.ptr(idx, C.c_uint32), _lib.ptr(sk, C.c_double)))
        return stat, idx, sk

    def condensed_euclidean(self, data):
        """condensed_pairwise_distance (cluster_utils.rs:7-34) -> f64 [n(n-1)/2]"""
        data = _lib.as_f64(data)
        n, d = data.shape
        out = np.zeros(n * (n - 1) // 2, np.float64)
        self._chk(self.lib.rg_condensed_euclidean(self.h, _lib.ptr(data, C.c_double), n, d, _lib.ptr(out, C.c_double)))
        return out

    def silhouette(self, data, labels, n_clusters: int):
        """silhouette_score (cluster_utils.rs:38-120) from the points -> (score, per-point values)"""
        data = _lib.as_f64(data)
        labels = np.ascontiguousarray(labels, np.int32)
        n, d = data.shape
        score, pp = C.c_double(0.0), np.zeros(n, np.float64)
        self._chk(self.lib.rg_silhouette(self.h, _lib.ptr(data, C.c_double), n, d, _lib.ptr(labels, C.c_int32), n_clusters,
                                         C.byref(score), _lib.ptr(pp, C.c_double)))
        return score.value, pp

    # ------------------------------------------------------------------ graph post-ops
    def graph_mutual(self, idx, keep_n_edges: int):
        idx = np.ascontiguousarray(idx, np.uint32)
        keep = np.zeros(idx.shape, np.uint8)
        self._chk(self.lib.rg_graph_mutual(self.h, _lib.ptr(idx, C.c_uint32), idx.shape[0], idx.shape[1], keep_n_edges,
                                           _lib.ptr(keep, C.c_uint8)))
        return keep

    def graph_mutual_static(self, idx, keep_n_edges: int):
        """rg_graph_mutual_static: the same mask with the device staging in the library's fixed static array
        (no device allocation inside the call; at most 2**20 edges)."""
        idx = np.ascontiguousarray(idx, np.uint32)
        keep = np.zeros(idx.shape, np.uint8)
        self._chk(self.lib.rg_graph_mutual_static(self.h, _lib.ptr(idx, C.c_uint32), idx.shape[0], idx.shape[1],
                                                  keep_n_edges, _lib.ptr(keep, C.c_uint8)))
        return keep

    def graph_intersect(self, base, other, keep_n_edges: int):
        base = np.ascontiguousarray(base, np.uint32)
        other = np.ascontiguousarray(other, np.uint32)
        keep = np.zeros(base.shape, np.uint8)
        self._chk(self.lib.rg_graph_intersect(self.h, _lib.ptr(base, C.c_uint32), _lib.ptr(other, C.c_uint32),
                                              base.shape[0], base.shape[1], keep_n_edges, _lib.ptr(keep, C.c_uint8)))
        return keep


class Multi:
    """rg_multi: n GPUs behind one caller thread (one worker thread + one handle per device inside the library)."""

    def __init__(self, devices):
        self.lib = _lib.load()
        devs = (C.c_int32 * len(devices))(*devices)
        m = C.c_void_p()
        rc = self.lib.rg_multi_create(devs, len(devices), C.byref(m))
        if rc:
            raise RosellaGpuError(rc, (self.lib.rg_multi_last_error(None) or b"").decode())
        self.m = m
        self.devices = list(devices)

    def close(self):
        if getattr(self, "m", None):
            self.lib.rg_multi_destroy(self.m)
            self.m = None

    __del__ = close

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    def _chk(self, rc: int):
        if rc:
            raise RosellaGpuError(rc, (self.lib.rg_multi_last_error(self.m) or b"").decode())

    def handle(self, i: int) -> Context:
        return Context(self.devices[i], _borrowed=self.lib.rg_multi_handle(self.m, i))

    def knn_build_sketched(self, tnf, cov, hashes, offsets, k: int, scale: float = 0.1, variant: int = VARIANT_RUST,
                           which_mask: int = 3):
        tnf, cov = _lib.as_f64(tnf), _lib.as_f64(cov)
        hashes, offsets = np.ascontiguousarray(hashes, np.uint64), np.ascontiguousarray(offsets, np.uint64)
        n, s = cov.shape[0], cov.shape[1] // 2
        d = tnf.shape[1] if tnf is not None else NBINS
        ti, td = np.zeros((n, k), np.uint32), np.zeros((n, k), np.float32)
        di, dd = np.zeros((n, k), np.uint32), np.zeros((n, k), np.float32)
        self._chk(self.lib.rg_multi_knn_build_sketched(self.m, _lib.ptr(tnf, C.c_double), _lib.ptr(cov, C.c_double), n, d, s, k,
                                                       variant, which_mask, _lib.ptr(hashes, C.c_uint64),
                                                       _lib.ptr(offsets, C.c_uint64), scale, _lib.ptr(ti, C.c_uint32),
                                                       _lib.ptr(td, C.c_float), _lib.ptr(di, C.c_uint32), _lib.ptr(dd, C.c_float)))
        return (ti, td), (di, dd)

    def knn_build(self, tnf, cov, k: int, variant: int = VARIANT_RUST, which_mask: int = 3):
        tnf, cov = _lib.as_f64(tnf), _lib.as_f64(cov)
        n = (tnf if tnf is not None else cov).shape[0]
        d = tnf.shape[1] if tnf is not None else NBINS
        s = cov.shape[1] // 2 if cov is not None else 1
        ti, td = np.zeros((n, k), np.uint32), np.zeros((n, k), np.float32)
        di, dd = np.zeros((n, k), np.uint32), np.zeros((n, k), np.float32)
        self._chk(self.lib.rg_multi_knn_build(self.m, _lib.ptr(tnf, C.c_double), _lib.ptr(cov, C.c_double), n, d, s, k, variant,
                                              which_mask, _lib.ptr(ti, C.c_uint32), _lib.ptr(td, C.c_float),
                                              _lib.ptr(di, C.c_uint32), _lib.ptr(dd, C.c_float)))
        return (ti, td), (di, dd)

    def knn_build_mutual(self, tnf, cov, k: int, keep_n_edges: int, variant: int = VARIANT_RUST):
        tnf, cov = _lib.as_f64(tnf), _lib.as_f64(cov)
        n, d, s = tnf.shape[0], tnf.shape[1], cov.shape[1] // 2
        off = np.zeros(n + 1, np.uint64)
        nbr, w = np.zeros(n * k, np.uint32), np.zeros(n * k, np.float32)
        disc, nd = np.zeros(n, np.uint32), C.c_uint64(0)
        self._chk(self.lib.rg_multi_knn_build_mutual(self.m, _lib.ptr(tnf, C.c_double), _lib.ptr(cov, C.c_double), n, d, s, k,
                                                     variant, keep_n_edges, _lib.ptr(off, C.c_uint64), _lib.ptr(nbr, C.c_uint32),
                                                     _lib.ptr(w, C.c_float), C.byref(nd), _lib.ptr(disc, C.c_uint32)))
        e = int(off[-1])
        return off, nbr[:e].copy(), w[:e].copy(), disc[:nd.value].copy()

    def tnf_count_ascii(self, seq, offsets, out=None):
        seq = np.ascontiguousarray(seq, np.uint8)
        offsets = np.ascontiguousarray(offsets, np.uint64)
        n = len(offsets) - 1
        counts, nk, freqs = out if out is not None else (np.zeros((n, NBINS), np.uint32), np.zeros(n, np.uint64),
                                                         np.zeros((n, NBINS), np.float64))
        self._chk(self.lib.rg_multi_tnf_count_ascii(self.m, _lib.ptr(seq, C.c_uint8), _lib.ptr(offsets, C.c_uint64), n,
                                                    _lib.ptr(counts, C.c_uint32), _lib.ptr(nk, C.c_uint64),
                                                    _lib.ptr(freqs, C.c_double)))
        return counts, nk, freqs

    def sketch_contigs(self, seq, offsets, kmer_length: int = 31, scale: float = 0.1, seed: int = 42):
        """rg_multi_sketch_contigs + rg_multi_sketch_fetch -> (hashes, sketch_offsets, n_valid_kmers)."""
        seq = np.ascontiguousarray(seq, np.uint8)
        offsets = np.ascontiguousarray(offsets, np.uint64)
        n = len(offsets) - 1
        sk_off = np.zeros(n + 1, np.uint64)
        nk = np.zeros(max(n, 1), np.uint64)
        total = C.c_uint64(0)
        self._chk(self.lib.rg_multi_sketch_contigs(self.m, _lib.ptr(seq, C.c_uint8) if len(seq) else None, _lib.ptr(offsets, C.c_uint64),
                                                   n, kmer_length, scale, seed, _lib.ptr(sk_off, C.c_uint64),
                                                   _lib.ptr(nk, C.c_uint64), C.byref(total)))
        hashes = np.zeros(max(total.value, 1), np.uint64)
        self._chk(self.lib.rg_multi_sketch_fetch(self.m, _lib.ptr(hashes, C.c_uint64), len(hashes)))
        return hashes[: total.value], sk_off, nk[:n]
