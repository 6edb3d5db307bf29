"""ctypes view of librosella_host.so — the C++ mirror of the reference's Rust interface for the hot path
(rosella_b200/host/rosella_host.hpp: KmerFrequencyTable, count_kmers, EmbedderEngine::build_mutual_kgraph,
mutual_knn / intersect, KGraph).  Used by the tests so that they read like tests of the reference's own API;
C++ callers include rosella_host.hpp, the Rust shim binds include/rosella_gpu.h directly."""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "librosella_host.so")
_lib = None


class HostError(RuntimeError):
    pass


def load() -> C.CDLL:
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise ImportError(f"{LIB_PATH} is missing: build it with `python -m rosella_b200.build`")
    C.CDLL(os.path.join(_HERE, "librosella_gpu.so"), mode=C.RTLD_GLOBAL)
    lib = C.CDLL(LIB_PATH)
    vp, sz, dp = C.c_void_p, C.c_size_t, C.POINTER(C.c_double)
    sig = {
        "rh_last_error": (C.c_char_p, []),
        "rh_format_f64": (C.c_int, [C.c_double, C.c_char_p, sz]),
        "rh_clamp_n_neighbours": (sz, [sz, sz]),
        "rh_table_new": (vp, [dp, sz, sz, C.POINTER(C.c_char_p)]),
        "rh_table_free": (None, [vp]),
        "rh_table_write": (C.c_int, [vp, C.c_char_p]),
        "rh_table_read": (vp, [C.c_char_p, C.c_int]),
        "rh_table_nrows": (sz, [vp]), "rh_table_ncols": (sz, [vp]), "rh_table_kmer_size": (sz, [vp]),
        "rh_table_data": (dp, [vp]), "rh_table_name": (C.c_char_p, [vp, sz]), "rh_table_path": (C.c_char_p, [vp]),
        "rh_table_filter_by_index": (C.c_long, [vp, C.POINTER(sz), sz]),
        "rh_table_filter_by_name": (C.c_long, [vp, C.POINTER(C.c_char_p), sz]),
        "rh_clr": (C.c_int, [dp, sz, sz]),
        "rh_fastx_parse": (vp, [C.c_char_p]), "rh_fastx_free": (None, [vp]), "rh_fastx_n": (sz, [vp]),
        "rh_fastx_id": (C.c_char_p, [vp, sz]), "rh_fastx_seq": (C.POINTER(C.c_uint8), [vp]),
        "rh_fastx_offsets": (C.POINTER(C.c_uint64), [vp]),
        "rh_count_kmers": (vp, [C.c_int, C.c_char_p, C.c_char_p, C.c_int]),
        "rh_sketch_contigs": (vp, [C.c_int, C.c_char_p, sz, sz, C.c_double, C.c_uint64]),
        "rh_sketches_new": (vp, [C.POINTER(C.c_char_p), C.POINTER(C.c_uint64), C.POINTER(C.c_uint64), sz]),
        "rh_sketches_filter_by_index": (C.c_int, [vp, C.POINTER(sz), sz]),
        "rh_sketches_free": (None, [vp]), "rh_sketches_n": (sz, [vp]), "rh_sketches_name": (C.c_char_p, [vp, sz]),
        "rh_sketches_offsets": (C.POINTER(C.c_uint64), [vp]), "rh_sketches_hashes": (C.POINTER(C.c_uint64), [vp]),
        "rh_sketches_num_valid_kmers": (C.POINTER(C.c_uint64), [vp]), "rh_sketches_seq_length": (C.POINTER(C.c_uint64), [vp]),
        "rh_cov_from_file": (vp, [C.c_char_p, C.c_int]),
        "rh_cov_new": (vp, [dp, sz, sz, C.POINTER(C.c_char_p), C.POINTER(sz), C.POINTER(C.c_char_p)]),
        "rh_cov_free": (None, [vp]), "rh_cov_write": (C.c_int, [vp, C.c_char_p]), "rh_cov_merge": (C.c_int, [vp, vp]),
        "rh_cov_filter_by_length": (C.c_long, [vp, sz]), "rh_cov_nrows": (sz, [vp]), "rh_cov_ncols": (sz, [vp]),
        "rh_cov_n_samples": (sz, [vp]), "rh_cov_data": (dp, [vp]), "rh_cov_average_depths": (dp, [vp]),
        "rh_cov_lengths": (C.POINTER(sz), [vp]), "rh_cov_name": (C.c_char_p, [vp, sz]), "rh_cov_sample": (C.c_char_p, [vp, sz]),
        "rh_cov_path": (C.c_char_p, [vp]),
        "rh_recover_new": (vp, [dp, sz, sz, dp, sz, C.POINTER(C.c_char_p), C.POINTER(C.c_uint64), C.POINTER(C.c_uint64), sz, C.c_int,
                                C.c_uint32, C.c_double]),
        "rh_recover_from_files": (vp, [C.c_int, C.c_char_p, C.c_char_p, C.c_char_p, C.c_int, C.c_char_p, sz, sz, C.c_int]),
        "rh_recover_cov_cols": (sz, [vp]), "rh_recover_tnf_cols": (sz, [vp]),
        "rh_recover_free": (None, [vp]), "rh_recover_n_contigs": (sz, [vp]), "rh_recover_prefilter_rounds": (sz, [vp]),
        "rh_recover_n_filtered": (sz, [vp]), "rh_recover_is_filtered": (C.c_int, [vp, C.c_char_p]),
        "rh_recover_name": (C.c_char_p, [vp, sz]), "rh_recover_tnf_name": (C.c_char_p, [vp, sz]),
        "rh_recover_cov_row": (dp, [vp, sz]), "rh_recover_tnf_row": (dp, [vp, sz]), "rh_recover_sketch_len": (C.c_long, [vp, sz]),
        "rh_recover_retrieve": (C.c_long, [vp, C.POINTER(sz), sz, dp, dp, C.POINTER(C.c_long)]),
        "rh_recover_filter_contigs": (C.c_int, [vp, C.POINTER(sz), sz]),
        "rh_recover_prefilter_nodes": (vp, [C.c_int, vp, sz]),
        "rh_recover_evaluate_subset_graph": (vp, [C.c_int, vp, C.POINTER(sz), sz, sz]),
        "rh_build_mutual_kgraph": (vp, [C.c_int, dp, sz, dp, sz, sz, sz, sz, sz, sz, C.c_int, C.c_uint32,
                                        C.POINTER(C.c_uint64), C.POINTER(C.c_uint64), C.c_double]),
        "rh_graph_free": (None, [vp]), "rh_graph_nb_nodes": (sz, [vp]), "rh_graph_max_nbng": (sz, [vp]),
        "rh_graph_n_disconnected": (sz, [vp]), "rh_graph_disconnected": (C.POINTER(sz), [vp]),
        "rh_graph_degree": (sz, [vp, C.c_int, sz]),
        "rh_graph_neighbours": (C.c_int, [vp, C.c_int, sz, C.POINTER(sz), dp]),
        "rh_graph_data_id_from_idx": (C.c_long, [vp, sz]),
    }
    for name, (res, args) in sig.items():
        f = getattr(lib, name)
        f.restype, f.argtypes = res, args
    _lib = lib
    return lib


def _err() -> HostError:
    return HostError(load().rh_last_error().decode())


def format_f64(v: float) -> str:
    buf = C.create_string_buffer(64)
    if load().rh_format_f64(v, buf, 64):
        raise _err()
    return buf.value.decode()


def clamp_n_neighbours(n_included: int, n_neighbours: int) -> int:
    return load().rh_clamp_n_neighbours(n_included, n_neighbours)


def clr(table: np.ndarray) -> np.ndarray:
    t = np.ascontiguousarray(table, np.float64).copy()
    if load().rh_clr(t.ctypes.data_as(C.POINTER(C.c_double)), t.shape[0], t.shape[1]):
        raise _err()
    return t


class KmerFrequencyTable:
    """kmer_counting.rs:211-360"""

    def __init__(self, handle):
        if not handle:
            raise _err()
        self._h = handle

    @classmethod
    def new(cls, table: np.ndarray, names: list[str]):
        t = np.ascontiguousarray(table, np.float64)
        arr = (C.c_char_p * len(names))(*[s.encode() for s in names])
        return cls(load().rh_table_new(t.ctypes.data_as(C.POINTER(C.c_double)), t.shape[0], t.shape[1], arr))

    @classmethod
    def read(cls, path: str, apply_clr: bool = False):
        return cls(load().rh_table_read(path.encode(), int(apply_clr)))

    def write(self, path: str):
        if load().rh_table_write(self._h, path.encode()):
            raise _err()

    @property
    def kmer_table(self) -> np.ndarray:
        lib = load()
        n, m = lib.rh_table_nrows(self._h), lib.rh_table_ncols(self._h)
        if n * m == 0:
            return np.zeros((n, m))
        return np.ctypeslib.as_array(lib.rh_table_data(self._h), shape=(n, m)).copy()

    @property
    def contig_names(self) -> list[str]:
        lib = load()
        return [lib.rh_table_name(self._h, i).decode() for i in range(lib.rh_table_nrows(self._h))]

    @property
    def kmer_size(self) -> int:
        return load().rh_table_kmer_size(self._h)

    @property
    def table_path(self) -> str:
        return load().rh_table_path(self._h).decode()

    def filter_by_index(self, indices) -> int:
        idx = (C.c_size_t * len(indices))(*sorted(indices))
        r = load().rh_table_filter_by_index(self._h, idx, len(indices))
        if r < 0:
            raise _err()
        return r

    def filter_by_name(self, names) -> int:
        arr = (C.c_char_p * len(names))(*[s.encode() for s in names])
        r = load().rh_table_filter_by_name(self._h, arr, len(names))
        if r < 0:
            raise _err()
        return r

    def __del__(self):
        if getattr(self, "_h", None) and _lib is not None:
            _lib.rh_table_free(self._h)
            self._h = None


def parse_fastx_file(path: str):
    """-> (ids, seq u8, offsets u64) as needletail::parse_fastx_file yields the records"""
    lib = load()
    h = lib.rh_fastx_parse(path.encode())
    if not h:
        raise _err()
    try:
        n = lib.rh_fastx_n(h)
        ids = [lib.rh_fastx_id(h, i).decode() for i in range(n)]
        offs = np.ctypeslib.as_array(lib.rh_fastx_offsets(h), shape=(n + 1,)).copy()
        seq = np.ctypeslib.as_array(lib.rh_fastx_seq(h), shape=(int(offs[-1]),)).copy() if offs[-1] else np.zeros(0, np.uint8)
        return ids, seq, offs
    finally:
        lib.rh_fastx_free(h)


def count_kmers(assembly: str, output_directory: str, no_flight: bool = False, device: int = 0) -> KmerFrequencyTable:
    """count_kmers(m, n_contigs) (kmer_counting.rs:13-16) on the GPU"""
    return KmerFrequencyTable(load().rh_count_kmers(device, assembly.encode(), output_directory.encode(), int(no_flight)))


def _sketches_out(lib, h):
    n = lib.rh_sketches_n(h)
    names = [lib.rh_sketches_name(h, i).decode() for i in range(n)]
    offs = np.ctypeslib.as_array(lib.rh_sketches_offsets(h), (n + 1,)).copy()
    total = int(offs[-1])
    hashes = np.ctypeslib.as_array(lib.rh_sketches_hashes(h), (total,)).copy() if total else np.zeros(0, np.uint64)
    nk = np.ctypeslib.as_array(lib.rh_sketches_num_valid_kmers(h), (n,)).copy() if n else np.zeros(0, np.uint64)
    sl = np.ctypeslib.as_array(lib.rh_sketches_seq_length(h), (n,)).copy() if n else np.zeros(0, np.uint64)
    return names, hashes, offs, nk, sl


def sketch_result_filter_by_index(names, hashes, offsets, indices_to_remove):
    """ContigSketchResult::filter_by_index (contig_sketcher.rs:158-176) on caller data -> (names, hashes, offsets)"""
    lib = load()
    arr = (C.c_char_p * len(names))(*[s.encode() for s in names])
    offsets = np.ascontiguousarray(offsets, np.uint64)
    hashes = np.ascontiguousarray(hashes, np.uint64) if len(hashes) else np.zeros(1, np.uint64)
    h = lib.rh_sketches_new(arr, offsets.ctypes.data_as(C.POINTER(C.c_uint64)), hashes.ctypes.data_as(C.POINTER(C.c_uint64)), len(names))
    if not h:
        raise _err()
    try:
        idx = (C.c_size_t * len(indices_to_remove))(*sorted(indices_to_remove))
        if lib.rh_sketches_filter_by_index(h, idx, len(indices_to_remove)):
            raise _err()
        out = _sketches_out(lib, h)
    finally:
        lib.rh_sketches_free(h)
    return out[0], out[1], out[2]


def sketch_contigs(assembly: str, min_contig_size: int = 1500, kmer_size: int = 31, sketch_scale: float = 0.1, seed: int = 42,
                   device: int = 0):
    """sketch_contigs(m) (contig_sketcher.rs:11-92) on the GPU
    -> (contig_names, hashes u64[...], offsets u64[n+1], num_valid_kmers u64[n], seq_length u64[n])"""
    lib = load()
    h = lib.rh_sketch_contigs(device, assembly.encode(), min_contig_size, kmer_size, sketch_scale, seed)
    if not h:
        raise _err()
    try:
        return _sketches_out(lib, h)
    finally:
        lib.rh_sketches_free(h)


class KGraphResult:
    """(KGraph<f64>, disconnected_nodes) of EmbedderEngine::build_mutual_kgraph plus the TNF kgraph of the build"""

    def __init__(self, handle):
        if not handle:
            raise _err()
        self._h = handle

    @property
    def nb_nodes(self) -> int:
        return load().rh_graph_nb_nodes(self._h)

    @property
    def max_nbng(self) -> int:
        return load().rh_graph_max_nbng(self._h)

    @property
    def disconnected_nodes(self) -> list[int]:
        lib = load()
        n = lib.rh_graph_n_disconnected(self._h)
        p = lib.rh_graph_disconnected(self._h)
        return [int(p[i]) for i in range(n)]

    def neighbours(self, node: int, tnf: bool = False):
        lib = load()
        deg = lib.rh_graph_degree(self._h, int(tnf), node)
        nodes = (C.c_size_t * max(deg, 1))()
        w = (C.c_double * max(deg, 1))()
        if lib.rh_graph_neighbours(self._h, int(tnf), node, nodes, w):
            raise _err()
        return np.array(nodes[:deg], np.int64), np.array(w[:deg], np.float64)

    def data_id_from_idx(self, idx: int) -> int:
        return load().rh_graph_data_id_from_idx(self._h, idx)

    def __del__(self):
        if getattr(self, "_h", None) and _lib is not None:
            _lib.rh_graph_free(self._h)
            self._h = None


SHORT_READ, LONG_READ, SHORT_BAM, LONG_BAM = 0, 1, 2, 3          # MappingMode, coverm_engine.rs:217-222


class CoverageTable:
    """src/coverage/coverage_table.rs: coverm's TSV <-> the n x 2S table the coverage metric reads"""

    def __init__(self, handle):
        if not handle:
            raise _err()
        self._h = handle

    @classmethod
    def from_file(cls, path: str, mode: int = SHORT_BAM):
        return cls(load().rh_cov_from_file(path.encode(), mode))

    @classmethod
    def new(cls, table, names, lengths, samples):
        t = np.ascontiguousarray(table, np.float64)
        na = (C.c_char_p * len(names))(*[s.encode() for s in names])
        sa = (C.c_char_p * len(samples))(*[s.encode() for s in samples])
        la = (C.c_size_t * len(lengths))(*[int(x) for x in lengths])
        return cls(load().rh_cov_new(t.ctypes.data_as(C.POINTER(C.c_double)), t.shape[0], t.shape[1], na, la, sa))

    def __del__(self):
        if getattr(self, "_h", None):
            load().rh_cov_free(self._h)
            self._h = None

    def write(self, path: str):
        if load().rh_cov_write(self._h, path.encode()):
            raise _err()

    def merge(self, other: "CoverageTable"):
        if load().rh_cov_merge(self._h, other._h):
            raise _err()

    def filter_by_length(self, min_contig_size: int) -> int:
        r = load().rh_cov_filter_by_length(self._h, min_contig_size)
        if r < 0:
            raise _err()
        return r

    @property
    def table(self) -> np.ndarray:
        lib = load()
        n, c = lib.rh_cov_nrows(self._h), lib.rh_cov_ncols(self._h)
        return np.ctypeslib.as_array(lib.rh_cov_data(self._h), (n * c,)).copy().reshape(n, c) if n * c else np.zeros((n, c))

    @property
    def average_depths(self) -> np.ndarray:
        lib = load()
        n = lib.rh_cov_nrows(self._h)
        return np.ctypeslib.as_array(lib.rh_cov_average_depths(self._h), (n,)).copy() if n else np.zeros(0)

    @property
    def contig_lengths(self):
        lib = load()
        n = lib.rh_cov_nrows(self._h)
        return [lib.rh_cov_lengths(self._h)[i] for i in range(n)]

    @property
    def contig_names(self):
        lib = load()
        return [lib.rh_cov_name(self._h, i).decode() for i in range(lib.rh_cov_nrows(self._h))]

    @property
    def sample_names(self):
        lib = load()
        return [lib.rh_cov_sample(self._h, i).decode() for i in range(lib.rh_cov_n_samples(self._h))]

    @property
    def output_path(self) -> str:
        return load().rh_cov_path(self._h).decode()


class RecoverEngine:
    """The graph driver of RecoverEngine (recover_engine.rs:407-543): tables + prefilter_nodes / filter_contigs /
    retrieve_contig_data; sketches = (hashes u64, offsets u64[n + 1]) or None."""

    def __init__(self, cov, tnf, names, n_neighbours: int = 100, apply_filters: bool = False, variant: int = 0, sketches=None,
                 sketch_scale: float = 0.1):
        c = np.ascontiguousarray(cov, np.float64)
        t = np.ascontiguousarray(tnf, np.float64)
        dp, up = C.POINTER(C.c_double), C.POINTER(C.c_uint64)
        arr = (C.c_char_p * len(names))(*[s.encode() for s in names])
        so = sh = None
        if sketches is not None:
            sh_a = np.ascontiguousarray(sketches[0], np.uint64) if len(sketches[0]) else np.zeros(1, np.uint64)
            so_a = np.ascontiguousarray(sketches[1], np.uint64)
            sh, so = sh_a.ctypes.data_as(up), so_a.ctypes.data_as(up)
        self._c, self._d = c.shape[1], t.shape[1]
        self._h = load().rh_recover_new(c.ctypes.data_as(dp), c.shape[0], c.shape[1], t.ctypes.data_as(dp), t.shape[1], arr, so, sh,
                                        n_neighbours, int(apply_filters), variant, sketch_scale)
        if not self._h:
            raise _err()

    @classmethod
    def from_files(cls, assembly: str, output_directory: str, coverage_file: str, mode: int = 2, kmer_frequency_file=None,
                   min_contig_size: int = 1500, n_neighbours: int = 100, apply_filters: bool = False, device: int = 0):
        """RecoverEngine::new (recover_engine.rs:93-131) without the coverm subprocess"""
        self = cls.__new__(cls)
        lib = load()
        self._h = lib.rh_recover_from_files(device, assembly.encode(), output_directory.encode(), coverage_file.encode(), mode,
                                            kmer_frequency_file.encode() if kmer_frequency_file else None, min_contig_size,
                                            n_neighbours, int(apply_filters))
        if not self._h:
            raise _err()
        self._c, self._d = lib.rh_recover_cov_cols(self._h), lib.rh_recover_tnf_cols(self._h)
        return self

    def __del__(self):
        if getattr(self, "_h", None):
            load().rh_recover_free(self._h)
            self._h = None

    @property
    def n_contigs(self) -> int:
        return load().rh_recover_n_contigs(self._h)

    @property
    def prefilter_rounds(self) -> int:
        return load().rh_recover_prefilter_rounds(self._h)

    @property
    def n_filtered(self) -> int:
        return load().rh_recover_n_filtered(self._h)

    def is_filtered(self, name: str) -> bool:
        return bool(load().rh_recover_is_filtered(self._h, name.encode()))

    def names(self):
        lib = load()
        out = [lib.rh_recover_name(self._h, i).decode() for i in range(self.n_contigs)]
        assert out == [lib.rh_recover_tnf_name(self._h, i).decode() for i in range(self.n_contigs)]
        return out

    def tables(self):
        """-> (cov [n, c], tnf [n, d], sketch lengths or None) as the engine holds them now"""
        lib, n = load(), self.n_contigs
        cov = np.array([np.ctypeslib.as_array(lib.rh_recover_cov_row(self._h, i), (self._c,)).copy() for i in range(n)]).reshape(n, self._c)
        tnf = np.array([np.ctypeslib.as_array(lib.rh_recover_tnf_row(self._h, i), (self._d,)).copy() for i in range(n)]).reshape(n, self._d)
        sl = [lib.rh_recover_sketch_len(self._h, i) for i in range(n)]
        return cov, tnf, (None if n and sl[0] < 0 else sl)

    def retrieve_contig_data(self, indices):
        idx = sorted(set(int(i) for i in indices))
        arr = (C.c_size_t * len(idx))(*idx)
        cov0, tnf0 = np.zeros(len(idx)), np.zeros(len(idx))
        sl = (C.c_long * max(len(idx), 1))()
        cnt = load().rh_recover_retrieve(self._h, arr, len(idx), cov0.ctypes.data_as(C.POINTER(C.c_double)),
                                         tnf0.ctypes.data_as(C.POINTER(C.c_double)), sl)
        if cnt < 0:
            raise _err()
        return cov0[:cnt], tnf0[:cnt], list(sl)[:cnt]

    def filter_contigs(self, nodes):
        arr = (C.c_size_t * len(nodes))(*[int(i) for i in nodes])
        if load().rh_recover_filter_contigs(self._h, arr, len(nodes)):
            raise _err()

    def evaluate_subset_graph(self, indices, keep_n_edges: int, device: int = 0) -> "KGraphResult":
        """graph half of evaluate_subset (:410-435); .disconnected_nodes holds ORIGINAL row indices"""
        idx = sorted(set(int(i) for i in indices))
        arr = (C.c_size_t * len(idx))(*idx)
        return KGraphResult(load().rh_recover_evaluate_subset_graph(device, self._h, arr, len(idx), keep_n_edges))

    def prefilter_nodes(self, rounds: int = 1, device: int = 0) -> "KGraphResult":
        return KGraphResult(load().rh_recover_prefilter_nodes(device, self._h, rounds))


def build_mutual_kgraph(tnf: np.ndarray, cov: np.ndarray, keep_n_edges: int, n_neighbours: int, max_layers: int = 16,
                        ef_construction: int = 200, apply_filters: bool = False, variant: int = 0, device: int = 0,
                        sketches=None, sketch_scale: float = 0.1):
    """EmbedderEngine::new(contig_data).build_mutual_kgraph(...) (embedder.rs:31-86); sketches = (hashes u64, offsets u64
    [n + 1]): the ContigInformation.sketch members (embedder.rs:96)"""
    t = np.ascontiguousarray(tnf, np.float64)
    c = np.ascontiguousarray(cov, np.float64)
    dp, up = C.POINTER(C.c_double), C.POINTER(C.c_uint64)
    sh = so = None
    if sketches is not None:
        sh_a, so_a = np.ascontiguousarray(sketches[0], np.uint64), np.ascontiguousarray(sketches[1], np.uint64)
        sh, so = sh_a.ctypes.data_as(up), so_a.ctypes.data_as(up)
    return KGraphResult(load().rh_build_mutual_kgraph(device, t.ctypes.data_as(dp), t.shape[1], c.ctypes.data_as(dp),
                                                      c.shape[1], t.shape[0], keep_n_edges, n_neighbours, max_layers,
                                                      ef_construction, int(apply_filters), variant, sh, so, sketch_scale))
