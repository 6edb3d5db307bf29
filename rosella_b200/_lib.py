"""ctypes binding of librosella_gpu.so (include/rosella_gpu.h).  No CPU fallback: a missing library
or a missing sm_100 device raises."""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "librosella_gpu.so")
NBINS = 136
VARIANT_RUST, VARIANT_FLIGHT = 0, 1
GRAPH_TNF, GRAPH_DEPTH, GRAPH_TNF_EUCLID = 0, 1, 2

# every symbol include/rosella_gpu.h declares (tests check the .so exports exactly these)
SYMBOLS = [
    "rg_abi_version", "rg_create", "rg_destroy", "rg_last_error", "rg_sync", "rg_stream", "rg_set_option",
    "rg_tnf_canonical_kmers", "rg_tnf_count_ascii", "rg_tnf_packed_words", "rg_tnf_pack_host",
    "rg_tnf_count_packed", "rg_tnf_load_packed", "rg_tnf_load_ascii", "rg_tnf_count_resident", "rg_tnf_fetch",
    "rg_knn_load", "rg_knn_load_sketches", "rg_sketch_contigs", "rg_sketch_fetch", "rg_knn_search", "rg_knn_search_dev", "rg_knn_build",
    "rg_knn_kprime", "rg_knn_depth_partial_dev", "rg_knn_depth_merge_dev", "rg_knn_build_mutual",
    "rg_graph_mutual", "rg_graph_mutual_static", "rg_graph_intersect", "rg_get_timings", "rg_knn_flagged_rows",
    "rg_knn_refilter", "rg_knn_build_subset", "rg_rank_stat", "rg_condensed_euclidean", "rg_silhouette",
    "rg_comm_unique_id", "rg_comm_init", "rg_comm_destroy", "rg_comm_shard", "rg_knn_load_dist", "rg_knn_build_dist_dev",
    "rg_knn_build_dist", "rg_get_dist_timings", "rg_get_dist_timings_of",
    "rg_multi_create", "rg_multi_destroy", "rg_multi_last_error", "rg_multi_size", "rg_multi_handle", "rg_multi_knn_build", "rg_multi_knn_build_sketched",
    "rg_multi_knn_build_mutual", "rg_multi_tnf_count_ascii", "rg_multi_sketch_contigs", "rg_multi_sketch_fetch",
]


class Timings(C.Structure):
    _fields_ = [("h2d_ms", C.c_double), ("transform_ms", C.c_double), ("select_ms", C.c_double),
                ("rescore_ms", C.c_double), ("d2h_ms", C.c_double), ("tnf_count_ms", C.c_double),
                ("kernel_launches", C.c_uint64), ("h2d_bytes", C.c_uint64), ("d2h_bytes", C.c_uint64),
                ("rows_flagged", C.c_uint64), ("rows_exact", C.c_uint64), ("extra_pairs", C.c_uint64)]

    def as_dict(self):
        return {k: getattr(self, k) for k, _ in self._fields_}


class DistTimings(C.Structure):
    _fields_ = [("load_h2d_ms", C.c_double), ("load_bcast_ms", C.c_double), ("load_transform_ms", C.c_double),
                ("partial_ms", C.c_double), ("exchange_ms", C.c_double), ("finish_ms", C.c_double), ("gather_ms", C.c_double)]

    def as_dict(self):
        return {k: getattr(self, k) for k, _ in self._fields_}


class RosellaGpuError(RuntimeError):
    def __init__(self, code: int, msg: str):
        super().__init__(f"librosella_gpu error {code}: {msg}")
        self.code = code


_lib = None


def load() -> C.CDLL:
    global _lib
    if _lib is not None:
        return _lib
    # The library dlopens NCCL on first multi-GPU use and takes a copy the process already holds.  PyTorch bundles its own
    # libnccl.so.2 (newer than the system's) and must see that one: import torch first when it is installed, so that the
    # harness, the tests and bench.py never end up with two NCCL versions fighting over one SONAME.
    try:
        import torch  # noqa: F401
    except ImportError:
        pass
    if not os.path.exists(LIB_PATH):
        raise ImportError(
            f"{LIB_PATH} is missing: build it with `python -m rosella_b200.build` "
            "(there is no CPU fallback for the GPU path)")
    lib = C.CDLL(LIB_PATH)
    vp, u8p, u32p, u64p, f32p, f64p = (C.c_void_p, C.POINTER(C.c_uint8), C.POINTER(C.c_uint32),
                                       C.POINTER(C.c_uint64), C.POINTER(C.c_float), C.POINTER(C.c_double))
    sig = {
        "rg_abi_version": (C.c_int32, []),
        "rg_create": (C.c_int32, [C.c_int32, C.POINTER(vp)]),
        "rg_destroy": (None, [vp]),
        "rg_last_error": (C.c_char_p, [vp]),
        "rg_sync": (C.c_int32, [vp]),
        "rg_stream": (vp, [vp]),
        "rg_set_option": (C.c_int32, [vp, C.c_char_p, C.c_int64]),
        "rg_tnf_canonical_kmers": (C.c_int32, [u8p]),
        "rg_tnf_count_ascii": (C.c_int32, [vp, u8p, u64p, C.c_uint64, u32p, u64p, f64p]),
        "rg_tnf_packed_words": (C.c_uint64, [u8p, u64p, C.c_uint64]),
        "rg_tnf_pack_host": (C.c_int32, [u8p, u64p, C.c_uint64, u64p, u32p, u64p, C.c_int32]),
        "rg_tnf_count_packed": (C.c_int32, [vp, u64p, u32p, u64p, C.c_uint64, u32p, u64p, f64p]),
        "rg_tnf_load_packed": (C.c_int32, [vp, u64p, u32p, u64p, C.c_uint64]),
        "rg_tnf_load_ascii": (C.c_int32, [vp, u8p, u64p, C.c_uint64]),
        "rg_tnf_count_resident": (C.c_int32, [vp]),
        "rg_tnf_fetch": (C.c_int32, [vp, u32p, u64p, f64p]),
        "rg_knn_load": (C.c_int32, [vp, f64p, f64p, C.c_uint64, C.c_uint32, C.c_uint32, C.c_uint32]),
        "rg_knn_load_sketches": (C.c_int32, [vp, u64p, u64p, C.c_uint64, C.c_double]),
        "rg_sketch_contigs": (C.c_int32, [vp, u8p, u64p, C.c_uint64, C.c_uint32, C.c_double, C.c_uint64, u64p, u64p, u64p]),
        "rg_sketch_fetch": (C.c_int32, [vp, u64p, C.c_uint64]),
        "rg_knn_search": (C.c_int32, [vp, C.c_uint32, C.c_uint32, C.c_uint64, C.c_uint64, u32p, f32p, u32p]),
        "rg_knn_search_dev": (C.c_int32, [vp, C.c_uint32, C.c_uint32, C.c_uint64, C.c_uint64, vp, vp, vp]),
        "rg_knn_kprime": (C.c_uint32, [vp, C.c_uint32, C.c_uint32]),
        "rg_knn_depth_partial_dev": (C.c_int32, [vp, C.c_uint32, C.c_uint32, C.c_uint32, vp, vp]),
        "rg_knn_depth_merge_dev": (C.c_int32, [vp, C.c_uint32, C.c_uint32, vp, vp, C.c_uint64, C.c_uint64, vp, vp, vp]),
        "rg_knn_build": (C.c_int32, [vp, f64p, f64p, C.c_uint64, C.c_uint32, C.c_uint32, C.c_uint32, C.c_uint32,
                                     C.c_uint32, u32p, f32p, u32p, f32p]),
        "rg_knn_build_mutual": (C.c_int32, [vp, f64p, f64p, C.c_uint64, C.c_uint32, C.c_uint32, C.c_uint32, C.c_uint32,
                                            C.c_uint32, C.POINTER(C.c_uint64), u32p, f32p, C.POINTER(C.c_uint64), u32p]),
        "rg_graph_mutual": (C.c_int32, [vp, u32p, C.c_uint64, C.c_uint32, C.c_uint32, u8p]),
        "rg_graph_mutual_static": (C.c_int32, [vp, u32p, C.c_uint64, C.c_uint32, C.c_uint32, u8p]),
        "rg_graph_intersect": (C.c_int32, [vp, u32p, u32p, C.c_uint64, C.c_uint32, C.c_uint32, u8p]),
        "rg_get_timings": (C.c_int32, [vp, C.POINTER(Timings)]),
        "rg_knn_flagged_rows": (C.c_int32, [vp, u32p, f32p, f32p, C.c_uint64, C.POINTER(C.c_uint64)]),
        "rg_rank_stat": (C.c_int32, [vp, f64p, f64p, f64p, C.c_uint64, C.c_uint32, C.c_uint32, C.c_uint32, f64p, u32p, f64p]),
        "rg_condensed_euclidean": (C.c_int32, [vp, f64p, C.c_uint64, C.c_uint32, f64p]),
        "rg_silhouette": (C.c_int32, [vp, f64p, C.c_uint64, C.c_uint32, C.POINTER(C.c_int32), C.c_uint32, f64p, f64p]),
        "rg_knn_refilter": (C.c_int32, [vp, C.c_uint32, u64p, u32p, f32p, u64p, u32p]),
        "rg_knn_build_subset": (C.c_int32, [vp, f64p, f64p, C.c_uint64, C.c_uint32, C.c_uint32, u64p, C.c_uint64, C.c_uint32,
                                            C.c_uint32, C.c_uint32, u32p, f32p, u32p, f32p]),
        "rg_comm_unique_id": (C.c_int32, [u8p]),
        "rg_comm_init": (C.c_int32, [vp, C.c_int32, C.c_int32, u8p]),
        "rg_comm_destroy": (C.c_int32, [vp]),
        "rg_comm_shard": (C.c_int32, [C.c_uint64, C.c_int32, C.c_int32, u64p, u64p]),
        "rg_knn_load_dist": (C.c_int32, [vp, f64p, f64p, C.c_uint64, C.c_uint32, C.c_uint32, C.c_uint32, C.c_int32]),
        "rg_knn_build_dist_dev": (C.c_int32, [vp, C.c_uint32, C.c_uint32, C.c_uint32, vp, vp]),
        "rg_knn_build_dist": (C.c_int32, [vp, C.c_uint32, C.c_uint32, C.c_uint32, u32p, f32p]),
        "rg_get_dist_timings": (C.c_int32, [vp, C.POINTER(DistTimings)]),
        "rg_get_dist_timings_of": (C.c_int32, [vp, C.c_uint32, C.POINTER(DistTimings)]),
        "rg_multi_create": (C.c_int32, [C.POINTER(C.c_int32), C.c_int32, C.POINTER(vp)]),
        "rg_multi_destroy": (None, [vp]),
        "rg_multi_last_error": (C.c_char_p, [vp]),
        "rg_multi_size": (C.c_int32, [vp]),
        "rg_multi_handle": (vp, [vp, C.c_int32]),
        "rg_multi_knn_build": (C.c_int32, [vp, f64p, f64p, C.c_uint64, C.c_uint32, C.c_uint32, C.c_uint32, C.c_uint32,
                                           C.c_uint32, u32p, f32p, u32p, f32p]),
        "rg_multi_knn_build_sketched": (C.c_int32, [vp, f64p, f64p, C.c_uint64, C.c_uint32, C.c_uint32, C.c_uint32, C.c_uint32,
                                                    C.c_uint32, u64p, u64p, C.c_double, u32p, f32p, u32p, f32p]),
        "rg_multi_knn_build_mutual": (C.c_int32, [vp, f64p, f64p, C.c_uint64, C.c_uint32, C.c_uint32, C.c_uint32, C.c_uint32,
                                                  C.c_uint32, u64p, u32p, f32p, u64p, u32p]),
        "rg_multi_tnf_count_ascii": (C.c_int32, [vp, u8p, u64p, C.c_uint64, u32p, u64p, f64p]),
        "rg_multi_sketch_contigs": (C.c_int32, [vp, u8p, u64p, C.c_uint64, C.c_uint32, C.c_double, C.c_uint64, u64p, u64p, u64p]),
        "rg_multi_sketch_fetch": (C.c_int32, [vp, u64p, C.c_uint64]),
    }
    for name, (res, args) in sig.items():
        fn = getattr(lib, name)
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib


def ptr(a, ctype):
    """ctypes pointer to a C-contiguous numpy array (None -> NULL)."""
    if a is None:
        return None
    assert a.flags["C_CONTIGUOUS"]
    return a.ctypes.data_as(C.POINTER(ctype))


def as_f64(a):
    return None if a is None else np.ascontiguousarray(a, np.float64)
