"""CPU legs of bench.py: the oracle (oracle/) timed on the box's host cores.  This is the one place
besides tests/ and smoke() that may execute oracle/ — as the reported baseline, never as the product.

No reference binary can be built (the reference is Rust only and no Rust toolchain exists here or on
the GPU box: DESIGN.md §3), so kind = "port": the f64 C restatement of the reference's metric code
(coverage_calculator.rs:588-666, kmer_counting.rs:366-414) evaluated exhaustively with OpenMP."""
from __future__ import annotations

import os
import time

import numpy as np


def host_threads() -> int:
    """Cores this process may run on.  NOT omp_get_max_threads(): torch.distributed.run exports OMP_NUM_THREADS=1 to
    every rank, which made the round-1 reference arm run on one thread at N > 1."""
    try:
        return max(1, len(os.sched_getaffinity(0)))
    except AttributeError:
        return max(1, os.cpu_count() or 1)


def knn_sample(tnf, cov, k: int, budget_s: float, variant: int = 0, threads: int = 0):
    """Exhaustive exact kNN (both graphs) for a bounded number of query rows against ALL rows."""
    import oracle
    threads = threads or host_threads()
    n = tnf.shape[0]
    rng = np.random.default_rng(1)

    def once(rows):
        q = np.sort(rng.choice(n, rows, replace=False)).astype(np.uint64)
        t = time.perf_counter()
        oracle.knn(cov, oracle.DEPTH, k, variant, qrows=q, threads=threads)
        oracle.knn(tnf, oracle.TNF, k, variant, qrows=q, threads=threads)
        return time.perf_counter() - t

    t1 = once(threads)                                   # calibration: one row per thread
    rows = threads * max(1, int(budget_s / max(t1, 1e-3)))
    rows = min(rows, n)
    t = once(rows)
    return {"value": rows / t, "unit": "contigs/s", "cores": threads, "kind": "port",
            "sample": f"{rows} query rows x all {n} reference rows, both graphs (depth + TNF), exhaustive f64, "
                      f"OpenMP {threads} threads, {t:.1f} s",
            "rows": rows, "seconds": t}


def tnf_sample(seq, offs, budget_s: float, threads: int = 0):
    """Faithful single-thread restatement of KmerCounter::run (what the reference does) and a LUT/OpenMP
    version on all cores, on a prefix of the contig set."""
    import oracle
    threads = threads or host_threads()
    n = len(offs) - 1
    # prefix of ~64 Mbp for calibration
    cut = int(np.searchsorted(offs, min(int(offs[-1]), 32_000_000)))
    cut = max(1, min(cut, n))
    t = time.perf_counter()
    oracle.tnf_count(seq, offs[:cut + 1])
    t_f = time.perf_counter() - t
    bp = int(offs[cut])
    faithful = bp / t_f / 1e9
    t = time.perf_counter()
    oracle.tnf_count(seq, offs, fast=True, threads=threads)
    t_fast = time.perf_counter() - t
    return {"faithful_single_thread_gbps": faithful, "faithful_sample_bp": bp,
            "lut_openmp_gbps": int(offs[-1]) / t_fast / 1e9, "lut_openmp_threads": threads,
            "lut_openmp_sample_bp": int(offs[-1])}
