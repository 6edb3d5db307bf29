"""CPU restatement of flight's rank statistic — TEST INFRASTRUCTURE ONLY (see oracle/oracle.h).

Follows flight/flight/distance.py: makeRanks :47-90 (pair weights len_i * len_j in condensed order), makeRankStat :92-108
(w = max(1/3, S / (S + 4)), 1/2 for one sample; dists = fun(cov_ranks) ** w * fun(kmer_ranks)), choose_rank_method :127-136
(pdist(cov, coverage_distance) unweighted, pdist(tnf, 'euclidean') weighted), _fractional_rank :168-201.  The caller
(flight/flight/rosella/rosella.py:337-350) passes fun = a / max(a).  Pinned against the reference's own output:
tests/golden/flight_rankstat.npz (oracle/make_golden.py: make_rankstat)."""
from __future__ import annotations

import numpy as np


def fractional_rank(ar: np.ndarray, weights: np.ndarray | None = None) -> np.ndarray:
    """distance.py:168-201: ranks with ties averaged; with weights, cumulative weights stand in for positions."""
    size = ar.size
    perm = ar.argsort()
    aux = ar[perm]
    flag = np.concatenate((aux[1:] != aux[:-1], [True]))            # ends of streaks of equal values
    if weights is None:
        rflag = np.flatnonzero(flag).astype(np.double) + 1
    else:
        rflag = weights[perm].cumsum().astype(np.double)[flag]
    rflag = np.concatenate((rflag[:1] + 1, rflag[1:] + rflag[:-1] + 1)) * 0.5
    iflag = np.concatenate(([0.], np.cumsum(flag[:-1]))).astype(int)
    r = np.empty(size, dtype=np.double)
    r[perm] = rflag[iflag]
    return r


def pair_weights(lengths: np.ndarray) -> np.ndarray:
    """distance.py:53-58: len_i * len_j for i < j in condensed order, as doubles."""
    n = len(lengths)
    w = np.empty(n * (n - 1) // 2, dtype=np.double)
    k = 0
    for i in range(n - 1):
        w[k:k + n - 1 - i] = lengths[i] * lengths[i + 1:n]
        k += n - 1 - i
    return w


def stat_from_distances(dc: np.ndarray, dk: np.ndarray, lengths: np.ndarray, s: int):
    """distance.py:92-136 from the two condensed distance vectors -> (cov_ranks, kmer_ranks, stat)."""
    cov_ranks = fractional_rank(dc)
    kmer_ranks = fractional_rank(dk, pair_weights(lengths))
    w = max(1 / 3, s / (s + 4))
    if s == 1:
        w = 1 / 2
    return cov_ranks, kmer_ranks, (cov_ranks / cov_ranks.max()) ** w * (kmer_ranks / kmer_ranks.max())


def distances(cov: np.ndarray, tnf: np.ndarray):
    """-> (coverage_distance, Euclidean TNF distance), condensed (i < j), f64."""
    import ctypes as C

    import oracle
    from scipy.spatial.distance import pdist
    cov = np.ascontiguousarray(cov, np.float64)
    n = cov.shape[0]
    dc = np.empty(n * (n - 1) // 2)
    # metrics.py:429-443 coverage_distance: identical rows -> 0, else the geometric mean of metabat_distance
    oracle.lib().orc_coverage_distance_condensed(cov.ctypes.data_as(C.POINTER(C.c_double)), C.c_uint64(n), C.c_uint32(cov.shape[1] // 2),
                                                 dc.ctypes.data_as(C.POINTER(C.c_double)), C.c_int(0))
    return dc, pdist(tnf, metric="euclidean")


def rank_stat(cov: np.ndarray, tnf: np.ndarray, lengths: np.ndarray):
    """-> (cov_ranks, kmer_ranks, stat), all condensed (i < j)."""
    dc, dk = distances(cov, tnf)
    return stat_from_distances(dc, dk, lengths, cov.shape[1] // 2)
