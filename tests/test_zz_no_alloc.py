"""rg_graph_mutual_static — the allocation-free variant of rg_graph_mutual (device staging in a fixed static
__device__ array inside the library) — against rg_graph_mutual through the C ABI.  Integer work: bit-exact.

Collected last on purpose: a failure here hides no other test.  The static array is shared by
every call on the GPU and keeps the previous call's contents, so the calls come in a row with shrinking and changing
inputs: stale lists or a stale mask from the call before would show up as a difference.
"""
import numpy as np
import pytest

import oracle
from rosella_b200 import gpu

pytestmark = pytest.mark.gpu

PAD = 0xFFFFFFFF
MAX_EDGES = 1 << 20          # RG_STATIC_MAX_EDGES, include/rosella_gpu.h


def _lists(n, k, seed, pad_frac=0.05):
    """kNN-shaped index lists: no self edges, some reciprocal edges, padded tails on a few rows."""
    rng = np.random.default_rng(seed)
    idx = rng.integers(0, n - 1, size=(n, k), dtype=np.int64)
    idx += idx >= np.arange(n)[:, None]                       # skip the row itself
    # plant reciprocal edges so that the mask has both outcomes past keep_n_edges
    rows = rng.integers(0, n, size=n // 2)
    for r in rows:
        nb = idx[r, k - 1]
        idx[nb, rng.integers(0, k)] = r
    idx = idx.astype(np.uint32)
    short = rng.random(n) < pad_frac
    cut = rng.integers(0, k, size=n)
    idx[short[:, None] & (np.arange(k)[None, :] >= cut[:, None])] = PAD
    return idx


def test_static_matches_allocating_entry_point_on_calls_in_a_row(ctx):
    cases = [(4096, 32, 5, 1), (1500, 20, 0, 2), (1500, 20, 20, 3), (37, 7, 2, 4), (MAX_EDGES // 64, 64, 8, 5), (300, 12, 3, 6)]
    for n, k, keep_n, seed in cases:
        idx = _lists(n, k, seed)
        want = ctx.graph_mutual(idx, keep_n)
        got = ctx.graph_mutual_static(idx, keep_n)
        assert got.dtype == np.uint8 and got.shape == want.shape
        assert np.array_equal(got, want), (n, k, keep_n, int((got != want).sum()))
        # the same input again: the array now holds this call's own lists and mask
        assert np.array_equal(ctx.graph_mutual_static(idx, keep_n), want)
    # and once against the oracle's restatement of mutual_knn (nearest_neighbour_graph.rs:12-35)
    idx = _lists(800, 16, 9)
    assert np.array_equal(ctx.graph_mutual_static(idx, 4), oracle.mutual_knn(idx, 4))


def test_static_on_a_second_handle_and_size_limit(ctx):
    idx = _lists(2000, 24, 11)
    want = ctx.graph_mutual(idx, 6)
    other = gpu.Context(0)                                    # another handle on the same GPU shares the array
    try:
        assert np.array_equal(other.graph_mutual_static(idx, 6), want)
        assert np.array_equal(ctx.graph_mutual_static(idx[:100], 6), ctx.graph_mutual(idx[:100], 6))
        big = np.zeros((MAX_EDGES // 32 + 1, 32), np.uint32)
        with pytest.raises(gpu.RosellaGpuError, match="RG_STATIC_MAX_EDGES"):
            other.graph_mutual_static(big, 4)
        # the refused call left the handle and the array usable
        assert np.array_equal(other.graph_mutual_static(idx, 6), want)
    finally:
        other.close()
