"""condensed_pairwise_distance and silhouette_score (src/clustering/cluster_utils.rs:7-120) on the
device against the oracle's restatement (which agrees with scikit-learn's silhouette_samples to 1e-15 on clean data)."""
import numpy as np
import pytest

import oracle
from rosella_b200 import gpu

pytestmark = pytest.mark.gpu


def _blobs(rng, n_per, centres, dim):
    x = np.concatenate([rng.normal(0.0, 0.4, (n_per, dim)) + np.asarray(c, float) for c in centres])
    return x, np.repeat(np.arange(len(centres)), n_per).astype(np.int32)


def test_condensed_euclidean_is_bit_exact(ctx):
    rng = np.random.default_rng(1)
    x = rng.normal(size=(700, 5))
    x[13] = x[12]                       # distance 0
    x[40, 2] = np.nan                   # NaN -> 0 (:24-26)
    got = ctx.condensed_euclidean(x)
    assert np.array_equal(got, oracle.condensed_euclidean(x))
    assert got[np.isnan(got)].size == 0


@pytest.mark.parametrize("dim", [2, 7])
def test_silhouette_matches_the_restatement(ctx, dim):
    rng = np.random.default_rng(dim)
    centres = rng.normal(0.0, 3.0, (9, dim))
    x, lab = _blobs(rng, 120, centres, dim)
    perm = rng.permutation(len(x))
    x, lab = x[perm], lab[perm]
    lab[rng.random(len(lab)) < 0.05] = -1           # outliers belong to no cluster
    lab[lab == 8] = -1; lab[5] = 8                  # a cluster of one: a = 0 (:66-69)
    score, pp = ctx.silhouette(x, lab, 9)
    want, wpp = oracle.silhouette(oracle.condensed_euclidean(x), lab, 9)
    assert np.allclose(pp, wpp, rtol=1e-12, atol=1e-15) and abs(score - want) <= 1e-12 * abs(want)
    assert (pp[lab < 0] == 0).all() and pp[5] == 1.0
    assert ctx.silhouette(x, np.zeros(len(x), np.int32), 1)[0] == 0.0  # a single cluster scores 0 (:39-41)
    with pytest.raises(gpu.RosellaGpuError, match="outside"):
        ctx.silhouette(x, np.full(len(x), 9, np.int32), 9)


def test_silhouette_without_the_distance_matrix_at_200k(ctx):
    """200k points: the reference's condensed vector would hold 2e10 doubles (160 GB); here nothing O(n^2) is stored."""
    rng = np.random.default_rng(3)
    centres = rng.normal(0.0, 6.0, (50, 2))
    x, lab = _blobs(rng, 4000, centres, 2)
    score, pp = ctx.silhouette(x, lab, 50)
    t = ctx.timings()
    assert 0.2 < score < 1.0 and np.isfinite(pp).all()
    sub = rng.choice(len(x), 40, replace=False)     # spot check against direct sums
    for i in sub:
        d = np.sqrt(((x - x[i]) ** 2).sum(axis=1))
        a = d[lab == lab[i]].sum() / (np.sum(lab == lab[i]) - 1)
        b = min(d[lab == c].mean() for c in range(50) if c != lab[i])
        assert abs(pp[i] - (b - a) / max(a, b)) < 1e-10
    print(f"silhouette of 200k points, 50 clusters: {t['select_ms']:.1f} ms")
