"""flight's rank statistic (rg_rank_stat) against the reference's own output (tests/golden/flight_rankstat.npz, made by
oracle/make_golden.py from flight/flight/distance.py) and against the CPU restatement (oracle/rankstat.py)."""
import os

import numpy as np
import pytest

from oracle import rankstat
from rosella_b200 import synth

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("s", [1, 3, 10])
def test_rank_stat_reproduces_the_reference(ctx, golden_dir, s):
    g = np.load(os.path.join(golden_dir, "flight_rankstat.npz"))
    cov, tnf, lengths = g[f"cov_s{s}"], g[f"tnf_s{s}"], g[f"len_s{s}"]
    stat, idx, sk = ctx.rank_stat(tnf, cov, lengths, k=10)
    _, _, want = rankstat.rank_stat(cov, tnf, lengths)
    # Ranks are decided by exact ties.  With few samples whole groups of pairs tie structurally (a present row against
    # every absent row), and whether d(a, b) and d(b, a) agree in the last bit differs between numba's fastmath code,
    # glibc and the device's libm — a group is then averaged as one streak or as two.  Everything else is equal to
    # 1e-12; a split group moves its members by at most half the group (0.3 % of the 4005 ranks here).
    ref = g[f"stat_s{s}"]
    for other in (want, ref):
        assert np.median(np.abs(stat - other) / other) < 1e-12 and np.abs(stat - other).max() < 0.02
    if s == 10:
        assert np.allclose(stat, want, rtol=1e-12, atol=0)
    n = len(cov)
    full = np.full((n, n), np.inf)
    iu = np.triu_indices(n, 1)
    full[iu] = stat
    full.T[iu] = stat
    for r in range(n):
        order = np.lexsort((np.arange(n), full[r]))[:10]
        assert list(order) == list(idx[r]) and np.array_equal(full[r][order], sk[r])


def test_rank_stat_5000_against_the_restatement(ctx):
    """equal to scipy.pdist + _fractional_rank at N = 5000 (12.5M pairs)."""
    n, s = 5000, 10
    tnf, cov, _ = synth.knn_tables(n, s, seed=33, flight=True)
    rng = np.random.default_rng(1)
    lengths = np.rint(np.exp(rng.uniform(np.log(2000.0), np.log(500000.0), n)))
    stat, idx, sk = ctx.rank_stat(tnf, cov, lengths, k=30)
    cr, kr, want = rankstat.rank_stat(cov, tnf, lengths)
    close = np.isclose(stat, want, rtol=1e-11, atol=0)
    # device erfc / log / exp differ from glibc's in the last bits: two coverage distances closer than that may swap ranks
    assert close.mean() > 0.9999 and np.abs(stat - want).max() < 1e-5
    t = ctx.timings()
    print(f"n={n}: distances {t['select_ms']:.1f} ms, ranks + statistic {t['rescore_ms']:.1f} ms")
    assert (np.diff(sk, axis=1) >= 0).all() and (idx != np.arange(n, dtype=np.uint32)[:, None]).all()


def test_out_of_memory_is_an_error_return(ctx):
    """RG_ERR_OOM, not a crash: 200 000 rows need 9 arrays of 2e10 doubles; the handle stays usable afterwards."""
    from rosella_b200 import gpu
    n = 200_000
    tnf = np.zeros((n, 136)); cov = np.ones((n, 2)); lengths = np.full(n, 5000.0)
    with pytest.raises(gpu.RosellaGpuError) as e:
        ctx.rank_stat(tnf, cov, lengths, k=5, condensed=False)
    assert e.value.code == -4 and "cudaMalloc" in str(e.value)
    t, c, _ = synth.knn_tables(300, 2, seed=3, flight=True)
    stat, _, _ = ctx.rank_stat(t, c, np.full(300, 1000.0))
    assert np.isfinite(stat).all()
