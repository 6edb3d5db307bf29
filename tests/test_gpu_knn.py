"""GPU parity: exact kNN graphs through the C ABI vs the exhaustive f64 oracle.

Tolerances (stated per SURVEY §8c).  The library selects K' > k candidates per row in FP32 and then
re-evaluates them with the reference's f64 arithmetic, truncated to f32 as embedder.rs:125,:170 do:
  * a reported distance must equal the oracle's f32 distance of that (row, neighbour) pair within
    TAU_REL = 2.4e-7 relative (2 f32 ulps: device vs glibc erfc/log/exp differ in the last f64 bits,
    which can straddle an f32 rounding boundary);
  * every returned neighbour must lie within TAU of the oracle's k-th distance, and the k-th distances
    must agree within TAU (ties at the boundary may swap);
  * tie-aware recall must be EXACTLY 1.0: the FP32 select phase is certified per row (margin between the k-th f64
    distance and the K'-th select distance against the row's measured select error), FP32 collisions of f64-distinct
    means / variances are patched in as extra candidates, and uncertified rows are re-evaluated exhaustively in f64
    (knn_rescore.cu, knn_exact.cu).  "Tie-aware": an oracle neighbour whose distance equals the k-th
    within TAU may be replaced by its tie partner (device vs glibc libm, last-bit differences).
"""
import os

import numpy as np
import pytest

import oracle
from rosella_b200 import gpu, synth

pytestmark = pytest.mark.gpu

TAU_REL = 2.4e-7
PAD = 0xFFFFFFFF


def _pair_dist(data, which, variant, i, j):
    if which == oracle.TNF_EUCLID:
        v = oracle.tnf_euclidean(data[i], data[j])
        return np.float32(3.402823466e+38) if np.isnan(v) else np.float32(v)
    if which == oracle.TNF:
        v = oracle.rho_flight(data[i], data[j]) if variant == oracle.FLIGHT else oracle.kmer_correlation(data[i], data[j])
        if np.isnan(v):
            return np.float32(0.0 if variant == oracle.RUST else 2.0)
    else:
        v = oracle.metabat(data[i], data[j], variant)
        if np.isnan(v):
            return np.float32(1.0)
    return np.float32(v)


def check_against_oracle(data, which, variant, k, idx, dist, q_begin=0, min_recall=1.0, tau_rel=TAU_REL):
    nq = idx.shape[0]
    qrows = np.arange(q_begin, q_begin + nq, dtype=np.uint64)
    oidx, odist, _ = oracle.knn(data, which, k, variant, qrows=qrows)
    # + 1e-15 absolute: a per-sample d is a difference of cdf values near 0.5, so device and glibc erfc disagree by
    # ~1e-16 absolute whatever d is (visible only on the 1e-10 distances of the near-duplicate tests)
    tau = lambda x: tau_rel * np.maximum(np.abs(x), 1e-30) + 1e-15
    hits = total = 0
    for r in range(nq):
        valid = idx[r] != PAD
        assert valid.sum() == (oidx[r] != PAD).sum()
        assert (idx[r][valid] != q_begin + r).all()                      # self excluded
        assert len(set(idx[r][valid].tolist())) == valid.sum()           # no duplicates
        d = dist[r][valid]
        assert (np.diff(d) >= 0).all()                                   # ascending
        same = idx[r] == oidx[r]
        assert np.all(np.abs(dist[r][same & valid] - odist[r][same & valid]) <= tau(odist[r][same & valid]))
        kth = odist[r][valid][-1] if valid.any() else 0.0
        for p in np.where(valid & ~same)[0]:
            od = _pair_dist(data, which, variant, q_begin + r, int(idx[r][p]))
            assert abs(od - dist[r][p]) <= tau(od), (r, p, od, dist[r][p])
            assert od <= kth + tau(kth), (r, p, od, kth)
        if valid.any():
            assert abs(d[-1] - kth) <= tau(kth)
        # tie-aware recall: an oracle neighbour counts as found if present, or if its distance ties the k-th
        got = set(idx[r][valid].tolist())
        for p in np.where(oidx[r] != PAD)[0]:
            total += 1
            hits += (int(oidx[r][p]) in got) or (odist[r][p] >= kth - tau(kth))
    recall = hits / max(total, 1)
    assert recall >= min_recall, recall
    exact = float((idx == oidx).mean())
    return recall, exact


@pytest.mark.parametrize("variant", [gpu.VARIANT_RUST, gpu.VARIANT_FLIGHT])
def test_knn_synthetic_3000(ctx, variant):
    tnf, cov, _ = synth.knn_tables(3000, 10, flight=variant == gpu.VARIANT_FLIGHT)
    (ti, td), (di, dd) = ctx.knn_build(tnf, cov, 30, variant)
    r1, e1 = check_against_oracle(tnf, oracle.TNF, variant, 30, ti, td)
    r2, e2 = check_against_oracle(cov, oracle.DEPTH, variant, 30, di, dd)
    print(f"variant {variant}: tnf recall {r1:.5f} exact-position {e1:.5f}; depth recall {r2:.5f} exact-position {e2:.5f}")
    assert e1 > 0.99 and e2 > 0.98


@pytest.mark.parametrize("variant", [gpu.VARIANT_RUST, gpu.VARIANT_FLIGHT])
@pytest.mark.parametrize("block,chunk,k,n,samples", [(256, 1024, 30, 3000, 6), (992, 992, 30, 3000, 6), (320, 352, 30, 3000, 6),
                                                     (256, 512, 100, 1500, 4), (224, 480, 200, 1100, 3),
                                                     (256, 512, 30, 1000, 131)])
def test_depth_symmetric_build(ctx, variant, block, chunk, k, n, samples):
    """Whole-graph coverage builds evaluate each unordered pair once (knn_depth_sym.cu): the lists must equal the batch
    kernel's bit for bit and pass the oracle checks, including rows in different blocks that share their variances
    (the equal-variance branch whose sign depends on which row is the query, coverage_calculator.rs:638-647), for
    K' = 64 / 128 / 256 and for sample counts that are staged in chunks."""
    tnf, cov, _ = synth.knn_tables(n, samples, seed=21, flight=variant == gpu.VARIANT_FLIGHT)
    a, b = 100, n - 500
    cov[b, 1::2] = cov[a, 1::2]                 # equal variances across blocks, different means
    cov[b + 1, 1::2] = cov[a, 1::2]
    cov[a + 1] = cov[n - 400]                   # duplicate rows across blocks
    cov[7] = 0.0
    cov[n - 1, 1::2] = 0.0
    ctx.knn_load(None, cov, variant)
    try:
        ctx.set_option("depth_build", 1)
        bi, bd, bnv = ctx.knn_search(gpu.GRAPH_DEPTH, k)
        ctx.set_option("depth_build", 0)
        ctx.set_option("sym_block", block)
        ctx.set_option("sym_chunk", chunk)
        launches0 = ctx.timings()["kernel_launches"]
        si, sd, snv = ctx.knn_search(gpu.GRAPH_DEPTH, k)
    finally:
        for key in ("depth_build", "sym_block", "sym_chunk"):
            ctx.set_option(key, 0)
    nb = -(-n // block)
    assert ctx.timings()["kernel_launches"] - launches0 >= 3 * nb                     # the block path did run
    assert np.array_equal(si, bi) and np.array_equal(sd, bd) and np.array_equal(snv, bnv)
    for r0 in (a - 10, b - 10, n - 10):
        check_against_oracle(cov, oracle.DEPTH, variant, k, si[r0:r0 + 10], sd[r0:r0 + 10], q_begin=r0)


def test_knn_sharded_query_ranges(ctx):
    """Rank shards [q_begin, q_end) that are not tile aligned concatenate to the full result."""
    tnf, cov, _ = synth.knn_tables(1500, 5, seed=11)
    ctx.knn_load(tnf, cov, gpu.VARIANT_RUST)
    full_i, full_d, nv = ctx.knn_search(gpu.GRAPH_DEPTH, 10)
    assert (nv == 10).all()
    parts = [ctx.knn_search(gpu.GRAPH_DEPTH, 10, b, e) for b, e in [(0, 37), (37, 1000), (1000, 1500)]]
    assert np.array_equal(np.concatenate([p[0] for p in parts]), full_i)
    assert np.array_equal(np.concatenate([p[1] for p in parts]), full_d)
    check_against_oracle(cov, oracle.DEPTH, oracle.RUST, 10, parts[1][0][:200], parts[1][1][:200], q_begin=37)
    ti, td, _ = ctx.knn_search(gpu.GRAPH_TNF, 10, 777, 901)
    check_against_oracle(tnf, oracle.TNF, oracle.RUST, 10, ti, td, q_begin=777)
    e_i, e_d, _ = ctx.knn_search(gpu.GRAPH_TNF, 10, 5, 5)
    assert e_i.shape == (0, 10)


def test_knn_tnf_euclidean(ctx, golden_dir):
    """flight's second TNF index (rosella/embedding.py:163, metrics.tnf_euclidean): Euclidean distance of the rows as
    passed, for both TNF flavours (CLR rows have mean 0; raw frequency rows do not), and the reference's own
    numba outputs for all pairs."""
    for flight in (False, True):
        tnf, _, _ = synth.knn_tables(2500, 2, seed=21, flight=flight)
        ctx.knn_load(tnf, None, gpu.VARIANT_FLIGHT if flight else gpu.VARIANT_RUST)
        idx, dist, nv = ctx.knn_search(gpu.GRAPH_TNF_EUCLID, 30)
        assert (nv == 30).all()
        r, e = check_against_oracle(tnf, oracle.TNF_EUCLID, oracle.FLIGHT, 30, idx, dist)
        assert e > 0.99
    freq = np.exp(tnf[:800]); freq /= freq.sum(axis=1, keepdims=True)      # uncentred rows: the mean term matters
    freq[:, 0] += np.linspace(0.0, 0.05, 800)
    ctx.knn_load(freq, None, gpu.VARIANT_FLIGHT)
    idx, dist, _ = ctx.knn_search(gpu.GRAPH_TNF_EUCLID, 15, 100, 500)
    check_against_oracle(freq, oracle.TNF_EUCLID, oracle.FLIGHT, 15, idx, dist, q_begin=100)
    g = np.load(os.path.join(golden_dir, "flight_metrics.npz"))
    x, ref = g["tnf_clr"], g["tnf_euclidean"]
    ctx.knn_load(x, None, gpu.VARIANT_FLIGHT)
    idx, dist, _ = ctx.knn_search(gpu.GRAPH_TNF_EUCLID, len(x) - 1)
    for i in range(len(x)):
        want = ref[i][idx[i]].astype(np.float32)
        assert np.all(np.abs(dist[i] - want) <= TAU_REL * np.abs(want) + 1e-12)


@pytest.mark.parametrize("k,n", [(1, 300), (48, 400), (50, 2100), (52, 700), (53, 500), (100, 700), (200, 600)])
def test_knn_k_range(ctx, k, n):
    tnf, cov, _ = synth.knn_tables(n, 3, seed=5 + k)
    (ti, td), (di, dd) = ctx.knn_build(tnf, cov, k)
    check_against_oracle(tnf, oracle.TNF, oracle.RUST, k, ti, td)
    check_against_oracle(cov, oracle.DEPTH, oracle.RUST, k, di, dd)


@pytest.mark.parametrize("n,s,k", [(700, 100, 50), (500, 131, 30), (300, 230, 10), (200, 400, 5)])
def test_knn_many_samples(ctx, n, s, k):
    """BASELINE.json configs[4] shape (100 samples, k = 50) and beyond: the depth kernel stages a reference tile
    in chunks of samples (4 x 25, 3 x 44, 2 x 115, 80 x 5 here) and carries the accumulators across them."""
    tnf, cov, _ = synth.knn_tables(n, s, seed=40 + s)
    ctx.knn_load(None, cov, gpu.VARIANT_RUST)
    di, dd, _ = ctx.knn_search(gpu.GRAPH_DEPTH, k)
    check_against_oracle(cov, oracle.DEPTH, oracle.RUST, k, di, dd)
    if s == 100:
        ctx.knn_load(None, cov, gpu.VARIANT_FLIGHT)
        di, dd, _ = ctx.knn_search(gpu.GRAPH_DEPTH, k)
        check_against_oracle(cov, oracle.DEPTH, oracle.FLIGHT, k, di, dd)


def test_knn_fewer_rows_than_k(ctx):
    tnf, cov, _ = synth.knn_tables(10, 2, seed=3)
    (ti, td), (di, dd) = ctx.knn_build(tnf, cov, 30)
    assert (ti[:, 9:] == PAD).all() and np.isinf(td[:, 9:]).all() and (ti[:, :9] != PAD).all()
    check_against_oracle(tnf, oracle.TNF, oracle.RUST, 30, ti, td)
    check_against_oracle(cov, oracle.DEPTH, oracle.RUST, 30, di, dd)


def test_knn_degenerate_rows(ctx):
    """Duplicates, all-zero coverage, constant TNF rows, a TNF row holding -inf (ln 0 in the CLR)."""
    tnf, cov, _ = synth.knn_tables(400, 4, seed=9)
    tnf[5] = tnf[4]; cov[5] = cov[4]            # duplicate contig
    cov[7] = 0.0                                 # absent everywhere
    cov[8, 1::2] = 0.0                           # zero variance
    tnf[9] = 0.0; tnf[10] = 0.0                  # constant rows: both norms 0 -> distance 0 between them
    tnf[11, 3] = -np.inf                         # flagged row: distance 1 to everything
    cov[12] = cov[13] * 1.0000001               # nearly identical coverage
    for variant in (gpu.VARIANT_RUST, gpu.VARIANT_FLIGHT):
        (ti, td), (di, dd) = ctx.knn_build(tnf, cov, 30, variant)
        check_against_oracle(tnf, oracle.TNF, variant, 30, ti, td)
        check_against_oracle(cov, oracle.DEPTH, variant, 30, di, dd)
    (ti, td), _ = ctx.knn_build(tnf, cov, 30, gpu.VARIANT_RUST)
    assert ti[9, 0] == 10 and td[9, 0] == 0.0 and ti[4, 0] == 5
    assert (td[11] == 1.0).all() and list(ti[11]) == list(range(11)) + list(range(12, 31))   # ties by index


def test_flight_golden_vectors_through_the_gpu(ctx, golden_dir):
    """All-neighbour lists (k = n-1) reproduce the reference's own numba outputs."""
    g = np.load(os.path.join(golden_dir, "flight_metrics.npz"))
    for s in (1, 3, 10):
        cov, ref = g[f"cov_s{s}"], g[f"metabat_nn_s{s}"]
        n = len(cov)
        ctx.knn_load(None, cov, gpu.VARIANT_FLIGHT)
        idx, dist, _ = ctx.knn_search(gpu.GRAPH_DEPTH, n - 1)
        for i in range(n):
            want = ref[i][idx[i]].astype(np.float32)
            assert np.all(np.abs(dist[i] - want) <= TAU_REL * np.abs(want)), (s, i)
    x, ref = g["tnf_clr"], g["rho"]
    ctx.knn_load(x, None, gpu.VARIANT_FLIGHT)
    idx, dist, _ = ctx.knn_search(gpu.GRAPH_TNF, len(x) - 1)
    for i in range(len(x)):
        want = ref[i][idx[i]].astype(np.float32)
        assert np.all(np.abs(dist[i] - want) <= TAU_REL * np.abs(want) + 1e-12)


def test_errors(ctx):
    tnf, cov, _ = synth.knn_tables(64, 2, seed=1)
    bad = cov.copy(); bad[3, 1] = np.nan
    with pytest.raises(gpu.RosellaGpuError, match="non-finite"):
        ctx.knn_load(tnf, bad)
    with pytest.raises(gpu.RosellaGpuError, match="no tables loaded"):
        ctx.knn_search(gpu.GRAPH_TNF, 5, 0, 10)
    ctx.knn_load(tnf, None)
    with pytest.raises(gpu.RosellaGpuError, match="no coverage table"):
        ctx.knn_search(gpu.GRAPH_DEPTH, 5, 0, 10)
    with pytest.raises(gpu.RosellaGpuError, match="k=0"):
        ctx.knn_search(gpu.GRAPH_TNF, 0, 0, 10)
    with pytest.raises(gpu.RosellaGpuError, match="outside"):
        ctx.knn_search(gpu.GRAPH_TNF, 5, 0, 65)


def test_graph_ops(ctx):
    tnf, cov, _ = synth.knn_tables(500, 3, seed=2)
    (ti, _), (di, _) = ctx.knn_build(tnf, cov, 12)
    for keep_n in (0, 5, 12):
        assert np.array_equal(ctx.graph_mutual(di, keep_n), oracle.mutual_knn(di, keep_n))
        assert np.array_equal(ctx.graph_intersect(di, ti, keep_n), oracle.intersect(di, ti, keep_n))


@pytest.mark.parametrize("keep_n_edges", [0, 3, 8, 12])
def test_build_mutual_kgraph_on_device(ctx, keep_n_edges):
    """rg_knn_build_mutual (intersect + mutual_knn applied on the device, CSR out) against the oracle's restatement of
    embedder.rs:68-84 fed with the lists of rg_knn_build."""
    tnf, cov, _ = synth.knn_tables(1200, 3, seed=12)
    (ti, _), (di, dd) = ctx.knn_build(tnf, cov, 12)
    off, nbr, w, disc = ctx.knn_build_mutual(tnf, cov, 12, keep_n_edges)
    o_off, o_nbr, o_w, o_disc = oracle.build_mutual_kgraph(di, dd, ti, keep_n_edges)
    assert np.array_equal(off, o_off) and np.array_equal(nbr, o_nbr) and np.array_equal(w, o_w)
    assert np.array_equal(disc, o_disc)
    if keep_n_edges == 0:
        assert 0 < off[-1] < di.size                  # the filters did drop edges
    if keep_n_edges >= 12:
        assert off[-1] == di.size and len(disc) == 0


def test_config1_full_build_100k(ctx):
    """BASELINE.json configs[1]: 100k contigs x (136 TNF + 10 samples), k = 30, the whole graph on one B200 through
    rg_knn_build; three blocks of rows are checked against the exhaustive f64 oracle, all rows against the
    size-independent properties of a kNN list."""
    n, s, k = 100_000, 10, 30
    tnf, cov, _ = synth.knn_tables(n, s)
    (ti, td), (di, dd) = ctx.knn_build(tnf, cov, k)
    for idx, dist in ((ti, td), (di, dd)):
        assert idx.shape == (n, k) and (idx != PAD).all() and (idx < n).all()
        assert (idx != np.arange(n, dtype=np.uint32)[:, None]).all()             # self excluded
        assert (np.diff(dist, axis=1) >= 0).all()                                  # ascending
        srt = np.sort(idx, axis=1)
        assert (srt[:, 1:] != srt[:, :-1]).all()                                   # no duplicates
    for qb in (0, 50_016, n - 64):
        r1, _ = check_against_oracle(tnf, oracle.TNF, oracle.RUST, k, ti[qb:qb + 64], td[qb:qb + 64], q_begin=qb)
        r2, _ = check_against_oracle(cov, oracle.DEPTH, oracle.RUST, k, di[qb:qb + 64], dd[qb:qb + 64], q_begin=qb)
        assert r1 == 1.0 and r2 == 1.0


def test_config2_query_batch_1m(ctx):
    """BASELINE.json configs[2] at full reference size: 1M contigs x (136 TNF + 50 samples), k = 30.  A batch of query
    rows is searched against all 1M references (what one bench step does); 32 of them are checked against the
    exhaustive f64 oracle, and a shifted batch must reproduce the overlapping rows bit for bit."""
    n, s, k = 1_000_000, 50, 30
    tnf, cov, _ = synth.knn_tables(n, s)
    ctx.knn_load(tnf, cov, gpu.VARIANT_RUST)
    qb = 700_000
    for which, data, ow in ((gpu.GRAPH_DEPTH, cov, oracle.DEPTH), (gpu.GRAPH_TNF, tnf, oracle.TNF)):
        idx, dist, nv = ctx.knn_search(which, k, qb, qb + 1024)
        assert (nv == k).all() and (np.diff(dist, axis=1) >= 0).all()
        recall, _ = check_against_oracle(data, ow, oracle.RUST, k, idx[:32], dist[:32], q_begin=qb)
        assert recall == 1.0
        idx2, dist2, _ = ctx.knn_search(which, k, qb + 96, qb + 96 + 512)          # another CTA / panel alignment
        assert np.array_equal(idx2, idx[96:608]) and np.array_equal(dist2, dist[96:608])


@pytest.mark.parametrize("s", [1, 3, 10])
def test_rust_variant_golden_vectors_through_the_gpu(ctx, golden_dir, s):
    """RG_VARIANT_RUST distances (the variant bench.py times) against the reference's own numba metabat_distance_nn run
    on (mean - 9e-7, var - 9e-7): the same function as coverage_calculator.rs:588-666 wherever flight's clamp and the
    equal-variance sign rule cannot fire (oracle/make_golden.py: make_rust_pin)."""
    g = np.load(os.path.join(golden_dir, "rust_pin.npz"))
    cov, ref = g[f"cov_s{s}"], g[f"metabat_rust_mainbranch_s{s}"]
    main = g[f"noclamp_s{s}"] & g[f"noneq_s{s}"]
    eq = g[f"noclamp_s{s}"] & ~g[f"noneq_s{s}"]
    n = len(cov)
    ctx.knn_load(None, cov, gpu.VARIANT_RUST)
    idx, dist, _ = ctx.knn_search(gpu.GRAPH_DEPTH, n - 1)
    checked = 0
    for i in range(n):
        want = ref[i][idx[i]].astype(np.float32)
        m = main[i][idx[i]]
        assert np.all(np.abs(dist[i][m] - want[m]) <= TAU_REL * np.abs(want[m])), (s, i)
        e = eq[i][idx[i]]
        assert np.all((np.abs(dist[i][e] - want[e]) <= TAU_REL * np.abs(want[e])) | (dist[i][e] == 1.0))
        checked += int(m.sum())
    assert checked > 700


def _exact_positions(data, which, variant, k, idx, dist):
    oi, od, _ = oracle.knn(data, which, k, variant)
    return float((idx == oi).mean()), oi, od


@pytest.mark.parametrize("samples", [1, 2, 6])
def test_certificate_near_duplicate_clusters(ctx, samples):
    """Thousands of near-duplicate coverage rows: inside a cluster the distances are ~1e-5, far below the
    absolute error of the FP32 select (6.7e-7 per sample from the erf fit alone), and the cluster is larger than K'.
    The certificate must reject those rows and the exhaustive f64 pass must return the oracle's lists."""
    rng = np.random.default_rng(100 + samples)
    n, k = 4000, 30
    _, cov, _ = synth.knn_tables(n, samples, seed=50 + samples)
    for c in range(6):                                     # 6 clusters of 300 rows, relative noise 1e-5
        base = cov[c].copy()
        base[0::2] = np.maximum(base[0::2], 3.0 + c)
        base[1::2] = base[0::2] * 2.5
        rows = np.arange(500 + 300 * c, 800 + 300 * c)
        cov[rows] = base * (1.0 + 1e-5 * rng.standard_normal((300, 2 * samples)))
    ctx.knn_load(None, cov, gpu.VARIANT_RUST)
    di, dd, _ = ctx.knn_search(gpu.GRAPH_DEPTH, k)
    t = ctx.timings()
    assert t["rows_flagged"] >= 1000 and t["rows_exact"] == t["rows_flagged"], t
    # inside a cluster every per-sample d is a difference of cdf values that cancels to ~1e-10 of their size, so the
    # f64 metric itself is conditioned at ~1e-6 relative there (device erfc vs glibc erfc): tau_rel 4e-6 for this data
    recall, _ = check_against_oracle(cov, oracle.DEPTH, oracle.RUST, k, di, dd, tau_rel=4e-6)
    assert recall == 1.0
    # the hazard is real: without the certificate the same search loses neighbours inside the clusters
    ctx.set_option("certify", 0)
    try:
        ctx.knn_load(None, cov, gpu.VARIANT_RUST)
        ui, ud, _ = ctx.knn_search(gpu.GRAPH_DEPTH, k)
    finally:
        ctx.set_option("certify", 1)
    oi, _, _ = oracle.knn(cov, oracle.DEPTH, k, oracle.RUST)
    lost = sum(len(set(oi[r]) - set(ui[r])) for r in range(500, 2300))
    print(f"S={samples}: flagged {t['rows_flagged']}, neighbours lost without the certificate: {lost}")


def test_fp32_collisions_of_distinct_f64_values_are_patched(ctx):
    """Means (variances) that differ in f64 but round to one FP32 value flip `a_mean != b_mean`
    (coverage_calculator.rs:619) / the p1-p2 choice (:638) in the FP32 select: the pair must still come out exactly as
    the f64 metric ranks it (knn_exact.cu adds it as an extra candidate of both rows)."""
    n, s, k = 3000, 5, 30
    _, cov, _ = synth.knn_tables(n, s, seed=77)
    a = np.arange(100, 300)
    b = a + 1000
    cov[b] = cov[a]
    cov[b, 0::2] *= 1.0 + 3e-10                 # means: FP32-equal, f64-distinct -> true distance ~0, select says 1.0
    c, d = np.arange(400, 500), np.arange(1700, 1800)
    cov[d, 1::2] = np.maximum(cov[c, 1::2], 2.0) * (1.0 + 2e-10)   # variances collide, means differ
    cov[c, 1::2] = np.maximum(cov[c, 1::2], 2.0)
    dense = np.arange(2400, 2600)               # one FP32 value, 200 f64 values: longer than the pair enumeration allows
    cov[dense] = cov[2399]
    cov[dense, 0] = 7.25 * (1.0 + 1e-12 * np.arange(200))
    ctx.knn_load(None, cov, gpu.VARIANT_RUST)
    t0 = ctx.timings()
    assert t0["extra_pairs"] >= 2 * 200
    di, dd, _ = ctx.knn_search(gpu.GRAPH_DEPTH, k)
    t = ctx.timings()
    assert t["rows_exact"] >= 200               # the dense group went through the exhaustive pass
    recall, _ = check_against_oracle(cov, oracle.DEPTH, oracle.RUST, k, di, dd)
    assert recall == 1.0
    # the twin with the colliding means is the nearest neighbour in f64 (the select kernel alone would rank it last)
    assert (di[a, 0] == b).mean() > 0.9
    # same through the symmetric block build
    try:
        ctx.set_option("depth_build", 2); ctx.set_option("sym_block", 512); ctx.set_option("sym_chunk", 1024)
        si, sd, _ = ctx.knn_search(gpu.GRAPH_DEPTH, k)
    finally:
        for key in ("depth_build", "sym_block", "sym_chunk"):
            ctx.set_option(key, 0)
    assert np.array_equal(si, di) and np.array_equal(sd, dd)


def test_certificate_tnf_duplicates(ctx):
    """A cluster of identical and near-identical TNF rows larger than K' = 48: ties at distance 0 / ~1e-9."""
    rng = np.random.default_rng(5)
    n, k = 3000, 30
    tnf, _, _ = synth.knn_tables(n, 2, seed=91)
    tnf[1000:1100] = tnf[999]                                   # 100 identical rows
    tnf[1200:1400] = tnf[1199] + 1e-7 * rng.standard_normal((200, 136))
    ctx.knn_load(tnf, None, gpu.VARIANT_RUST)
    ti, td, _ = ctx.knn_search(gpu.GRAPH_TNF, k)
    t = ctx.timings()
    assert t["rows_flagged"] >= 100 and t["rows_exact"] == t["rows_flagged"]
    recall, _ = check_against_oracle(tnf, oracle.TNF, oracle.RUST, k, ti, td)
    assert recall == 1.0
    assert list(ti[1000, :5]) == [999, 1001, 1002, 1003, 1004]  # ties at 0 by index


def test_negative_mean_is_rejected(ctx):
    _, cov, _ = synth.knn_tables(64, 2, seed=1)
    cov[5, 2] = -0.5
    with pytest.raises(gpu.RosellaGpuError, match="negative mean"):
        ctx.knn_load(None, cov)


def test_depth_graph_with_the_minhash_factor(ctx):
    """DepthDistance::eval (embedder.rs:133-171) = metabat x (1 - min_jaccard): with sketches loaded the coverage graph
    must equal the oracle's exhaustive kNN of the PRODUCT metric — pairs that share hashes move up (or to distance 0 for
    identical sketches), pairs of empty sketches sit at distance 0 (sketch_distances.rs:109-111), everything else is
    unchanged — through the batch kernel, the symmetric build and a query sub-range."""
    n, s, k = 3000, 4, 30
    _, cov, _ = synth.knn_tables(n, s, seed=61)
    hashes, offs = synth.sketches(n, seed=9, mean_size=40, share_frac=0.25, empty_frac=0.004)
    lists = [hashes[int(offs[i]):int(offs[i + 1])] for i in range(n)]
    ctx.knn_load(None, cov, gpu.VARIANT_RUST)
    plain_i, plain_d, _ = ctx.knn_search(gpu.GRAPH_DEPTH, k)
    ctx.knn_load_sketches(hashes, offs, 0.1)
    assert ctx.timings()["extra_pairs"] > 500
    di, dd, nv = ctx.knn_search(gpu.GRAPH_DEPTH, k)
    oi, od = oracle.knn_depth_sketch(cov, hashes, offs, k, oracle.RUST, 0.1)
    assert (nv == k).all() and (np.diff(dd, axis=1) >= 0).all()
    assert (di != plain_i).any(axis=1).sum() > 100            # the factor does change the graph
    changed = 0
    for r in range(n):
        if np.array_equal(di[r], oi[r]):
            assert np.all(np.abs(dd[r] - od[r]) <= TAU_REL * np.abs(od[r]) + 1e-15)
            continue
        changed += 1
        kth = od[r, -1]
        got = set(di[r].tolist())
        for p in range(k):                                     # tie-aware: every oracle neighbour is present or ties the k-th
            assert int(oi[r, p]) in got or od[r, p] >= kth * (1 - TAU_REL), (r, p)
        for p in range(k):                                     # and every reported distance is the product metric's
            want = oracle.depth_distance_sketch(cov[r], cov[di[r, p]], lists[r], lists[di[r, p]], oracle.RUST, 0.1)
            assert abs(want - dd[r, p]) <= TAU_REL * abs(want) + 1e-15
    assert changed < 0.01 * n
    empty = np.where(np.diff(offs.astype(np.int64)) == 0)[0]
    assert len(empty) >= 2 and dd[empty[0], 0] == 0.0 and di[empty[0], 0] in empty
    try:                                                       # symmetric block build and a sub-range: same lists
        ctx.set_option("depth_build", 2); ctx.set_option("sym_block", 512); ctx.set_option("sym_chunk", 1024)
        si, sd, _ = ctx.knn_search(gpu.GRAPH_DEPTH, k)
    finally:
        for key in ("depth_build", "sym_block", "sym_chunk"):
            ctx.set_option(key, 0)
    assert np.array_equal(si, di) and np.array_equal(sd, dd)
    pi, pd_, _ = ctx.knn_search(gpu.GRAPH_DEPTH, k, 1000, 1700)
    assert np.array_equal(pi, di[1000:1700]) and np.array_equal(pd_, dd[1000:1700])
    ctx.knn_load(None, cov, gpu.VARIANT_RUST)                 # a new load drops the sketches
    again_i, again_d, _ = ctx.knn_search(gpu.GRAPH_DEPTH, k)
    assert np.array_equal(again_i, plain_i) and np.array_equal(again_d, plain_d)
    bad = hashes.copy(); bad[1], bad[0] = bad[0], bad[1]
    with pytest.raises(gpu.RosellaGpuError, match="strictly increasing"):
        ctx.knn_load_sketches(bad, offs, 0.1)


def test_full_n_oracle_check_20k(ctx):
    """every row of a 20k x (136 + 10) build through rg_knn_build — the symmetric block path for the coverage
    graph (three row blocks), the tensor-core path for the TNF graph — against the exhaustive f64 oracle: tie-aware
    recall exactly 1.0 over all 600 000 edges of each graph."""
    n, s, k = 20_000, 10, 30
    tnf, cov, _ = synth.knn_tables(n, s, seed=2026)
    (ti, td), (di, dd) = ctx.knn_build(tnf, cov, k)
    r1, e1 = check_against_oracle(tnf, oracle.TNF, oracle.RUST, k, ti, td)
    r2, e2 = check_against_oracle(cov, oracle.DEPTH, oracle.RUST, k, di, dd)
    t = ctx.timings()
    print(f"20k full-N: tnf recall {r1} positions {e1:.6f}; depth recall {r2} positions {e2:.6f}; flagged {t['rows_flagged']}")
    assert r1 == 1.0 and r2 == 1.0 and e1 > 0.999 and e2 > 0.999
