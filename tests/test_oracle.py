"""CPU tests: the oracle against the golden vectors (tests/golden/, generated from the reference by
oracle/make_golden.py) and against the known-answer values of SURVEY.md §4."""
import os

import numpy as np
import pytest

import oracle


@pytest.fixture(scope="module")
def tnf_fix(golden_dir):
    return np.load(os.path.join(golden_dir, "tnf_fixtures.npz"))


@pytest.fixture(scope="module")
def flight(golden_dir):
    return np.load(os.path.join(golden_dir, "flight_metrics.npz"))


def test_canonical_kmers_known_answers(tnf_fix):
    names = oracle.canonical_kmers()
    assert len(names) == 136 and len(set(names)) == 136
    assert names[:12] == ["AAAA", "AAAC", "AAAG", "AAAT", "AACA", "AACC", "AACG", "AACT", "AAGA", "AAGC", "AAGG", "AAGT"]
    assert names[-6:] == ["TCAA", "TCCA", "TCGA", "TGAA", "TGCA", "TTAA"]
    comp = str.maketrans("ACGT", "TGCA")
    assert sum(1 for k in names if k == k[::-1].translate(comp)) == 16     # palindromes
    assert names == list(tnf_fix["canonical"])
    lut = oracle.code_to_bin()
    assert lut.max() == 135 and len(np.unique(lut)) == 136


def test_tnf_counts_match_reference_model(tnf_fix):
    for fast in (False, True):
        counts, nk, freqs = oracle.tnf_count(tnf_fix["seq"], tnf_fix["offsets"], fast=fast)
        assert np.array_equal(counts, tnf_fix["counts"])
        assert np.array_equal(nk, tnf_fix["n_kmers"])
    # SURVEY §4 known answers on tests/data/ben/random10000.fna
    assert nk[0] == 9997 and list(counts[0, :6]) == [88, 87, 65, 71, 89, 104]
    assert counts[0].max() == 104 and counts[0].min() == 27
    assert nk[1] == 1997 and list(counts[1, :6]) == [17, 12, 13, 25, 15, 15]
    assert all(nk[2:11] == 4)
    # 0/0 -> NaN row, as kmer_counting.rs:104-107
    empty = np.where(nk == 0)[0]
    assert len(empty) >= 3 and np.isnan(freqs[empty]).all()
    ok = nk > 0
    assert np.allclose(freqs[ok].sum(axis=1), 1.0)


def test_normalize_semantics():
    assert oracle.normalize(b"acgtnACGTN-") == b"ACGTNACGTN-"
    assert oracle.normalize(b"uU.~ \t\r\nRYKMbdhv*") == b"TT--NNNNNNNNN"


@pytest.mark.parametrize("s", [1, 3, 10])
def test_metabat_flight_variant_is_pinned(flight, s):
    cov = flight[f"cov_s{s}"]
    ref = flight[f"metabat_nn_s{s}"]
    got = np.array([[oracle.metabat(a, b, oracle.FLIGHT) for b in cov] for a in cov])
    # numba compiles the reference with fastmath; 1e-12 relative is far below any f32 effect
    assert np.allclose(got, ref, rtol=1e-12, atol=0)
    # coverage_distance = same metric except identical rows -> 0 (metrics.py:429-431)
    cd = flight[f"coverage_distance_s{s}"]
    same = np.array([[np.array_equal(a, b) for b in cov] for a in cov])
    assert np.allclose(got[~same], cd[~same], rtol=1e-12, atol=0)
    assert (cd[same] == 0).all()


def test_rho_and_euclid_pinned(flight):
    x = flight["tnf_clr"]
    rho = np.array([[oracle.rho_flight(a, b) for b in x] for a in x])
    eu = np.array([[oracle.tnf_euclidean(a, b) for b in x] for a in x])
    off = ~np.eye(len(x), dtype=bool)
    assert np.allclose(rho[off], flight["rho"][off], rtol=1e-12, atol=1e-15)
    assert np.allclose(eu, flight["tnf_euclidean"], rtol=1e-12, atol=1e-13)
    # Rust proportionality distance is exactly half of flight's rho
    kc = np.array([[oracle.kmer_correlation(a, b) for b in x] for a in x])
    assert np.allclose(2 * kc[off], flight["rho"][off], rtol=1e-12, atol=1e-15)


def test_rust_variant_semantics(flight):
    cov = flight["cov_s3"]
    d = np.array([[oracle.metabat(a, b, oracle.RUST) for b in cov] for a in cov])
    assert ((d >= 0) & (d <= 1)).all() and not np.isnan(d).any()
    assert d[0, 1] == 1.0 and d[6, 6] == 1.0          # identical rows: no contributing sample (:665)
    # equal-variance branch keeps the sign (:647): a negative d makes the pair NaN -> 1.0, so the
    # Rust metric is not symmetric
    assert (d != d.T).any()
    a = np.array([5.0, 0.5]); b = np.array([9.0, 0.25])       # both variances clamp to 1
    assert oracle.metabat(a, b, oracle.RUST) < 1.0 and oracle.metabat(b, a, oracle.RUST) == 1.0


def test_kmer_correlation_guards():
    z = np.zeros(136); c = np.ones(136); r = np.random.default_rng(0).normal(size=136)
    assert oracle.kmer_correlation(z, c) == 0.0           # both centred norms are 0 (:383)
    assert oracle.kmer_correlation(z, r) == 1.0           # dot == 0 (:385)
    bad = r.copy(); bad[3] = -np.inf
    assert oracle.kmer_correlation(bad, r) == 1.0         # NaN -> 1 (:410)
    assert abs(oracle.kmer_correlation(r, r)) < 1e-15


def test_exhaustive_knn_matches_bruteforce():
    rng = np.random.default_rng(7)
    x = rng.normal(size=(60, 136))
    idx, dist, d64 = oracle.knn(x, oracle.TNF, 5)
    full = np.array([[oracle.kmer_correlation(a, b) for b in x] for a in x]).astype(np.float32)
    np.fill_diagonal(full, np.inf)
    for i in range(60):
        order = np.lexsort((np.arange(60), full[i]))[:5]
        assert list(order) == list(idx[i])
        assert np.array_equal(full[i][order], dist[i])
    # fewer candidates than k: padding
    idx, dist, _ = oracle.knn(x[:3], oracle.TNF, 5)
    assert (idx[:, 2:] == 0xFFFFFFFF).all() and np.isinf(dist[:, 2:]).all()


def test_graph_ops():
    idx = np.array([[1, 2], [0, 2], [1, 0xFFFFFFFF]], np.uint32)
    keep = oracle.mutual_knn(idx, 0)
    assert keep.tolist() == [[1, 0], [1, 1], [1, 0]]
    assert oracle.mutual_knn(idx, 1).tolist() == [[1, 0], [1, 1], [1, 0]]
    other = np.array([[2, 1], [2, 0xFFFFFFFF], [0, 1]], np.uint32)
    assert oracle.intersect(idx, other, 0).tolist() == [[0, 1], [1, 1], [1, 0]]


def test_build_mutual_kgraph_restatement_matches_list_semantics():
    """oracle.build_mutual_kgraph against a literal list-of-lists walk of embedder.rs:68-84 /
    nearest_neighbour_graph.rs:12-72 with the commented-out assignments (:33, :70) applied."""
    from rosella_b200 import synth
    tnf, cov, _ = synth.knn_tables(120, 3, seed=4)
    k = 8
    di, dd, _ = oracle.knn(cov, oracle.DEPTH, k)
    ti, _, _ = oracle.knn(tnf, oracle.TNF, k)
    depth = [[(int(j), float(w)) for j, w in zip(di[i], dd[i])] for i in range(120)]
    tnf_l = [[int(j) for j in ti[i]] for i in range(120)]
    for keep_n_edges in (0, 2, 6, 8):
        keep_first = max(keep_n_edges, 5)
        inter = [[e for p, e in enumerate(nb) if p < keep_first or node in tnf_l[e[0]]] for node, nb in enumerate(depth)]
        mutual = [[e for p, e in enumerate(nb) if p < keep_n_edges or any(r[0] == node for r in inter[e[0]])]
                  for node, nb in enumerate(inter)]
        off, nbr, w, disc = oracle.build_mutual_kgraph(di, dd, ti, keep_n_edges)
        assert [int(off[i + 1] - off[i]) for i in range(120)] == [len(m) for m in mutual]
        assert nbr.tolist() == [e[0] for m in mutual for e in m]
        assert np.array_equal(w, np.array([e[1] for m in mutual for e in m], np.float32))
        assert disc.tolist() == [i for i, m in enumerate(mutual) if not m]


@pytest.mark.parametrize("s", [1, 3, 10])
def test_metabat_rust_variant_main_branch_is_pinned(golden_dir, s):
    """MetabatDistance::distance (coverage_calculator.rs:588-666) against the reference's own numba
    metabat_distance_nn run on (mean - 9e-7, var - 9e-7) — where neither flight's clamp nor the equal-variance sign
    rule can fire the two are the same function (oracle/make_golden.py: make_rust_pin)."""
    g = np.load(os.path.join(golden_dir, "rust_pin.npz"))
    cov, ref = g[f"cov_s{s}"], g[f"metabat_rust_mainbranch_s{s}"]
    noclamp, noneq = g[f"noclamp_s{s}"], g[f"noneq_s{s}"]
    got = np.array([[oracle.metabat(a, b, oracle.RUST) for b in cov] for a in cov])
    main = noclamp & noneq
    assert main.sum() > 800
    assert np.allclose(got[main], ref[main], rtol=1e-9, atol=0)
    # pairs with an equal-variance sample: flight takes abs(), Rust keeps the sign -> the same value or NaN -> 1.0
    eq = noclamp & ~noneq
    same = np.isclose(got[eq], ref[eq], rtol=1e-9, atol=0)
    assert (same | (got[eq] == 1.0)).all() and same.any() and (~same).any()


def _raw_distance_model(q, r, scale):
    """Literal walk of raw_distance (src/sketch/sketch_distances.rs:61-126) on Python ints."""
    i = j = common = 0
    while i < len(q) and j < len(r):
        if q[i] < r[j]: i += 1
        elif q[i] > r[j]: j += 1
        else: common += 1; i += 1; j += 1
    if scale > 0:
        max_hash = (2 ** 64 - 1) // int(1.0 / scale)
        while i < len(q) and q[i] < max_hash: i += 1
        while j < len(r) and r[j] < max_hash: j += 1
    if i == 0 and j == 0:
        return 1.0
    return common / min(i, j) if min(i, j) else float("nan")


def test_min_jaccard_restatement_matches_the_literal_walk():
    from rosella_b200 import synth
    hashes, offs = synth.sketches(300, seed=11, mean_size=25, share_frac=0.3, empty_frac=0.03)
    lists = [[int(x) for x in hashes[int(offs[i]):int(offs[i + 1])]] for i in range(300)]
    rng = np.random.default_rng(0)
    seen_shared = seen_nan = seen_tail = 0
    for a, b in rng.integers(0, 300, (4000, 2)):
        want = _raw_distance_model(lists[a], lists[b], 0.1)
        got = oracle.min_jaccard(lists[a], lists[b], 0.1)
        assert (np.isnan(want) and np.isnan(got)) or got == want
        seen_shared += 0 < want < 1 if want == want else 0
        seen_nan += want != want
        seen_tail += any(x >= synth.MAX_HASH_SCALE_0_1 for x in lists[a] + lists[b])
    assert seen_shared > 10 and seen_nan > 5 and seen_tail > 50
    assert oracle.min_jaccard([], [], 0.1) == 1.0                       # :109-111
    # DepthDistance::eval (embedder.rs:133-171): metabat x (1 - min_jaccard) as f32, NaN -> 1
    cov = np.array([[5.0, 9.0, 2.0, 3.0], [7.0, 12.0, 0.0, 0.0]])
    m = oracle.metabat(cov[0], cov[1], oracle.RUST)
    assert oracle.depth_distance_sketch(cov[0], cov[1], [1, 2, 3, 4], [2, 4, 9], oracle.RUST) == np.float32(m * float(np.float32(1 - 2 / 3)))
    assert oracle.depth_distance_sketch(cov[0], cov[1], [1, 2], [7, 8], oracle.RUST) == np.float32(m)
    assert oracle.depth_distance_sketch(cov[0], cov[1], [], [7, 8], oracle.RUST) == np.float32(m)      # NaN -> factor 1
    assert oracle.depth_distance_sketch(cov[0], cov[1], [], [], oracle.RUST) == 0.0                   # both empty: min_jaccard = 1


@pytest.mark.parametrize("s", [1, 3, 10])
def test_rank_statistic_restatement_is_pinned(golden_dir, s):
    """oracle/rankstat.py against ProfileDistanceEngine.makeRanks / makeRankStat run from the reference itself
    (flight/flight/distance.py:47-108; oracle/make_golden.py: make_rankstat).  The rank step is pinned EXACTLY on the
    reference's own distance vectors; the distances agree to 1e-12.  Ranks computed from the oracle's distances can differ
    from the reference's where numba's fastmath code returns d(a, b) != d(b, a) in the last bit: with one sample, the
    pairs (present row, any absent row) tie exactly here and only almost there, so a tie group is averaged differently."""
    from oracle import rankstat
    g = np.load(os.path.join(golden_dir, "flight_rankstat.npz"))
    cov, tnf, lengths = g[f"cov_s{s}"], g[f"tnf_s{s}"], g[f"len_s{s}"]
    cr, kr, st = rankstat.stat_from_distances(g[f"cov_dist_s{s}"], g[f"kmer_dist_s{s}"], lengths, s)
    assert np.array_equal(cr, g[f"cov_ranks_s{s}"]) and np.array_equal(kr, g[f"kmer_ranks_s{s}"])
    assert np.allclose(st, g[f"stat_s{s}"], rtol=1e-14, atol=0)
    assert len(np.unique(cr)) < len(cr)                  # the fixture holds tied coverage distances
    dc, dk = rankstat.distances(cov, tnf)
    assert np.allclose(dc, g[f"cov_dist_s{s}"], rtol=1e-12, atol=0) and np.allclose(dk, g[f"kmer_dist_s{s}"], rtol=1e-13, atol=1e-15)
    cr2, kr2, st2 = rankstat.rank_stat(cov, tnf, lengths)
    groups = np.unique(g[f"cov_dist_s{s}"].round(12), return_counts=True)[1]
    assert np.abs(cr2 - cr).max() <= groups[groups < 300].max() and np.array_equal(kr2, kr)
    assert np.median(np.abs(st2 - st) / st) < 1e-12


def test_silhouette_restatement_matches_the_textbook_definition():
    """oracle.silhouette (cluster_utils.rs:38-120 with `a` over all other members) against scikit-learn on clean clusters,
    and the reference's corner cases: one cluster -> 0, a cluster of one -> a = 0 -> s = 1, outliers (label -1) unscored,
    the score divides by ALL rows (:119)."""
    from scipy.spatial.distance import pdist
    from sklearn.metrics import silhouette_samples
    rng = np.random.default_rng(0)
    x = np.concatenate([rng.normal(c, 0.3, (40, 2)) for c in ((0, 0), (3, 0), (0, 4))])
    lab = np.repeat([0, 1, 2], 40).astype(np.int32)
    d = oracle.condensed_euclidean(x)
    assert np.array_equal(d, pdist(x))
    score, pp = oracle.silhouette(d, lab, 3)
    ss = silhouette_samples(x, lab)
    assert np.allclose(pp, ss, rtol=1e-12, atol=1e-14) and abs(score - ss.mean()) < 1e-14
    assert oracle.silhouette(d, np.zeros(120, np.int32), 1)[0] == 0.0
    lab2 = lab.copy(); lab2[:10] = -1; lab2[10] = 3
    score2, pp2 = oracle.silhouette(d, lab2, 4)
    assert (pp2[:10] == 0).all() and pp2[10] == 1.0 and abs(score2 - pp2.sum() / 120) < 1e-15


# ---------------------------------------------------------------------------------------------------------------
# sketch_contigs (contig_sketcher.rs:47-92): finch's scaled MinHash restated in oracle/oracle_minhash.c

def test_murmur3_x64_128_pinned_to_applebys_source(tmp_path):
    """The hash function itself is pinned: known answer for b"hello", and every length 0..40 / several seeds against
    Austin Appleby's MurmurHash3.cpp (public domain), which scikit-learn vendors on this image."""
    import ctypes as C
    import subprocess
    assert oracle.murmur3_x64_128(b"hello", 0) == (0xcbd8a7b341bd9b02, 0x5b1e906a48ae1d19)
    assert oracle.murmur3_x64_128(b"", 0) == (0, 0)
    import sklearn
    src = os.path.join(os.path.dirname(sklearn.__file__), "utils", "src", "MurmurHash3.cpp")
    if not os.path.exists(src):
        pytest.skip("no MurmurHash3.cpp on this image")
    so = str(tmp_path / "libmm3.so")
    subprocess.check_call(["g++", "-O2", "-shared", "-fPIC", "-I", os.path.dirname(src), "-o", so, src])
    ref = C.CDLL(so)
    rng = np.random.default_rng(5)
    for length in list(range(0, 41)) + [63, 64, 65, 255]:
        key = bytes(rng.integers(0, 256, length, dtype=np.uint8))
        for seed in (0, 42, 0xdeadbeef):
            out = (C.c_uint64 * 2)()
            ref.MurmurHash3_x64_128(key, length, C.c_uint32(seed), out)
            assert oracle.murmur3_x64_128(key, seed) == (out[0], out[1]), (length, seed)


def _revcomp(b: bytes) -> bytes:
    return b.translate(bytes.maketrans(b"ACGTacgt", b"TGCAtgca"))[::-1]


def test_sketch_contigs_restatement_properties():
    rng = np.random.default_rng(9)
    alpha = np.frombuffer(b"ACGT", np.uint8)
    recs = [bytes(alpha[rng.integers(0, 4, n)]) for n in (5000, 1500, 40, 31, 30, 0)]
    recs.append(b"ACGT" * 2000)                                   # 2 distinct canonical 31-mers: fewer than kmers_to_sketch
    recs.append(recs[0][:2000] + b"N" * 500 + recs[0][2000:4000])
    recs.append(_revcomp(recs[0]))
    offs = np.zeros(len(recs) + 1, np.uint64); offs[1:] = np.cumsum([len(r) for r in recs])
    seq = np.frombuffer(b"".join(recs), np.uint8)
    h, so, nk, nl = oracle.sketch_contigs(seq, offs)
    max_hash = oracle.sketch_max_hash(0.1)
    assert max_hash == (2 ** 64 - 1) // 10
    sk = [h[int(so[i]):int(so[i + 1])] for i in range(len(recs))]
    for i, r in enumerate(recs):
        assert nl[i] == len(r)
        assert (np.diff(sk[i].astype(object)) > 0).all() if len(sk[i]) > 1 else True      # strictly ascending
    assert nk[0] == 5000 - 30 and nk[3] == 1 and nk[4] == 0 and nk[5] == 0
    # by hand for record 0: all canonical 31-mer hashes, distinct, ascending; <= max_hash, padded to 500
    r = recs[0]
    hs = set()
    for p in range(len(r) - 30):
        km = r[p:p + 31]
        hs.add(oracle.murmur3_x64_128(min(km, _revcomp(km)), 42)[0])
    hs = sorted(hs)
    below = sum(1 for x in hs if x <= max_hash)
    keep = below if below >= 500 else min(500, len(hs))
    assert [int(x) for x in sk[0]] == hs[:keep]
    # the reverse complement has the same canonical k-mers
    assert np.array_equal(sk[0], sk[8])
    # a record with 2 distinct canonical k-mers (ACGT is its own reverse complement) keeps both: kmers_to_sketch = 800
    # cannot be reached
    assert len(sk[6]) == 2 and nk[6] == 8000 - 30
    # one k-mer: kept whatever its hash (1 <= kmers_to_sketch = 3)
    assert len(sk[3]) == 1 and len(sk[4]) == 0


def test_device_kmer_hash_model_for_every_k():
    """sketch.cu murmur_kmer in integer arithmetic (same constants and bit tricks: 2-bit canonical k-mer -> four ASCII
    words through the nibble-spread + PRMT look-up, bytes past k zeroed, the zero-padded tail words) against
    MurmurHash3_x64_128 of the k-mer's ASCII bytes, for every k in 1..32.  The kernel itself is compared with the oracle
    on the GPU (tests/test_gpu_sketch.py) for seven of these k."""
    M = (1 << 64) - 1

    def byte_perm(x, y, s):
        src = [(x >> (8 * i)) & 0xFF for i in range(4)] + [(y >> (8 * i)) & 0xFF for i in range(4)]
        return sum(src[(s >> (4 * i)) & 0x7] << (8 * i) for i in range(4))

    def four(y):
        t = (y | (y << 4)) & 0x0F0F
        t = (t | (t << 2)) & 0x3333
        return byte_perm(byte_perm(0x54474341, 0, t), 0, 0x0123)

    def ascii8(x16):
        return four((x16 >> 8) & 0xFF) | (four(x16 & 0xFF) << 32)

    rotl = lambda x, r: ((x << r) | (x >> (64 - r))) & M

    def fmix(k):
        k ^= k >> 33; k = k * 0xff51afd7ed558ccd & M; k ^= k >> 33; k = k * 0xc4ceb9fe1a85ec53 & M; k ^= k >> 33
        return k

    def murmur_kmer(canon, k, seed):
        c1, c2 = 0x87c37b91114253d5, 0x4cf5ad432745937f
        x = (canon << (64 - 2 * k)) & M
        w = [ascii8((x >> 48) & 0xFFFF), ascii8((x >> 32) & 0xFFFF), ascii8((x >> 16) & 0xFFFF), ascii8(x & 0xFFFF)]
        for i in range(4):
            have = k - 8 * i
            w[i] = w[i] if have >= 8 else (0 if have <= 0 else w[i] & ((1 << (8 * have)) - 1))
        h = [seed, seed]

        def block(k1, k2):
            k1 = k1 * c1 & M; k1 = rotl(k1, 31); k1 = k1 * c2 & M; h[0] ^= k1
            h[0] = rotl(h[0], 27); h[0] = (h[0] + h[1]) & M; h[0] = (h[0] * 5 + 0x52dce729) & M
            k2 = k2 * c2 & M; k2 = rotl(k2, 33); k2 = k2 * c1 & M; h[1] ^= k2
            h[1] = rotl(h[1], 31); h[1] = (h[1] + h[0]) & M; h[1] = (h[1] * 5 + 0x38495ab5) & M

        if k >= 32:
            block(w[0], w[1]); block(w[2], w[3]); t1 = t2 = 0
        elif k >= 16:
            block(w[0], w[1]); t1, t2 = w[2], w[3]
        else:
            t1, t2 = w[0], w[1]
        t2 = t2 * c2 & M; t2 = rotl(t2, 33); t2 = t2 * c1 & M; h[1] ^= t2
        t1 = t1 * c1 & M; t1 = rotl(t1, 31); t1 = t1 * c2 & M; h[0] ^= t1
        h1, h2 = h[0] ^ k, h[1] ^ k
        h1 = (h1 + h2) & M; h2 = (h2 + h1) & M
        return (fmix(h1) + fmix(h2)) & M

    rng = np.random.default_rng(3)
    for k in range(1, 33):
        for seed in (42, 0, 0xdeadbeefcafe):
            codes = rng.integers(0, 4, k)
            canon = 0
            for c in codes:
                canon = (canon << 2) | int(c)
            kmer = bytes(b"ACGT"[int(c)] for c in codes)
            assert murmur_kmer(canon, k, seed) == oracle.murmur3_x64_128(kmer, seed)[0], (k, seed)
