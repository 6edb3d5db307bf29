"""The C++ host mirror of the reference's Rust interface (rosella_b200/host/rosella_host.hpp) — CPU tests of the
table type, CSV round trip, CLR, filters and the FASTA reader; GPU tests of count_kmers and
EmbedderEngine::build_mutual_kgraph against the oracle.  Reference lines are cited per test."""
import gzip
import math
import os
import random
import struct

import numpy as np
import pytest

import oracle
from rosella_b200 import host, synth


def _ryu_pretty(v: float) -> str:
    """Independent model of the text Rust's csv crate writes for an f64 (ryu pretty format)."""
    if math.isnan(v):
        return "NaN"
    if math.isinf(v):
        return "-inf" if v < 0 else "inf"
    sign = "-" if math.copysign(1.0, v) < 0 else ""
    if v == 0:
        return sign + "0.0"
    mant, exp = f"{abs(v):.17e}".split("e")           # exact enough to recover repr's digits below
    r = repr(abs(v))
    if "e" in r:
        m, e = r.split("e"); e = int(e)
    else:
        m, e = r, 0
    if "." in m:
        ip, fp = m.split(".")
    else:
        ip, fp = m, ""
    digits = (ip + fp).lstrip("0")
    k = e - len(fp)                                    # value = digits * 10^k
    stripped = digits.rstrip("0")
    k += len(digits) - len(stripped)
    digits = stripped or "0"
    n = len(digits)
    kk = n + k
    if 0 <= k and kk <= 16:
        return sign + digits + "0" * k + ".0"
    if 0 < kk <= 16:
        return sign + digits[:kk] + "." + digits[kk:]
    if -5 < kk <= 0:
        return sign + "0." + "0" * (-kk) + digits
    if n == 1:
        return sign + digits + "e" + str(kk - 1)
    return sign + digits[0] + "." + digits[1:] + "e" + str(kk - 1)


def test_format_f64_is_shortest_round_trip_in_ryu_layout():
    known = {1.0: "1.0", 0.1: "0.1", 0.3: "0.3", 1e-5: "0.00001", 1e-6: "1e-6", 1e16: "1e16", 1e15: "1000000000000000.0",
             1.5e300: "1.5e300", 123456789.125: "123456789.125", 1.2345678901234568e17: "1.2345678901234568e17",
             5e-324: "5e-324", -2.5: "-2.5", 0.0: "0.0", float("nan"): "NaN", float("inf"): "inf", float("-inf"): "-inf",
             0.007352941176470588: "0.007352941176470588"}
    for v, want in known.items():
        assert host.format_f64(v) == want, (v, host.format_f64(v), want)
    rng = random.Random(7)
    for _ in range(3000):
        v = struct.unpack("<d", struct.pack("<Q", rng.getrandbits(64)))[0]
        if math.isnan(v) or math.isinf(v):
            continue
        s = host.format_f64(v)
        assert float(s) == v and s == _ryu_pretty(v), (v, s, _ryu_pretty(v))
    for v in np.random.default_rng(1).dirichlet(np.ones(136)):      # the values the table really holds
        s = host.format_f64(float(v))
        assert float(s) == v and s == _ryu_pretty(float(v))


def test_table_write_read_round_trip_is_bit_exact(tmp_path):
    """KmerFrequencyTable::write / read (kmer_counting.rs:285-333): header-less CSV, f64 text round trip."""
    rng = np.random.default_rng(2)
    counts = rng.integers(0, 500, (40, 136)).astype(np.float64)
    counts[3] = 0.0                                                  # 0/0 row -> NaN (:104-107)
    with np.errstate(invalid="ignore"):
        freq = counts / counts.sum(axis=1, keepdims=True)
    names = [f"contig_{i} len={i * 7}" for i in range(39)] + ['odd,"name"']
    t = host.KmerFrequencyTable.new(freq, names)
    path = str(tmp_path / "kmer_frequencies.tsv")
    t.write(path)
    assert t.table_path == path
    first = open(path).readline().rstrip("\n").split(",")
    assert first[0] == names[0] and len(first) == 137                # no header, comma separated (:287-291)
    back = host.KmerFrequencyTable.read(path)
    assert back.contig_names == names and back.kmer_size == 4
    assert np.array_equal(back.kmer_table, freq, equal_nan=True)
    clr = host.KmerFrequencyTable.read(path, apply_clr=True).kmer_table        # the no_flight feature (:322-323)
    assert np.array_equal(clr, oracle.clr(freq), equal_nan=True)
    with pytest.raises(host.HostError):
        host.KmerFrequencyTable.read(str(tmp_path / "missing.tsv"))
    bad = tmp_path / "bad.tsv"
    bad.write_text("a,0.5,zzz\n")
    with pytest.raises(host.HostError, match="invalid float"):
        host.KmerFrequencyTable.read(str(bad))


def test_table_filters(tmp_path):
    """filter_by_index / filter_by_name (kmer_counting.rs:228-282)"""
    x = np.arange(12 * 136, dtype=np.float64).reshape(12, 136)
    names = [f"c{i}" for i in range(12)]
    t = host.KmerFrequencyTable.new(x, names)
    assert t.filter_by_index([1, 5, 11]) == 3
    keep = [i for i in range(12) if i not in (1, 5, 11)]
    assert t.contig_names == [names[i] for i in keep] and np.array_equal(t.kmer_table, x[keep])
    assert t.filter_by_name(["c0", "c7", "nope"]) == 2
    keep = [i for i in keep if i not in (0, 7)]
    assert t.contig_names == [names[i] for i in keep] and np.array_equal(t.kmer_table, x[keep])
    with pytest.raises(host.HostError, match="out of range"):
        t.filter_by_index([100])


def test_clr_and_k_clamp():
    x = np.random.default_rng(4).dirichlet(np.ones(136), 9)
    x = np.concatenate([x, x[:, :131] * 1.5], axis=1)[:, :131 + 136]    # a width that is not a multiple of 8
    # :345-355 with row.sum() in ndarray's order: 8 strided partial sums, (p0+p4)+(p1+p5)+(p2+p6)+(p3+p7), remainder
    def nd_sum(r):
        m = len(r) // 8 * 8
        p = [0.0] * 8
        for i in range(0, m, 8):
            for l in range(8):
                p[l] = p[l] + r[i + l]
        acc = 0.0
        for a, b in ((0, 4), (1, 5), (2, 6), (3, 7)):
            acc = acc + (p[a] + p[b])
        for v in r[m:]:
            acc = acc + v
        return acc
    want = np.array([[math.log(v / (nd_sum(r.tolist()) / len(r))) for v in r.tolist()] for r in x])
    assert np.array_equal(host.clr(x), want) and np.array_equal(oracle.clr(x), want)
    # recover_engine.rs:513-517
    assert host.clamp_n_neighbours(50, 30) == 25 and host.clamp_n_neighbours(299, 30) == 30
    assert host.clamp_n_neighbours(300, 30) == 30 and host.clamp_n_neighbours(10, 30) == 5


def test_contig_sketch_result_filter_by_index():
    """ContigSketchResult::filter_by_index (contig_sketcher.rs:158-176): names and sketches lose the same positions."""
    names = [f"c{i}" for i in range(6)]
    lists = [np.arange(i * 10, i * 10 + i, dtype=np.uint64) for i in range(6)]           # lengths 0..5
    offs = np.zeros(7, np.uint64); offs[1:] = np.cumsum([len(x) for x in lists])
    hashes = np.concatenate(lists)
    got_n, got_h, got_o = host.sketch_result_filter_by_index(names, hashes, offs, {0, 3, 5, 17})
    assert got_n == ["c1", "c2", "c4"]
    assert list(got_o) == [0, 1, 3, 7] and np.array_equal(got_h, np.concatenate([lists[1], lists[2], lists[4]]))
    got_n, got_h, got_o = host.sketch_result_filter_by_index(names, hashes, offs, set())
    assert got_n == names and np.array_equal(got_o, offs) and np.array_equal(got_h, hashes)


def test_recover_engine_tables_follow_filter_contigs():
    """RecoverEngine::retrieve_contig_data (recover_engine.rs:522-543) hands the included rows in row order; filter_contigs
    (:485-505) removes the same rows from the coverage table (coverage_table.rs:69-131), the TNF table and the sketches
    and collects the removed names."""
    n, s = 12, 2
    rng = np.random.default_rng(4)
    cov = rng.random((n, 2 * s)) + np.arange(n)[:, None]              # row i starts with i.xxx
    tnf = rng.random((n, 136)) + 100 * np.arange(n)[:, None]
    names = [f"contig_{i}" for i in range(n)]
    lists = [np.arange(i, dtype=np.uint64) + 1000 * i for i in range(n)]   # sketch of row i has i hashes
    offs = np.zeros(n + 1, np.uint64); offs[1:] = np.cumsum([len(x) for x in lists])
    eng = host.RecoverEngine(cov, tnf, names, n_neighbours=3, sketches=(np.concatenate(lists), offs))
    c0, t0, sl = eng.retrieve_contig_data({7, 2, 9, 2})
    assert list(c0) == [cov[2, 0], cov[7, 0], cov[9, 0]] and list(t0) == [tnf[2, 0], tnf[7, 0], tnf[9, 0]] and sl == [2, 7, 9]
    eng.filter_contigs([3, 0, 11])
    keep = [i for i in range(n) if i not in (0, 3, 11)]
    assert eng.n_contigs == 9 and eng.names() == [names[i] for i in keep] and eng.n_filtered == 3
    assert all(eng.is_filtered(names[i]) for i in (0, 3, 11)) and not eng.is_filtered(names[1])
    c, t, sl = eng.tables()
    assert np.array_equal(c, cov[keep]) and np.array_equal(t, tnf[keep]) and sl == keep
    eng.filter_contigs([8])                                           # positions are those of the filtered tables
    assert eng.names()[-1] == names[9] and eng.n_contigs == 8 and eng.is_filtered(names[10])
    plain = host.RecoverEngine(cov, tnf, names)                       # no sketches at all
    assert plain.retrieve_contig_data({1})[2] == [-1] and plain.tables()[2] is None


def _write_fasta(path, ids, seqs, width=60, gz=False):
    text = "".join(f">{i}\n" + "".join(s[p:p + width] + "\n" for p in range(0, len(s), width)) for i, s in zip(ids, seqs))
    if gz:
        with gzip.open(path, "wt") as f:
            f.write(text)
    else:
        open(path, "w").write(text)


def test_parse_fastx_file(tmp_path):
    rng = np.random.default_rng(5)
    ids = ["k141_1 flag=1 multi=3.0 len=300", "k141_2", "empty"]
    seqs = ["".join(rng.choice(list("ACGTNacgt"), 300)), "".join(rng.choice(list("ACGT"), 61)), ""]
    for gz in (False, True):
        p = str(tmp_path / ("a.fna.gz" if gz else "a.fna"))
        _write_fasta(p, ids, seqs, gz=gz)
        got_ids, seq, offs = host.parse_fastx_file(p)
        assert got_ids == ids and len(offs) == 4
        for i, s in enumerate(seqs):
            raw = bytes(seq[int(offs[i]):int(offs[i + 1])])
            assert oracle.normalize(raw).decode() == oracle.normalize(s.encode()).decode()
    fq = tmp_path / "r.fq"
    fq.write_text("@r1\nACGT\n+\nIIII\n@r2 x\nGGCC\n+\nIIII\n")
    got_ids, seq, offs = host.parse_fastx_file(str(fq))
    assert got_ids == ["r1", "r2 x"] and bytes(seq) == b"ACGTGGCC" and list(offs) == [0, 4, 8]
    with pytest.raises(host.HostError):
        host.parse_fastx_file(str(tmp_path / "missing.fna"))
    junk = tmp_path / "junk.txt"
    junk.write_text("hello\n")
    with pytest.raises(host.HostError, match="invalid start"):
        host.parse_fastx_file(str(junk))


@pytest.mark.gpu
def test_count_kmers_matches_the_oracle_and_reuses_the_table(tmp_path):
    """count_kmers (kmer_counting.rs:13-16, :41-123): FASTA -> GPU counts -> kmer_frequencies.tsv -> re-read."""
    seq, offs = synth.contigs(400_000, threads=2)
    ids = [f"contig_{i}" for i in range(len(offs) - 1)]
    seqs = [bytes(seq[int(offs[i]):int(offs[i + 1])]).decode() for i in range(len(ids))]
    fa = str(tmp_path / "assembly.fna")
    _write_fasta(fa, ids, seqs)
    out = str(tmp_path / "out")
    os.makedirs(out)
    t = host.count_kmers(fa, out)
    _, _, want = oracle.tnf_count(seq, offs, fast=True)
    assert t.contig_names == ids and np.array_equal(t.kmer_table, want, equal_nan=True)
    assert t.table_path == os.path.join(out, "kmer_frequencies.tsv") and os.path.exists(t.table_path)
    # second call reads the existing table (:46-50), with the no_flight CLR applied on read (:322-323)
    os.remove(fa)
    t2 = host.count_kmers(fa, out, no_flight=True)
    with np.errstate(divide="ignore", invalid="ignore"):
        assert np.array_equal(t2.kmer_table, oracle.clr(want), equal_nan=True)


@pytest.mark.gpu
def test_sketch_contigs_matches_the_oracle(tmp_path):
    """sketch_contigs (contig_sketcher.rs:11-92): FASTA (60-column lines) -> min_contig_size filter -> one sketch per
    record, against the oracle's restatement of finch on the records that pass the filter."""
    seq, offs = synth.contigs(300_000, threads=2, min_len=400, max_len=9000)
    ids = [f"k141_{i} len={int(offs[i + 1] - offs[i])}" for i in range(len(offs) - 1)]
    seqs = [bytes(seq[int(offs[i]):int(offs[i + 1])]).decode() for i in range(len(ids))]
    fa = str(tmp_path / "assembly.fna")
    _write_fasta(fa, ids, seqs)
    names, hashes, so, nk, sl = host.sketch_contigs(fa, min_contig_size=1500)
    got_ids, raw, roffs = host.parse_fastx_file(fa)                 # the raw records (with their line breaks)
    keep = [i for i in range(len(got_ids)) if len(seqs[i]) >= 1500]         # seq().len(): bases, not bytes with line breaks
    assert names == [got_ids[i] for i in keep] and 0 < len(keep) < len(ids)
    recs = [bytes(raw[int(roffs[i]):int(roffs[i + 1])]) for i in keep]
    koffs = np.zeros(len(recs) + 1, np.uint64); koffs[1:] = np.cumsum([len(r) for r in recs])
    wh, wo, wnk, wnl = oracle.sketch_contigs(np.frombuffer(b"".join(recs), np.uint8), koffs)
    assert np.array_equal(so, wo) and np.array_equal(hashes, wh) and np.array_equal(nk, wnk) and np.array_equal(sl, wnl)
    with pytest.raises(host.HostError):
        host.sketch_contigs(str(tmp_path / "missing.fna"))


@pytest.mark.gpu
def test_build_mutual_kgraph_matches_the_exhaustive_oracle():
    """EmbedderEngine::build_mutual_kgraph (embedder.rs:31-86): the returned graph is the depth kNN graph (the
    reference drops the intersect / mutual results, nearest_neighbour_graph.rs:33,:70); apply_filters applies them."""
    n, s, k = 900, 5, 20
    tnf, cov, _ = synth.knn_tables(n, s, seed=17)
    g = host.build_mutual_kgraph(tnf, cov, keep_n_edges=3, n_neighbours=k)
    assert g.nb_nodes == n and g.max_nbng == k and g.disconnected_nodes == [] and g.data_id_from_idx(7) == 7
    di, dd, _ = oracle.knn(cov, oracle.DEPTH, k)
    ti, td, _ = oracle.knn(tnf, oracle.TNF, k)
    agree = 0
    for node in range(n):
        nodes, w = g.neighbours(node)
        tn, tw = g.neighbours(node, tnf=True)
        assert len(nodes) == k and (np.diff(w) >= 0).all()
        agree += int(np.array_equal(nodes, di[node])) + int(np.array_equal(tn, ti[node]))
        m = nodes == di[node]
        assert np.allclose(w[m], dd[node][m].astype(np.float64), rtol=2.4e-7, atol=0)
    assert agree > 0.97 * 2 * n
    # filters applied: keep the first max(keep_n_edges, 5) edges of the depth list, then only edges whose reverse
    # is in the TNF graph (embedder.rs:68), then the same with mutual edges and keep_n_edges (:69)
    f = host.build_mutual_kgraph(tnf, cov, keep_n_edges=3, n_neighbours=k, apply_filters=True)
    lists = [g.neighbours(i)[0] for i in range(n)]
    tl = [set(g.neighbours(i, tnf=True)[0].tolist()) for i in range(n)]
    step1 = [[int(nb) for j, nb in enumerate(l) if j < 5 or i in tl[int(nb)]] for i, l in enumerate(lists)]
    s1 = [set(l) for l in step1]
    step2 = [[nb for j, nb in enumerate(l) if j < 3 or i in s1[nb]] for i, l in enumerate(step1)]
    for i in range(n):
        assert f.neighbours(i)[0].tolist() == step2[i]
    assert f.disconnected_nodes == [i for i in range(n) if not step2[i]]


@pytest.mark.gpu
def test_prefilter_nodes_rebuilds_until_no_node_is_disconnected():
    """RecoverEngine::prefilter_nodes (recover_engine.rs:464-482).  As the reference stands (filters dropped) an exact
    graph has no disconnected node and one build is enough.  With the filters applied, the loop removes the
    disconnected contigs from all tables, rebuilds, and from round `rounds` + 1 on keeps n_neighbours edges, which ends
    it.  Modelled here with the building block tested above (EmbedderEngine::build_mutual_kgraph on the filtered
    tables); the driver must land on the same graph, the same survivors and the same number of builds."""
    n, s, k = 700, 4, 12
    tnf, cov, _ = synth.knn_tables(n, s, seed=29)
    for j, row in enumerate((100, 350, 600)):                # three contigs far from everything: every distance is 1.0,
        cov[row, 0::2] = 1e4 * (j + 1)                       # so nobody lists them and none of their edges is mutual
        cov[row, 1::2] = 1.0
    names = [f"c{i}" for i in range(n)]
    eng = host.RecoverEngine(cov, tnf, names, n_neighbours=k)
    g = eng.prefilter_nodes(rounds=1)
    assert eng.prefilter_rounds == 1 and eng.n_contigs == n and eng.n_filtered == 0 and g.nb_nodes == n
    ref = host.build_mutual_kgraph(tnf, cov, keep_n_edges=0, n_neighbours=k)
    assert all(np.array_equal(g.neighbours(i)[0], ref.neighbours(i)[0]) for i in range(0, n, 7))
    for rounds in (1, 2):
        eng = host.RecoverEngine(cov, tnf, names, n_neighbours=k, apply_filters=True)
        g = eng.prefilter_nodes(rounds=rounds)
        rows, builds, current_round, keep_n = list(range(n)), 1, 1, 0
        m = host.build_mutual_kgraph(tnf[rows], cov[rows], keep_n_edges=0, n_neighbours=k, apply_filters=True)
        assert {100, 350, 600} <= set(m.disconnected_nodes)
        while m.disconnected_nodes:
            gone = set(m.disconnected_nodes)
            rows = [r for p, r in enumerate(rows) if p not in gone]
            current_round += 1
            if current_round > rounds:
                keep_n += k
            m = host.build_mutual_kgraph(tnf[rows], cov[rows], keep_n_edges=keep_n, n_neighbours=k, apply_filters=True)
            builds += 1
        assert eng.prefilter_rounds == builds and eng.n_contigs == len(rows) and eng.names() == [names[r] for r in rows]
        assert eng.n_filtered == n - len(rows) and all(eng.is_filtered(names[r]) for r in set(range(n)) - set(rows))
        assert g.nb_nodes == len(rows) and g.disconnected_nodes == []
        for i in range(len(rows)):
            a, wa = g.neighbours(i); b, wb = m.neighbours(i)
            assert np.array_equal(a, b) and np.array_equal(wa, wb)
    # evaluate_subset's graph (:410-435): node i = the i-th included row, disconnected nodes come back as original rows
    eng = host.RecoverEngine(cov, tnf, names, n_neighbours=k, apply_filters=True)
    subset = sorted(set(range(40, 640, 2)) | {100, 350, 600})
    sg = eng.evaluate_subset_graph(subset, keep_n_edges=0)
    sm = host.build_mutual_kgraph(tnf[subset], cov[subset], keep_n_edges=0, n_neighbours=k, apply_filters=True)
    assert sg.nb_nodes == len(subset) and sg.disconnected_nodes == [subset[p] for p in sm.disconnected_nodes]
    assert {100, 350, 600} <= set(sg.disconnected_nodes)
    assert all(np.array_equal(sg.neighbours(i)[0], sm.neighbours(i)[0]) for i in range(0, len(subset), 5))
    # the k clamp of :513-517: fewer than 10 k rows -> min(n / 2, k)
    small = host.RecoverEngine(cov[:50], tnf[:50], names[:50], n_neighbours=30)
    assert small.prefilter_nodes().max_nbng == 25


@pytest.mark.gpu
def test_recover_engine_from_files(tmp_path):
    """RecoverEngine::new (recover_engine.rs:93-131) without the coverm subprocess: coverage file -> filter_by_length ->
    TNF table counted on the GPU (CLR applied) and filtered by the removed names -> sketches on the GPU; the three tables
    must describe the same contigs in the same order, and the graph driver runs on them."""
    seq, offs = synth.contigs(500_000, threads=2, min_len=600, max_len=9000, seed=41)
    n = len(offs) - 1
    ids = [f"k141_{i}" for i in range(n)]
    seqs = [bytes(seq[int(offs[i]):int(offs[i + 1])]).decode() for i in range(n)]
    lengths = [len(x) for x in seqs]
    fa = str(tmp_path / "assembly.fna")
    _write_fasta(fa, ids, seqs)
    _, cov, _ = synth.knn_tables(n, 2, seed=8)
    cf = str(tmp_path / "coverage.tsv")
    with open(cf, "w") as f:
        f.write("contigName\tcontigLen\ttotalAvgDepth\tassembly.fna/s1.bam\ts1-var\tassembly.fna/s2.bam\ts2-var\n")
        for i in range(n):
            f.write("\t".join([ids[i], str(lengths[i]), repr(float(cov[i, 0::2].mean()))] + [repr(float(x)) for x in cov[i]]) + "\n")
    out = str(tmp_path / "out" / "nested")
    eng = host.RecoverEngine.from_files(fa, out, cf, min_contig_size=1500, n_neighbours=10)
    keep = [i for i in range(n) if lengths[i] >= 1500]
    assert 0 < len(keep) < n and eng.n_contigs == len(keep) and eng.names() == [ids[i] for i in keep]
    c, t, sl = eng.tables()
    assert np.array_equal(c, np.hstack([cov[keep][:, 0::2], cov[keep][:, 1::2]]))          # rows as from_file hands them (F6)
    _, _, freq = oracle.tnf_count(seq, offs, fast=True)
    with np.errstate(divide="ignore", invalid="ignore"):
        assert np.array_equal(t, oracle.clr(freq)[keep], equal_nan=True)
    koffs = np.zeros(len(keep) + 1, np.uint64); koffs[1:] = np.cumsum([lengths[i] for i in keep])
    kseq = np.frombuffer("".join(seqs[i] for i in keep).encode(), np.uint8)
    _, wo, _, _ = oracle.sketch_contigs(kseq, koffs)
    assert sl == list(np.diff(wo.astype(np.int64)))
    assert os.path.exists(os.path.join(out, "kmer_frequencies.tsv"))
    g = eng.prefilter_nodes(rounds=1)
    assert g.nb_nodes == len(keep) and g.max_nbng == 10 and eng.prefilter_rounds == 1
    # a coverage file that lists other contigs than the assembly: the reference's assert_eq! messages come back as errors
    with open(cf, "a") as f:
        f.write("\t".join(["extra", "5000", "1.0", "1.0", "1.0", "1.0", "1.0"]) + "\n")
    with pytest.raises(host.HostError, match="TNF table row count"):
        host.RecoverEngine.from_files(fa, str(tmp_path / "out2"), cf, min_contig_size=1500)


@pytest.mark.gpu
def test_build_mutual_kgraph_with_sketches_is_depth_distance():
    """ContigInformation carries the contig's MinHash sketch (embedder.rs:96); with sketches the depth graph is
    DepthDistance in full, metabat x (1 - min_jaccard) (embedder.rs:133-171)."""
    n, s, k = 800, 4, 15
    tnf, cov, _ = synth.knn_tables(n, s, seed=23)
    hashes, offs = synth.sketches(n, seed=3, mean_size=30, share_frac=0.3)
    g = host.build_mutual_kgraph(tnf, cov, keep_n_edges=3, n_neighbours=k, sketches=(hashes, offs), sketch_scale=0.1)
    plain = host.build_mutual_kgraph(tnf, cov, keep_n_edges=3, n_neighbours=k)
    oi, od = oracle.knn_depth_sketch(cov, hashes, offs, k, oracle.RUST, 0.1)
    same = changed = 0
    for node in range(n):
        nodes, w = g.neighbours(node)
        same += int(np.array_equal(nodes, oi[node]))
        m = nodes == oi[node]
        assert np.allclose(w[m], od[node][m].astype(np.float64), rtol=2.4e-7, atol=1e-15)
        changed += int(not np.array_equal(nodes, plain.neighbours(node)[0]))
    assert same > 0.98 * n and changed > 20
