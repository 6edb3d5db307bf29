"""CoverageTable I/O of the C++ host mirror (host/coverage_table.cpp) against an independent Python restatement of
src/coverage/coverage_table.rs:135-377 (csv module + the reference's loops), CPU only.  The reference's behaviour is
kept as it is: rows come out of from_file as [cov_1..cov_S, var_1..var_S] and a long-read table of more than
one sample does not reshape."""
import csv
import math

import numpy as np
import pytest

from rosella_b200 import host


def _model_from_file(path, long_mode):
    """coverage_table.rs:135-284, literally"""
    with open(path, newline="") as f:
        rd = list(csv.reader(f, delimiter="\t"))
    header, records = rd[0], [r for r in rd[1:] if r]
    samples = []
    for h in header[2::3] if long_mode else header[3::2]:
        name = h.split()[0] if long_mode else h
        if "/" in h:
            name = name.split("/")[-1].replace(".bam", "")
        samples.append(name)
    table, names, lengths, depths = [], [], [], []
    for rec in records:
        names.append(rec[0]); lengths.append(int(rec[1]))
        vals = [float(x) for x in rec[(2 if long_mode else 3):]]
        cov, var = vals[0::2], vals[1::2]
        depths.append(sum(cov) / len(cov) if long_mode else float(rec[2]))
        table.append(cov + var)
    flat = [x for r in table for x in r]
    if len(flat) != len(names) * 2 * len(samples):
        raise ValueError("ShapeError")
    return np.array(flat).reshape(len(names), 2 * len(samples)), names, lengths, depths, samples


def _rust_fixed3(v):
    if math.isnan(v):
        return "NaN"
    if math.isinf(v):
        return "-inf" if v < 0 else "inf"
    return f"{v:.3f}"


def _short_tsv(path, names, lengths, table, headers, crlf=False):
    nl = "\r\n" if crlf else "\n"
    with open(path, "w", newline="") as f:
        f.write("\t".join(["contigName", "contigLen", "totalAvgDepth"] + headers) + nl)
        for n, ln, row in zip(names, lengths, table):
            f.write("\t".join([n, str(ln), repr(float(np.mean(row[0::2])))] + [repr(float(x)) for x in row]) + nl)


def test_from_file_short_layout_and_sample_names(tmp_path):
    rng = np.random.default_rng(1)
    n, s = 7, 3
    table = np.round(rng.gamma(2.0, 5.0, (n, 2 * s)), 4)
    table[2, 1] = 0.0; table[3, 0] = 1e-7; table[4, 5] = 12345678.25
    names = [f"k141_{i} flag=1" for i in range(n)]
    lengths = [int(x) for x in rng.integers(1000, 90000, n)]
    headers = ["assembly.fna/sampleA.bam Mean", "assembly.fna/sampleA.bam Variance", "sampleB", "sampleB-var",
               "ref/dir/sample.C.bam.bam", "x"]
    for crlf in (False, True):
        p = str(tmp_path / f"cov{int(crlf)}.tsv")
        _short_tsv(p, names, lengths, table, headers, crlf)
        t = host.CoverageTable.from_file(p, host.SHORT_BAM)
        want, wn, wl, wd, ws = _model_from_file(p, False)
        assert np.array_equal(t.table, want) and t.contig_names == wn == names and t.contig_lengths == wl == lengths
        assert np.array_equal(t.average_depths, np.array(wd)) and t.sample_names == ws and t.output_path == p
        assert ws == ["sampleA Mean", "sampleB", "sample.C"]
        # the coverages of a row first, then its variances — not the interleaved layout of the file
        assert np.array_equal(t.table[:, :s], table[:, 0::2]) and np.array_equal(t.table[:, s:], table[:, 1::2])
        assert host.CoverageTable.from_file(p, host.SHORT_READ).sample_names == ws


def test_from_file_long_layout_and_its_limits(tmp_path):
    p = str(tmp_path / "long1.tsv")
    with open(p, "w") as f:
        f.write("Contig\tLength\treads/s1.bam Mean\treads/s1.bam Variance\n")
        f.write("c1\t5000\t10.5\t3.25\nc2\t700\t0\t0\n")
    t = host.CoverageTable.from_file(p, host.LONG_READ)
    want, wn, wl, wd, ws = _model_from_file(p, True)
    assert np.array_equal(t.table, want) and t.sample_names == ws == ["s1"] and t.contig_lengths == [5000, 700]
    assert list(t.average_depths) == wd == [10.5, 0.0]
    # two samples: the repeated length column is read as a coverage value (coverage_table.rs:245-251) and the rows do not
    # reshape — the reference fails here, and so does the mirror
    p2 = str(tmp_path / "long2.tsv")
    with open(p2, "w") as f:
        f.write("Contig\tLength\ts1 Mean\ts1 Variance\tLength\ts2 Mean\ts2 Variance\n")
        f.write("c1\t5000\t10.5\t3.25\t5000\t7.0\t1.0\n")
    with pytest.raises(ValueError):
        _model_from_file(p2, True)
    with pytest.raises(host.HostError, match="ShapeError"):
        host.CoverageTable.from_file(p2, host.LONG_BAM)


def test_from_file_errors(tmp_path):
    def write(name, text):
        p = str(tmp_path / name)
        open(p, "w").write(text)
        return p
    head = "contigName\tcontigLen\ttotalAvgDepth\ts\ts-var\n"
    with pytest.raises(host.HostError, match="fields"):
        host.CoverageTable.from_file(write("a.tsv", head + "c1\t10\t1.0\t1.0\n"))
    with pytest.raises(host.HostError, match="digit"):
        host.CoverageTable.from_file(write("b.tsv", head + "c1\t10.5\t1.0\t1.0\t2.0\n"))
    with pytest.raises(host.HostError, match="float"):
        host.CoverageTable.from_file(write("c.tsv", head + "c1\t10\t1.0\tabc\t2.0\n"))
    with pytest.raises(host.HostError, match="float"):
        host.CoverageTable.from_file(write("d.tsv", head + "c1\t10\t1.0\t 1.0\t2.0\n"))
    with pytest.raises(host.HostError, match="cannot open"):
        host.CoverageTable.from_file(str(tmp_path / "missing.tsv"))
    t = host.CoverageTable.from_file(write("e.tsv", head + "c1\t10\t1.0\tinf\tNaN\n\n"))      # Rust parses these; blank line skipped
    assert math.isinf(t.table[0, 0]) and math.isnan(t.table[0, 1]) and t.contig_names == ["c1"]
    assert host.CoverageTable.from_file(write("f.tsv", head)).table.shape == (0, 2)


def test_write_round_trip_merge_and_length_filter(tmp_path):
    rng = np.random.default_rng(2)
    n = 9
    names = [f"c{i}" for i in range(n)]
    lengths = [500, 1500, 1499, 20000, 3000, 100, 1500, 7, 99999]
    a = host.CoverageTable.new(rng.gamma(2.0, 3.0, (n, 4)), names, lengths, ["s1", "s2"])
    b = host.CoverageTable.new(rng.gamma(2.0, 3.0, (n, 2)), names, lengths, ["s\t3"])     # a tab in a sample name: quoted on write
    ta, tb = a.table, b.table
    a.merge(b)
    assert a.sample_names == ["s1", "s2", "s\t3"] and np.array_equal(a.table, np.hstack([ta, tb]))
    assert np.allclose(a.average_depths, a.table[:, 0::2].sum(axis=1) / 3, rtol=0, atol=0)   # coverage_table.rs:300-305
    other = host.CoverageTable.new(tb, names[::-1], lengths, ["s4"])
    with pytest.raises(host.HostError, match="different contig names"):
        a.merge(other)
    p = str(tmp_path / "out.tsv")
    a.write(p)
    assert a.output_path == p
    with open(p, newline="") as f:
        rows = list(csv.reader(f, delimiter="\t"))
    assert rows[0] == ["contigName", "contigLen", "totalAvgDepth", "s1", "s1-var", "s2", "s2-var", "s\t3", "s\t3-var"]
    for i, r in enumerate(rows[1:]):
        assert r[0] == names[i] and r[1] == str(lengths[i]) and r[2] == _rust_fixed3(a.average_depths[i])
        assert r[3:] == [_rust_fixed3(v) for v in a.table[i]]
    back = host.CoverageTable.from_file(p, host.SHORT_BAM)                              # values at 3 decimals, rows de-interleaved
    rounded = np.array([[float(_rust_fixed3(v)) for v in row] for row in a.table])
    assert np.array_equal(back.table, np.hstack([rounded[:, 0::2], rounded[:, 1::2]]))
    assert back.sample_names == ["s1", "s2", "s\t3"]
    removed = a.filter_by_length(1500)                                                  # :38-52 strictly shorter goes
    keep = [i for i, ln in enumerate(lengths) if ln >= 1500]
    assert removed == n - len(keep) and a.contig_names == [names[i] for i in keep] and a.contig_lengths == [lengths[i] for i in keep]
    assert np.array_equal(a.table, np.hstack([ta, tb])[keep])
    # special values print as Rust prints them
    c = host.CoverageTable.new(np.array([[float("nan"), float("inf")], [-0.0004, 2.0005]]), ["x", "y"], [1, 2], ["s"])
    c.write(p)
    rows = list(csv.reader(open(p, newline=""), delimiter="\t"))
    assert rows[1][3:] == ["NaN", "inf"] and rows[2][3:] == ["-0.000", _rust_fixed3(2.0005)]


def test_write_then_read_property(tmp_path):
    """write -> from_file over generated tables (hypothesis, derandomised): every value comes back as the number its
    {:.3} text denotes, names with quotes / tabs survive the csv quoting, rows come back coverages-then-variances."""
    from hypothesis import given, settings, strategies as st

    name = st.text(alphabet=st.sampled_from(list("abcXYZ019_.|=\" \t")), min_size=1, max_size=12).filter(lambda x: x.strip() == x and x != "")
    value = st.one_of(st.floats(min_value=-1e6, max_value=1e9, allow_nan=False), st.sampled_from([0.0, -0.0, 0.0005, 0.0015, 1e-9, 123456.7895]))
    counter = [0]

    @settings(derandomize=True, max_examples=40, deadline=None)
    @given(st.integers(1, 5), st.integers(1, 3), st.data())
    def run(n, s, data):
        names = data.draw(st.lists(name, min_size=n, max_size=n, unique=True))
        samples = data.draw(st.lists(name.filter(lambda x: "/" not in x), min_size=s, max_size=s))
        table = np.array(data.draw(st.lists(st.lists(value, min_size=2 * s, max_size=2 * s), min_size=n, max_size=n)))
        lengths = data.draw(st.lists(st.integers(0, 10 ** 9), min_size=n, max_size=n))
        t = host.CoverageTable.new(table, names, lengths, samples)
        counter[0] += 1
        p = str(tmp_path / f"t{counter[0]}.tsv")
        t.write(p)
        back = host.CoverageTable.from_file(p)
        rounded = np.array([[float(_rust_fixed3(v)) for v in row] for row in table])
        assert np.array_equal(back.table, np.hstack([rounded[:, 0::2], rounded[:, 1::2]]))
        assert back.contig_names == names and back.contig_lengths == lengths and back.sample_names == samples
        want, wn, wl, wd, ws = _model_from_file(p, False)
        assert np.array_equal(back.table, want) and wn == names and ws == samples

    run()
