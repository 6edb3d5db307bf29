"""CPU tests of the host side: the C-ABI library loads and exports what include/rosella_gpu.h
declares, the host packer and canonical index agree with the oracle, and GPU entry points fail
loudly (no CPU fallback) when no device is present."""
import ctypes as C
import os
import re

import numpy as np
import pytest

import oracle
from rosella_b200 import _lib, gpu, synth

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_library_exports_every_declared_symbol():
    hdr = open(os.path.join(ROOT, "include", "rosella_gpu.h")).read()
    declared = set(re.findall(r"\b(rg_[a-z0-9_]+)\s*\(", hdr))
    declared -= {"rg_handle"}
    assert declared == set(_lib.SYMBOLS), declared ^ set(_lib.SYMBOLS)
    lib = _lib.load()
    for s in declared:
        assert hasattr(lib, s), s
    assert lib.rg_abi_version() == 2


def test_canonical_kmers_match_oracle():
    assert gpu.canonical_kmers() == oracle.canonical_kmers()


def _unpack(codes, valid, woff, rec):
    out = []
    for w in range(int(woff[rec]), int(woff[rec + 1])):
        for j in range(32):
            if (int(valid[w]) >> j) & 1:
                out.append("ACGT"[(int(codes[w]) >> (2 * j)) & 3])
            else:
                out.append("x")
    return "".join(out)


def test_pack_host_follows_needletail_normalize(golden_dir):
    g = np.load(os.path.join(golden_dir, "tnf_fixtures.npz"))
    seq, offsets = g["seq"], g["offsets"]
    codes, valid, woff = gpu.pack_host(seq, offsets, threads=3)
    assert len(woff) == len(offsets) and woff[-1] == len(codes) == len(valid)
    for rec in list(range(12)) + list(range(len(offsets) - 10, len(offsets) - 1)):
        raw = bytes(seq[int(offsets[rec]):int(offsets[rec + 1])])
        norm = oracle.normalize(raw).decode()
        want = "".join(c if c in "ACGT" else "x" for c in norm)
        got = _unpack(codes, valid, woff, rec)
        assert got[:len(want)] == want
        assert set(got[len(want):]) <= {"x"} and len(got) - len(want) < 32


def test_no_cpu_fallback_without_device():
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    lib = _lib.load()
    h = C.c_void_p()
    rc = lib.rg_create(0, C.byref(h))
    assert rc == -1 and not h.value                       # RG_ERR_NO_DEVICE
    assert b"no CPU fallback" in lib.rg_last_error(None)
    with pytest.raises(gpu.RosellaGpuError):
        gpu.Context(0)


def test_synth_shapes_and_determinism():
    tnf, cov, genome = synth.knn_tables(500, 4)
    tnf2, cov2, _ = synth.knn_tables(500, 4)
    assert tnf.shape == (500, 136) and cov.shape == (500, 8)
    assert np.array_equal(tnf, tnf2) and np.array_equal(cov, cov2)
    assert np.isfinite(tnf).all() and (cov >= 0).all()
    assert np.allclose(np.exp(tnf).mean(axis=1) * 0 + tnf.shape[1], 136)
    seq, offs = synth.contigs(200_000, threads=2)
    seq2, offs2 = synth.contigs(200_000, threads=1)
    assert np.array_equal(seq, seq2) and np.array_equal(offs, offs2)
    assert offs[-1] == len(seq) >= 200_000
    wrapped, woffs = synth.fasta_wrap(seq, offs)
    c1, n1, _ = oracle.tnf_count(seq, offs, fast=True)
    c2, n2, _ = oracle.tnf_count(wrapped, woffs, fast=True)
    assert np.array_equal(c1, c2) and np.array_equal(n1, n2)


def test_bench_reference_arm_contract():
    """`bench.py --impl reference` (the oracle port on the host cores) prints one JSON line with the contract's keys;
    run here at a toy size, the driver runs it at BASELINE.json configs[2]."""
    import json
    import subprocess
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    out = subprocess.run([sys.executable, os.path.join(root, "bench.py"), "--impl", "reference", "--n", "2000", "--samples", "3",
                          "--k", "10", "--steps", "1", "--warmup", "1", "--cpu-seconds", "1"],
                         capture_output=True, text=True, timeout=300, check=True).stdout
    lines = [ln for ln in out.splitlines() if ln.startswith("{")]
    assert len(lines) == 1
    d = json.loads(lines[0])
    assert d["impl"] == "reference" and d["unit"] == "contigs/s" and d["value"] > 0 and d["higher_is_better"] is True
    assert d["cpu_baseline"]["kind"] == "port" and d["cpu_baseline"]["cores"] >= 1 and d["cpu_baseline"]["value"] == d["value"]
    assert d["e2e"] == {"value": d["value"], "unit": "contigs/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}
    assert d["config"]["n_contigs"] == 2000 and "workload" in d["config"]


def test_reference_arm_ignores_omp_num_threads_and_mirrors_the_gpu_config():
    """torch.distributed.run exports OMP_NUM_THREADS=1 to every rank; the reference arm must still use all host cores
    (round 1 measured one thread at N > 1), and its `config` must be the GPU arm's."""
    import json
    import subprocess
    import sys
    import bench
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    env = dict(os.environ, OMP_NUM_THREADS="1", RANK="0", WORLD_SIZE="4")
    out = subprocess.run([sys.executable, os.path.join(root, "bench.py"), "--impl", "reference", "--gpus", "4", "--n", "40000",
                          "--samples", "3", "--k", "10", "--steps", "1", "--warmup", "0", "--cpu-seconds", "1"],
                         capture_output=True, text=True, timeout=300, check=True, env=env).stdout
    d = json.loads([ln for ln in out.splitlines() if ln.startswith("{")][0])
    assert d["cpu_baseline"]["cores"] == len(os.sched_getaffinity(0))
    assert f"OpenMP {d['cpu_baseline']['cores']} threads" in d["cpu_baseline"]["sample"]

    class A:
        n, samples, k, gpus, query_rows = 40000, 3, 10, 4, 0
    assert d["config"] == bench.workload_config(A, min(bench.DEFAULT_QUERY_ROWS, A.n // 4), 4)
    assert "reference_code" in d["cpu_baseline"]


def test_register_classifier_model_of_the_ascii_tnf_kernel():
    """tnf_count2.cu classify4 in integer arithmetic: `byte & 7` keys two PRMT tables; (byte ^ expected) & 0xDF == 0 must
    hold exactly for the ten bytes needletail's normalize(false) turns into A/C/G/T, and the packing multiply must put the
    four 2-bit codes in order.  A model of the device code (same constants), checked for every byte value at every
    position; the kernel itself is compared with the oracle on the GPU (tests/test_gpu_tnf.py)."""
    import random

    def byte_perm(x, y, s):
        src = [(x >> (8 * i)) & 0xFF for i in range(4)] + [(y >> (8 * i)) & 0xFF for i in range(4)]
        return sum(src[(s >> (4 * i)) & 0x7] << (8 * i) for i in range(4))

    def classify4(w):
        v = w & 0x07070707
        t = (v * 0x1001) & 0xFFFFFFFF
        sel = byte_perm(t, 0, 0x0031)
        exp = byte_perm(0x43FF41FF, 0x47FF5554, sel)
        cod = byte_perm(0x01000000, 0x02000303, sel)
        bad = byte_perm(w, 0, 0x2301) ^ exp
        return ((cod * 0x04014010) & 0xFFFFFFFF) >> 24, bad

    code = {ord(c): v for c, v in zip("ACGTUacgtu", [0, 1, 2, 3, 3, 0, 1, 2, 3, 3])}
    rnd = random.Random(1)
    bases = list(code)
    for pos in range(4):
        for b in range(256):
            for _ in range(8):
                bs = [rnd.choice(bases) for _ in range(4)]
                bs[pos] = b
                w = sum(x << (8 * i) for i, x in enumerate(bs))
                packed, bad = classify4(w)
                ok = (bad & 0xDFDFDFDF) == 0
                assert ok == all(x in code for x in bs), (bs, hex(bad))
                if ok:
                    assert packed == sum(code[x] << (2 * i) for i, x in enumerate(bs))
    for _ in range(50000):
        w = rnd.getrandbits(32)
        assert ((classify4(w)[1] & 0xDFDFDFDF) == 0) == all(((w >> (8 * i)) & 0xFF) in code for i in range(4))


def test_integration_doc_binds_every_entry_point():
    """INTEGRATION.md promises an `extern "C"` block 1:1 with include/rosella_gpu.h: every declared entry point has its
    `pub fn` there, and nothing else does."""
    hdr = open(os.path.join(ROOT, "include", "rosella_gpu.h")).read()
    declared = set(re.findall(r"\b(rg_[a-z0-9_]+)\s*\(", hdr)) - {"rg_handle"}
    doc = open(os.path.join(ROOT, "INTEGRATION.md")).read()
    bound = set(re.findall(r"pub fn (rg_[a-z0-9_]+)", doc))
    assert bound == declared, bound ^ declared
    # the two instrumentation structs mirror the header field for field
    for c_name, rust_name in (("rg_timings", "RgTimings"), ("rg_dist_timings", "RgDistTimings")):
        body = re.search(r"typedef struct \{([^}]*)\} " + c_name + ";", hdr).group(1)
        body = re.sub(r"/\*.*?\*/", "", body, flags=re.S)
        c_fields = [f.strip() for decl in body.split(";") if decl.strip() for f in decl.strip().split(None, 1)[1].split(",")]
        rust = re.search(r"pub struct " + rust_name + r" \{(.*?)\n\}", doc, re.S).group(1)
        rust = re.sub(r"//.*", "", rust)
        assert re.findall(r"pub (\w+):", rust) == c_fields, (c_name, c_fields)
