#!/usr/bin/env python
"""Whole-graph build through rg_multi_* — ONE process, ONE caller thread, N GPUs (the call a Rust pipeline thread makes,
recover_engine.rs:510-519) — at a BASELINE.json shape, with parity attached:
  * `bit_rows` rows of both graphs bit-compared with rg_knn_search through the batch kernel on device 0,
  * `sample_rows` random rows of both graphs saved to gpurun_out/<tag>_sample.npz for the exhaustive f64 oracle
    (benchmarks/verify_sample.py; run on the host cores afterwards so that N GPUs do not idle behind the CPU check),
  * TNF counting of a `--tnf-gbp` set split over the N GPUs (rg_multi_tnf_count_ascii), prefix checked against the oracle.
  python benchmarks/multi_build.py --gpus 8 --n 5000000 --samples 100 --k 50 --tag r2_cfg4_n8"""
import argparse
import json
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

import numpy as np  # noqa: E402

from rosella_b200 import gpu, synth  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--gpus", type=int, default=1)
ap.add_argument("--n", type=int, default=1_000_000)
ap.add_argument("--samples", type=int, default=50)
ap.add_argument("--k", type=int, default=30)
ap.add_argument("--tag", default="multi")
ap.add_argument("--sample-rows", type=int, default=64)
ap.add_argument("--bit-rows", type=int, default=4096)
ap.add_argument("--tnf-gbp", type=float, default=0.0)
ap.add_argument("--threads", type=int, default=min(32, os.cpu_count() or 8))
a = ap.parse_args()
out_dir = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "gpurun_out")
os.makedirs(out_dir, exist_ok=True)
n, s, k = a.n, a.samples, a.k
shapes = {(100_000, 10, 30): "configs[1]", (1_000_000, 50, 30): "configs[2]", (5_000_000, 100, 50): "configs[4]"}
rep = {"workload": f"both kNN graphs of all {n} rows x (136 TNF + {s} samples), k={k}, {a.gpus} GPU(s), one process / one caller "
                   f"thread (rg_multi_knn_build); BASELINE.json {shapes.get((n, s, k), 'custom shape')}",
       "n": n, "samples": s, "k": k, "gpus": a.gpus}
t0 = time.time()
big = n >= 2_000_000
tnf, cov, _ = synth.knn_tables_parallel(n, s, threads=a.threads) if big else synth.knn_tables(n, s)
rep["generator"] = "synth.knn_tables_parallel" if big else "synth.knn_tables"
rep["host_generate_s"] = round(time.time() - t0, 1)

m = gpu.Multi(list(range(a.gpus)))
wt, wc, _ = synth.knn_tables(40_000, 4, seed=1)
m.knn_build(wt, wc, 10)                                    # NCCL connections, kernel images: outside the timed region
ti, td = np.zeros((n, k), np.uint32), np.zeros((n, k), np.float32)
di, dd = np.zeros((n, k), np.uint32), np.zeros((n, k), np.float32)
import ctypes as C  # noqa: E402
from rosella_b200 import _lib  # noqa: E402
t0 = time.perf_counter()
rc = m.lib.rg_multi_knn_build(m.m, _lib.ptr(tnf, C.c_double), _lib.ptr(cov, C.c_double), n, 136, s, k, gpu.VARIANT_RUST, 3,
                              _lib.ptr(ti, C.c_uint32), _lib.ptr(td, C.c_float), _lib.ptr(di, C.c_uint32), _lib.ptr(dd, C.c_float))
wall = time.perf_counter() - t0
m._chk(rc)
rep["seconds"] = wall
rep["contigs_per_s"] = n / wall
hs = [m.handle(i) for i in range(a.gpus)]
ph = {"depth": [h.dist_timings(gpu.GRAPH_DEPTH) for h in hs], "tnf": [h.dist_timings(gpu.GRAPH_TNF) for h in hs]}
rep["phases_ms_max_over_gpus"] = {g: {key: max(p[key] for p in ph[g]) for key in ph[g][0]} for g in ph}
tms = [h.timings() for h in hs]
rep["rows_flagged"] = int(sum(t["rows_flagged"] for t in tms))
rep["rows_exact"] = int(sum(t["rows_exact"] for t in tms))
rep["extra_pairs"] = int(tms[0]["extra_pairs"])
# algorithmic flops of the whole build (SURVEY 8(d) convention: ordered pairs, no credit for symmetry)
absent = (cov[:, 0::2] <= 0).sum(axis=0).astype(np.float64)
flops = float(n) * n * 272.0 + float((float(n) * n - absent * absent).sum()) * 140.0
rep["algorithmic_tflops_per_s_all_gpus"] = flops / wall / 1e12

# ---- parity: bit-compare with the batch kernel on device 0 (its tables are still loaded)
h0 = hs[0]
h0._n = n
rng = np.random.default_rng(5)
qb = int(rng.integers(0, max(1, n - a.bit_rows)))
qe = min(n, qb + a.bit_rows)
h0.set_option("depth_build", 1)
bit = {}
for name, which, (idx, dst) in (("depth", gpu.GRAPH_DEPTH, (di, dd)), ("tnf", gpu.GRAPH_TNF, (ti, td))):
    bi, bd, _ = h0.knn_search(which, k, qb, qe)
    bit[name] = bool(np.array_equal(bi, idx[qb:qe]) and np.array_equal(bd, dst[qb:qe]))
h0.set_option("depth_build", 0)
rep["parity_bit_rows"] = {"rows": [qb, qe], "bit_identical_to_batch_kernel": bit}
props = {}
for name, (idx, dst) in (("depth", (di, dd)), ("tnf", (ti, td))):
    srt = np.sort(idx[:: max(1, n // 200_000)], axis=1)
    props[name] = {"all_valid": bool((idx < n).all()), "self_excluded": bool((idx != np.arange(n, dtype=np.uint32)[:, None]).all()),
                   "ascending": bool((np.diff(dst, axis=1) >= 0).all()), "no_duplicates_sampled": bool((srt[:, 1:] != srt[:, :-1]).all())}
rep["list_properties"] = props
rows = np.sort(rng.choice(n, min(a.sample_rows, n), replace=False)).astype(np.uint64)
np.savez(os.path.join(out_dir, f"{a.tag}_sample.npz"), rows=rows, n=n, samples=s, k=k, parallel_generator=big,
         depth_idx=di[rows], depth_dist=dd[rows], tnf_idx=ti[rows], tnf_dist=td[rows])
rep["oracle_sample"] = {"rows": int(len(rows)), "file": f"gpurun_out/{a.tag}_sample.npz", "checked_by": "benchmarks/verify_sample.py"}
del tnf, cov, ti, td, di, dd

# ---- TNF counting over the N GPUs
if a.tnf_gbp > 0:
    import torch
    import oracle
    t0 = time.time()
    seq, offs = synth.contigs(int(a.tnf_gbp * 1e9), threads=a.threads)
    gen_s = time.time() - t0
    pin = torch.empty(len(seq), dtype=torch.uint8, pin_memory=True)
    pin.numpy()[:] = seq
    seq_pin = pin.numpy()
    nrec = len(offs) - 1
    outs = (torch.empty((nrec, 136), dtype=torch.int32, pin_memory=True).numpy().view(np.uint32),
            torch.empty(nrec, dtype=torch.int64, pin_memory=True).numpy().view(np.uint64),
            torch.empty((nrec, 136), dtype=torch.float64, pin_memory=True).numpy())
    counts, nk, freqs = m.tnf_count_ascii(seq_pin, offs, out=outs)  # warm-up + parity
    cut = min(nrec, 300)
    ok = True
    for lo, hi in ((0, cut), (nrec - cut, nrec)):
        wc2, wnk, _ = oracle.tnf_count(seq, (offs[lo:hi + 1]), fast=True, threads=a.threads)
        ok &= bool(np.array_equal(counts[lo:hi], wc2) and np.array_equal(nk[lo:hi], wnk))
    c1, nk1, _ = hs[0].tnf_count_ascii(seq_pin[: int(offs[min(nrec, 2000)])], offs[: min(nrec, 2000) + 1])
    ok &= bool(np.array_equal(c1, counts[: min(nrec, 2000)]))
    times = []
    for _ in range(3):
        t = time.perf_counter()
        m.tnf_count_ascii(seq_pin, offs, out=outs)
        times.append(time.perf_counter() - t)
    bases = int(offs[-1])
    rep["tnf"] = {"value": bases / float(np.mean(times)) / 1e9, "unit": "Gbp/s", "bases": bases, "contigs": nrec, "gpus": a.gpus,
                  "seconds": float(np.mean(times)), "bit_exact_vs_oracle_prefix_and_suffix": ok, "host_generate_s": round(gen_s, 1),
                  "what": "rg_multi_tnf_count_ascii end to end: pinned host ASCII -> counts / n_kmers / freqs on the host, contiguous "
                          "ranges of equal bases per GPU, no exchange; outputs in reused pinned arrays"}
m.close()
print(json.dumps(rep), flush=True)
