#!/usr/bin/env python
"""Exhaustive f64 oracle check of the rows benchmarks/multi_build.py sampled from a timed whole-graph build.  Runs on
host cores only (no GPU): regenerates the same synthetic tables from their seeds, evaluates the sampled query rows
against ALL n rows with the oracle, and compares the saved lists (tie-aware recall, positions, distances).
  python benchmarks/verify_sample.py gpurun_out/r2_cfg4_n8_sample.npz > profiles/r2_cfg4_n8_parity.json"""
import json
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

import numpy as np  # noqa: E402

import oracle  # noqa: E402
from rosella_b200 import synth  # noqa: E402

g = np.load(sys.argv[1])
n, s, k = int(g["n"]), int(g["samples"]), int(g["k"])
rows = g["rows"].astype(np.uint64)
if len(sys.argv) > 2:
    rows = rows[: int(sys.argv[2])]
t0 = time.time()
threads = len(os.sched_getaffinity(0))
tnf, cov, _ = synth.knn_tables_parallel(n, s, threads=threads) if bool(g["parallel_generator"]) else synth.knn_tables(n, s)
rep = {"file": os.path.basename(sys.argv[1]), "n": n, "samples": s, "k": k, "oracle_rows": int(len(rows)), "host_threads": threads,
       "generate_s": round(time.time() - t0, 1), "graphs": {}}
TAU = 2.4e-7
for name, which, data in (("depth", oracle.DEPTH, cov), ("tnf", oracle.TNF, tnf)):
    t0 = time.time()
    oi, od, _ = oracle.knn(data, which, k, oracle.RUST, qrows=rows, threads=threads)
    gi, gd = g[f"{name}_idx"][: len(rows)], g[f"{name}_dist"][: len(rows)]
    hits = total = 0
    max_rel = 0.0
    for r in range(len(rows)):
        kth = od[r, -1]
        got = set(gi[r].tolist())
        for p in range(k):
            total += 1
            hits += (int(oi[r, p]) in got) or (od[r, p] >= kth * (1 - TAU))
        mm = gi[r] == oi[r]
        if mm.any():
            max_rel = max(max_rel, float(np.max(np.abs(gd[r][mm] - od[r][mm]) / np.maximum(od[r][mm], 1e-30))))
    rep["graphs"][name] = {"recall": hits / total, "positions_equal": float((gi == oi).mean()), "max_rel_dist_err": max_rel,
                           "oracle_seconds": round(time.time() - t0, 1)}
rep["recall"] = min(v["recall"] for v in rep["graphs"].values())
rep["ok"] = bool(rep["recall"] == 1.0 and all(v["max_rel_dist_err"] <= TAU for v in rep["graphs"].values()))
print(json.dumps(rep))
