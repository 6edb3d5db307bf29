"""Whole-graph coverage build: symmetric block-pair path (knn_depth_sym.cu) against the batch kernel, same handle.
usage: python benchmarks/sym_build.py [n] [samples] [k] -> one JSON line"""
import json
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from rosella_b200 import gpu, synth  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 200_000
s = int(sys.argv[2]) if len(sys.argv) > 2 else 50
k = int(sys.argv[3]) if len(sys.argv) > 3 else 30
_, cov, _ = synth.knn_tables(n, s)
ctx = gpu.Context(0)
ctx.knn_load(None, cov, gpu.VARIANT_RUST)
out = {"n": n, "samples": s, "k": k}
res = {}
for mode in ("batch", "sym", "sym"):
    os.environ["RG_DEPTH_BUILD"] = mode
    t0 = time.perf_counter()
    idx, dist, nv = ctx.knn_search(gpu.GRAPH_DEPTH, k)
    dt = time.perf_counter() - t0
    tm = ctx.timings()
    out[f"{mode}_wall_s"] = round(dt, 4)
    out[f"{mode}_select_ms"] = round(tm["select_ms"], 2)
    res[mode] = (idx, dist)
out["identical"] = bool(np.array_equal(res["batch"][0], res["sym"][0]) and np.array_equal(res["batch"][1], res["sym"][1]))
out["speedup_select"] = round(out["batch_select_ms"] / out["sym_select_ms"], 3)
print(json.dumps(out))
