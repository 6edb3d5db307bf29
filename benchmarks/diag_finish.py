#!/usr/bin/env python
"""Where the `finish` phase (merge + f64 re-evaluation + certificate + exhaustive pass) goes at the configs[4] shape, on
one GPU: a query shard of the size one of eight ranks owns, with and without the certificate."""
import json
import sys
import os
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from rosella_b200 import gpu, synth  # noqa: E402

n, s, k, rows = int(sys.argv[1]), int(sys.argv[2]), int(sys.argv[3]), int(sys.argv[4])
tnf, cov, _ = synth.knn_tables(n, s) if n <= 1_000_000 and s <= 50 else synth.knn_tables_parallel(n, s, threads=16)
ctx = gpu.Context(0)
out = {"n": n, "samples": s, "k": k, "query_rows": rows}
for certify, compact in ((1, 1), (1, 0), (0, 1)):
    ctx.set_option("certify", certify)
    ctx.set_option("exact_compact", compact)
    ctx.knn_load(None, cov)
    t = time.perf_counter()
    ctx.knn_search(gpu.GRAPH_DEPTH, k, 0, rows)
    wall = time.perf_counter() - t
    tm = ctx.timings()
    if certify and compact:
        import numpy as np
        fr, forced, dk, eps = ctx.knn_flagged_rows()
        out["flagged_rows"] = {"rows": fr.tolist()[:16], "dk": dk.tolist()[:16], "eps": eps.tolist()[:16], "forced": int(forced.sum())}
    out[f"certify{certify}_compact{compact}"] = {"wall_s": wall, "select_ms": tm["select_ms"], "rescore_ms": tm["rescore_ms"],
                                "rows_flagged": tm["rows_flagged"], "extra_pairs": tm["extra_pairs"], "load_transform_ms": tm["transform_ms"]}
print(json.dumps(out))
