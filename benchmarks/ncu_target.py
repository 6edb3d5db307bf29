#!/usr/bin/env python
"""Short, single-GPU command for ncu: one warm-up and one measured pass of every hot kernel.
  python benchmarks/ncu_target.py [--n 100000] [--samples 50] [--rows 9472] [--gbp 0.3]
Launch order per pass: depth_select, rescore(depth), tnf_select, rescore(tnf), tnf_count_packed, tnf_freq,
tnf_count_ascii, tnf_freq."""
import argparse
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

import numpy as np  # noqa: E402

from rosella_b200 import gpu, synth  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--n", type=int, default=100_000)
ap.add_argument("--samples", type=int, default=50)
ap.add_argument("--rows", type=int, default=9472)
ap.add_argument("--k", type=int, default=30)
ap.add_argument("--gbp", type=float, default=0.3)
ap.add_argument("--passes", type=int, default=2)
ap.add_argument("--sym-rows", type=int, default=0, help="then a whole-graph (symmetric) coverage build of this many rows")
a = ap.parse_args()

import torch  # noqa: E402

ctx = gpu.Context(0)
tnf, cov, _ = synth.knn_tables(a.n, a.samples)
ctx.knn_load(tnf, cov)
dev = torch.device("cuda", 0)
idx = torch.empty((a.rows, a.k), dtype=torch.int32, device=dev)
dst = torch.empty((a.rows, a.k), dtype=torch.float32, device=dev)
seq, offs = synth.contigs(int(a.gbp * 1e9), threads=8)
codes, valid, woff = gpu.pack_host(seq, offs)
for p in range(a.passes):
    ctx.knn_search_dev(gpu.GRAPH_DEPTH, a.k, 0, a.rows, idx.data_ptr(), dst.data_ptr())
    t1 = ctx.timings()
    ctx.knn_search_dev(gpu.GRAPH_TNF, a.k, 0, a.rows, idx.data_ptr(), dst.data_ptr())
    t2 = ctx.timings()
    ctx.tnf_load_packed(codes, valid, woff)
    ctx.tnf_count_resident()
    t3 = ctx.timings()
    ctx.tnf_load_ascii(seq, offs)
    ctx.tnf_count_resident()
    t4 = ctx.timings()
    print(f"pass {p}: depth_select {t1['select_ms']:.3f} ms rescore {t1['rescore_ms']:.3f} ms | tnf_select {t2['select_ms']:.3f} ms "
          f"rescore {t2['rescore_ms']:.3f} ms | tnf_count packed {t3['tnf_count_ms']:.3f} ms ascii {t4['tnf_count_ms']:.3f} ms "
          f"({int(offs[-1])} bp)")
if a.sym_rows:
    ctx.knn_load(None, cov[:a.sym_rows])
    ctx.knn_search(gpu.GRAPH_DEPTH, a.k)
    print(f"symmetric build of {a.sym_rows} rows: select {ctx.timings()['select_ms']:.3f} ms")
ctx.close()
