#!/usr/bin/env python
"""Short single-GPU target for ncu: the ASCII TNF kernel and the two sketch hash passes on the same resident records.
  ncu --set full ... --kernel-name 'regex:tnf_count_ascii2|sketch_hash' python benchmarks/ncu_k1_sketch.py [gbp]"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from rosella_b200 import gpu, synth  # noqa: E402

gbp = float(sys.argv[1]) if len(sys.argv) > 1 else 0.5
ctx = gpu.Context(0)
seq, offs = synth.contigs(int(gbp * 1e9), threads=8)
ctx.tnf_load_ascii(seq, offs)
for _ in range(2):
    ctx.tnf_count_resident()
ctx.sync()
h, so, nk = ctx.sketch_contigs(seq, offs)
print("bases", int(offs[-1]), "hashes", len(h))
