#!/usr/bin/env python
"""TNF counting leg alone (BASELINE.json configs[3]): python benchmarks/tnf_only.py [--gbp 10] [--steps 5]
The first-generation (shared-memory atomics) numbers are kept in profiles/r1d_tnf_4gbp_v1.json."""
import argparse
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

from benchmarks import tnf_arm  # noqa: E402
from rosella_b200 import gpu  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--gbp", type=float, default=10.0)
ap.add_argument("--steps", type=int, default=5)
ap.add_argument("--warmup", type=int, default=3)
a = ap.parse_args()
hbm = json.load(open(os.path.join(os.path.dirname(__file__), "..", "MEASURED_PEAKS.json")))["hbm_gbs"] \
    if os.path.exists(os.path.join(os.path.dirname(__file__), "..", "MEASURED_PEAKS.json")) else 6650.0
with gpu.Context(0) as ctx:
    t = tnf_arm.run(ctx, int(a.gbp * 1e9), a.steps, a.warmup, hbm, min(32, os.cpu_count() or 8))
    t.pop("sample")
    t["impl"] = "thread-private byte counters (tnf_count2.cu)"
    print(json.dumps(t))
