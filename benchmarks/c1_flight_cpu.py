#!/usr/bin/env python
"""CPU baseline C1: the reference's OWN runnable distance code, timed as the reference drives it.

flight/flight/distance.py:127-133 computes the two condensed distance matrices with
    scipy.spatial.distance.pdist(cov, metrics.coverage_distance)   and   pdist(tnf, metrics.rho)
single-threaded (rosella.py:341-348 passes use_multiple_processes=False).  This script imports metrics.py from
/root/reference (read-only, numba) and times exactly those calls at N = 2000, then extrapolates ∝ N² to the BASELINE
shapes — labelled as an extrapolation.  /root/reference does not exist on the GPU box, so this runs in the build
container only and its output is committed under profiles/ (bench.py quotes it under cpu_baseline.reference_code).

  PYTHONDONTWRITEBYTECODE=1 python benchmarks/c1_flight_cpu.py [--n 2000] > profiles/r2_c1_flight_cpu.json"""
import argparse
import importlib.util
import json
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from rosella_b200 import synth  # noqa: E402

REF = "/root/reference/flight/flight/metrics.py"


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--n", type=int, default=2000)
    a = ap.parse_args()
    if not os.path.exists(REF):
        print(json.dumps({"unavailable": f"{REF} is absent (the reference exists in the build container only)"}))
        return
    try:
        sys.dont_write_bytecode = True
        spec = importlib.util.spec_from_file_location("flight_metrics", REF)
        m = importlib.util.module_from_spec(spec)
        spec.loader.exec_module(m)
        from scipy.spatial.distance import pdist
    except Exception as e:  # noqa: BLE001
        print(json.dumps({"unavailable": f"{type(e).__name__}: {e}"}))
        return
    out = {"code": "scipy.spatial.distance.pdist with flight.metrics.coverage_distance / metrics.rho "
                   "(flight/flight/distance.py:127-133), single thread as rosella.py:341-348 drives it",
           "where": "build container (no GPU box has /root/reference)", "host_cores": len(os.sched_getaffinity(0)),
           "threads_used": 1, "n": a.n, "runs": {}}
    pairs = a.n * (a.n - 1) // 2
    for s in (10, 50, 100):
        tnf, cov, _ = synth.knn_tables(a.n, s, flight=True)
        pdist(cov[:64], m.coverage_distance)                       # numba compile outside the timed region
        pdist(tnf[:64], m.rho)
        t = time.perf_counter(); pdist(cov, m.coverage_distance); t_cov = time.perf_counter() - t
        t = time.perf_counter(); pdist(tnf, m.rho); t_rho = time.perf_counter() - t
        out["runs"][f"S{s}"] = {"coverage_distance_us_per_pair": 1e6 * t_cov / pairs, "rho_us_per_pair": 1e6 * t_rho / pairs,
                                "seconds": t_cov + t_rho}
    ext = {}
    for name, n, s in (("configs[1] 100k x (136+10)", 100_000, 10), ("configs[2] 1M x (136+50)", 1_000_000, 50),
                       ("configs[4] 5M x (136+100)", 5_000_000, 100)):
        r = out["runs"][f"S{s}"]
        sec = (r["coverage_distance_us_per_pair"] + r["rho_us_per_pair"]) * 1e-6 * n * (n - 1) / 2
        ext[name] = {"seconds_extrapolated": sec, "contigs_per_s_extrapolated": n / sec,
                     "note": "extrapolated ∝ N² from N = %d, unordered pairs, single thread" % a.n}
    out["extrapolation"] = ext
    print(json.dumps(out, indent=1))


if __name__ == "__main__":
    main()
