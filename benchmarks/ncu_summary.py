#!/usr/bin/env python
"""Summarise ncu reports (`ncu --set full ... -o X`) into one JSON for profiles/: per captured launch the metrics
the design discussion uses (durations, DRAM bytes, pipe utilisation, occupancy limits, stall samples).
  python benchmarks/ncu_summary.py OUT.json REPORT.ncu-rep [REPORT2.ncu-rep ...]"""
import csv
import io
import json
import subprocess
import sys

KEEP = ("Kernel Name", "Block Size", "Grid Size", "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "sm__throughput.avg.pct_of_peak_sustained_elapsed", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
        "launch__registers_per_thread", "launch__occupancy_limit_shared_mem", "launch__occupancy_limit_registers",
        "launch__shared_mem_per_block_dynamic", "sm__warps_active.avg.per_cycle_active",
        "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_tc.avg.pct_of_peak_sustained_active", "sm__pipe_tc_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_uniform.avg.pct_of_peak_sustained_active",
        "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active", "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "smsp__inst_executed.sum", "smsp__cycles_active.avg", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum",
        "lts__t_bytes.sum", "lts__t_sector_hit_rate.pct", "l1tex__t_bytes_pipe_lsu_mem_global_op_ld.sum",
        "l1tex__t_bytes_pipe_lsu_mem_global_op_st.sum")


def rows(rep):
    out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True, check=True).stdout
    start = out.index('"ID"')
    rd = list(csv.reader(io.StringIO(out[start:])))
    names, units, data = rd[0], rd[1], rd[2:]
    for r in data:
        d = {}
        for nme, unit, val in zip(names, units, r):
            if nme in KEEP or nme.startswith("smsp__pcsamp_warps_issue_stalled_") and not nme.endswith("_not_issued"):
                d[f"{nme} [{unit}]" if unit else nme] = val
        yield d


def source_hashes():
    """sha1 of every kernel source: bench.py refuses a roofline.traffic figure whose kernel file changed since the capture"""
    import glob
    import hashlib
    import os
    root = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "rosella_b200", "csrc")
    return {os.path.basename(f): hashlib.sha1(open(f, "rb").read()).hexdigest() for f in sorted(glob.glob(os.path.join(root, "*.cu*")))}


if __name__ == "__main__":
    res = {"reports": sys.argv[2:], "source_sha1": source_hashes(), "kernels": []}
    for rep in sys.argv[2:]:
        res["kernels"] += list(rows(rep))
    json.dump(res, open(sys.argv[1], "w"), indent=1)
    for k in res["kernels"]:
        t = [v for kk, v in k.items() if kk.startswith("gpu__time_duration")]
        print(k.get("Kernel Name", "?")[:90], t)
