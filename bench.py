#!/usr/bin/env python
"""bench.py — the driver's benchmark contract for the kNN-graph hot path (+ TNF counting as extra keys).

Workload (BASELINE.json configs[2]): synthetic 1M contigs x (136 TNF + 50 coverage samples), k=30.
A "step" = one pass of the hot path over one batch of `--query-rows` query contigs per GPU against ALL
n reference contigs: both kNN graphs (depth + TNF), i.e. FP32 select kernels + f64 re-evaluation, and
for N>1 the NCCL all-gather of the per-rank slabs.  Cost is linear in query rows, so
contigs/s = rows per step / step time is the throughput of the full 1M-row build.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--n", type=int, default=1_000_000, help="contigs (reference rows)")
    ap.add_argument("--samples", type=int, default=50)
    ap.add_argument("--k", type=int, default=30)
    ap.add_argument("--query-rows", type=int, default=0, help="query rows per GPU per step (0 = 2 waves of CTAs)")
    ap.add_argument("--tnf-gbp", type=float, default=-1.0, help="TNF counting set in Gbp (-1 = auto, 0 = skip)")
    ap.add_argument("--whole-graph-rows", type=int, default=200_000,
                    help="rows of the whole-graph coverage build timed beside the step (N=1 only; 0 = skip)")
    ap.add_argument("--no-whole-graph-full", action="store_true",
                    help="skip the one whole build of both graphs at the full shape (SURVEY 8(d) metric; ~55 s on one GPU)")
    ap.add_argument("--cpu-seconds", type=float, default=15.0, help="budget of the cpu_baseline sample")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    return ap.parse_args()


# ---------------------------------------------------------------------------------- clocks
class ClockSampler:
    """Samples nvidia-smi clocks / throttle reasons of one GPU while the timed region runs."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int):
        self.index, self.rows, self.proc = index, [], None

    def __enter__(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "200"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except OSError:
            self.proc = None
        return self

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def __exit__(self, *a):
        if self.proc:
            self.proc.terminate()
            try:
                self.proc.wait(timeout=5)
            except subprocess.TimeoutExpired:
                self.proc.kill()

    def summary(self):
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            try:
                sm.append(float(r[0])); mx.append(float(r[1]))
            except (ValueError, IndexError):
                continue
            for nme, v in zip(names, r[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(nme)
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        return {"sm_mhz": float(np.median(sm)), "sm_max_mhz": float(max(mx)), "reasons": sorted(reasons), "samples": len(sm)}


# ---------------------------------------------------------------------------------- workload
def algorithmic_flops(cov_q: np.ndarray, cov_all_absent: np.ndarray, n: int, graphs_tnf: int = 1) -> float:
    """SURVEY §8d convention: F = pairs * 2*136 * n_tnf_metrics + A * 140, A = contributing
    (ordered pair, sample) triples = sum_s (Q*n - zq_s * z_s), z = rows absent in sample s."""
    q = cov_q.shape[0]
    zq = (cov_q[:, 0::2] <= 0).sum(axis=0).astype(np.float64)
    a = float((q * float(n) - zq * cov_all_absent).sum())
    return float(q) * n * 272.0 * graphs_tnf + a * 140.0


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return {"hbm_gbs": d.get("hbm_gbs", 6650.0), "bf16_tflops": d.get("bf16_tflops", 1647.0),
                "source": "measured (MEASURED_PEAKS.json)"}
    return {"hbm_gbs": 6650.0, "bf16_tflops": 1647.0, "source": "fallback (B200_PROFILING.md)"}


def ncu_traffic(kernel: str, grid: str | None = None, min_bytes: float = 0.0):
    """DRAM bytes per launch (read + write) of `kernel` from the newest committed `ncu --set full` summary under
    profiles/ (benchmarks/ncu_summary.py); `grid` restricts to a launch of that grid size, `min_bytes` to a launch
    that moved at least that much (a capture of a smaller input cannot stand in).  -> (bytes, file) or (None, None)"""
    import glob
    import hashlib
    # the files whose code a capture of `kernel` measures: a summary taken before one of them changed is stale
    files = {"depth_select2_kernel": ("knn_depth2.cu", "knn_depth_common.cuh"), "tnf_select_tc_kernel": ("knn_tnf_tc.cu", "knn_tnf_common.cuh"),
             "tnf_count_packed2": ("tnf_count2.cu",)}.get(kernel, ())
    csrc = os.path.join(os.path.dirname(os.path.abspath(__file__)), "rosella_b200", "csrc")
    now = {f: hashlib.sha1(open(os.path.join(csrc, f), "rb").read()).hexdigest() for f in files if os.path.exists(os.path.join(csrc, f))}
    scale = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "Tbyte": 1e12}
    for path in sorted(glob.glob(os.path.join(os.path.dirname(os.path.abspath(__file__)), "profiles", "*_ncu_full_summary.json")), reverse=True):
        try:
            doc = json.load(open(path))
            kernels = doc.get("kernels", []) if isinstance(doc, dict) else doc
            then = doc.get("source_sha1") if isinstance(doc, dict) else None
        except (OSError, ValueError):
            continue
        if then is None or any(then.get(f) != h for f, h in now.items()):
            continue                                     # captured from other code than the one being timed
        for rec in kernels:
            if kernel not in rec.get("Kernel Name", "") or (grid and rec.get("Grid Size") != grid):
                continue
            total, ok = 0.0, 0
            for key, val in rec.items():
                if key.startswith(("dram__bytes_read.sum", "dram__bytes_write.sum")):
                    unit = key[key.index("[") + 1:-1] if "[" in key else "byte"
                    try:
                        v = float(val)
                    except ValueError:
                        continue
                    if v == v and unit in scale:
                        total += v * scale[unit]
                        ok += 1
            if ok == 2 and total >= min_bytes:
                return total, os.path.basename(path)
    return None, None


def fp32_peak_tflops():
    """Live FFMA micro-benchmark (rosella_b200/build/microbench); MEASURED_PEAKS.json has no FP32 entry."""
    exe = os.path.join(ROOT, "rosella_b200", "build", "microbench")
    try:
        out = subprocess.run([exe], capture_output=True, text=True, timeout=120).stdout
        d = json.loads(out)
        return d["ffma_tflops"], d
    except Exception as e:  # noqa: BLE001
        # fall back to the committed measurement of the same micro-benchmark on this pool's B200
        try:
            d = json.load(open(os.path.join(ROOT, "profiles", "r1_microbench_fp32.json")))
            return d["ffma_tflops"], {"error": str(e), "fallback": "profiles/r1_microbench_fp32.json"}
        except Exception:  # noqa: BLE001
            return None, {"error": str(e)}


def workload_config(args, q_rows, world):
    shapes = {(100_000, 10, 30): "configs[1]", (1_000_000, 50, 30): "configs[2]", (5_000_000, 100, 50): "configs[4]"}
    name = shapes.get((args.n, args.samples, args.k))
    name = f"BASELINE.json {name}" if name else "custom shape"
    ref_gb = args.n * (136 + 4 * args.samples) * 4 / 1e9
    return {"workload": f"synthetic {args.n} contigs x (136 TNF + {args.samples} coverage samples), k={args.k}: "
                        f"exact depth + TNF kNN graphs ({name})",
            "n_contigs": args.n, "tnf_bins": 136, "coverage_samples": args.samples, "k": args.k,
            "query_rows_per_gpu_per_step": q_rows, "sharding": f"query rows over {world} rank(s), references replicated",
            "variant": "rust (coverage_calculator.rs:582-666, kmer_counting.rs:366-414)",
            "l2": f"256 MB buffer zeroed between timed iterations; reference set ({ref_gb:.1f} GB FP32) exceeds L2"}


DEFAULT_QUERY_ROWS = 2 * (2 * 148 * 32)      # two waves of depth-select CTAs on the 148 SMs of a B200


def reference_code_c1():
    """Baseline C1: the reference's own flight distance code, timed single-threaded as the reference
    drives it.  /root/reference exists in the build container only, so the measurement is the committed
    profiles/r2_c1_flight_cpu.json (benchmarks/c1_flight_cpu.py), quoted with where it was taken."""
    p = os.path.join(ROOT, "profiles", "r2_c1_flight_cpu.json")
    try:
        d = json.load(open(p))
        return {"source": "profiles/r2_c1_flight_cpu.json", "where": d.get("where"), "host_cores": d.get("host_cores"),
                "threads_used": d.get("threads_used"), "n": d.get("n"), "us_per_pair": d.get("runs"),
                "extrapolation": d.get("extrapolation")}
    except (OSError, ValueError) as e:
        return {"unavailable": f"{type(e).__name__}: {e}"}


def main_reference(args):
    """--impl reference: the reference's own CPU path for this workload, timed on the host cores.
    The reference (Rust) cannot be built here or on the GPU box, so this is the oracle port."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    from benchmarks import cpu_arm
    from rosella_b200 import synth
    tnf, cov, _ = synth.knn_tables(args.n, args.samples)
    vals, samples = [], []
    budget = max(2.0, min(args.cpu_seconds, 120.0 / max(args.steps + args.warmup, 1)))
    for i in range(args.warmup + args.steps):
        r = cpu_arm.knn_sample(tnf, cov, args.k, budget)
        if i >= args.warmup:
            vals.append(r["value"]); samples.append(r)
    v = float(np.mean(vals))
    last = samples[-1]
    line = {"impl": "reference", "metric": "kNN-graph contigs/s", "value": v, "unit": "contigs/s",
            "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": 1e3 * last["seconds"], "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            # the SAME config as the GPU arm (the sample the CPU actually processed is in cpu_baseline.sample)
            "config": workload_config(args, min(args.query_rows or DEFAULT_QUERY_ROWS, args.n // max(args.gpus, 1)), max(args.gpus, 1)),
            "cpu_baseline": {"value": v, "unit": "contigs/s", "cores": last["cores"], "kind": "port",
                             "sample": last["sample"], "reference_code": reference_code_c1()},
            "e2e": {"value": v, "unit": "contigs/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line), flush=True)


def whole_graph(ctx, cov, rows, k, rank=0, world=1):
    """The drop-in call is a build of ALL rows: the coverage graph then takes the symmetric block path
    (knn_depth_sym.cu, every unordered pair once; on N GPUs block shares + all-to-all of partial lists inside the
    library, knn_dist.cu).  Timed here on the first `rows` rows of the step's table, beside the batch kernel /
    query-row shards on the same rows, with the lists compared bit for bit.  Collective: every rank calls it."""
    import torch
    from rosella_b200 import gpu
    s = cov.shape[1] // 2
    ctx.knn_load_dist(None, cov[:rows] if rank == 0 else None, rows, 136, s, gpu.VARIANT_RUST, 0)
    out = {"workload": f"coverage graph of all {rows} rows x {s} samples, k={k}, {world} GPU(s)"}
    res = {}
    idx = np.zeros((rows, k), np.uint32)
    dst = np.zeros((rows, k), np.float32)
    for mode in ("batch", "symmetric"):
        ctx.set_option("depth_build", 1 if mode == "batch" else 0)
        torch.cuda.synchronize()
        if world > 1:
            torch.distributed.barrier()
        t0 = time.perf_counter()
        ctx.knn_build_dist(gpu.GRAPH_DEPTH, k, gather=True, out=(idx, dst))
        dt = torch.tensor([time.perf_counter() - t0], dtype=torch.float64, device="cuda")
        if world > 1:
            torch.distributed.all_reduce(dt, op=torch.distributed.ReduceOp.MAX)
        out[f"{mode}_s"] = float(dt)
        res[mode] = (idx.copy(), dst.copy())
        if world == 1:
            out[f"{mode}_select_ms"] = ctx.timings()["select_ms"]
    ctx.set_option("depth_build", 0)
    out["identical"] = bool(np.array_equal(res["batch"][0], res["symmetric"][0]) and np.array_equal(res["batch"][1], res["symmetric"][1]))
    out["contigs_per_s_symmetric"] = rows / out["symmetric_s"]
    out["contigs_per_s_batch"] = rows / out["batch_s"]
    return out


def parity_sample(ctx, tnf, cov, k, lists, oracle_rows=64, bit_rows=4096, seed=3):
    """Parity attached to a timed whole-graph build (rank 0, after the clock stopped): `oracle_rows` random rows of both
    graphs against the exhaustive f64 oracle on the host cores (tie-aware recall must be 1.0; a position may differ only
    where two neighbours tie within 2 f32 ulps), and `bit_rows` contiguous rows bit-compared with rg_knn_search through
    the batch kernel (a different select path: no symmetry, no exchange)."""
    import oracle
    from benchmarks import cpu_arm
    from rosella_b200 import gpu
    n = tnf.shape[0]
    threads = cpu_arm.host_threads()         # not OMP_NUM_THREADS: torch.distributed.run sets it to 1 for every rank
    rng = np.random.default_rng(seed)
    rows = np.sort(rng.choice(n, min(oracle_rows, n), replace=False)).astype(np.uint64)
    out = {"oracle_rows": int(len(rows)), "bit_rows": 0, "graphs": {}}
    qb = int(rng.integers(0, max(1, n - bit_rows)))
    qe = min(n, qb + bit_rows)
    out["bit_rows"] = qe - qb
    t0 = time.perf_counter()
    ok = True
    for name, which, ow, data in (("depth", gpu.GRAPH_DEPTH, oracle.DEPTH, cov), ("tnf", gpu.GRAPH_TNF, oracle.TNF, tnf)):
        idx, dst = lists[name]
        oi, od, _ = oracle.knn(data, ow, k, oracle.RUST, qrows=rows, threads=threads)
        gi, gd = idx[rows], dst[rows]
        hits = total = 0
        max_rel = 0.0
        for r in range(len(rows)):
            kth = od[r, -1]
            got = set(gi[r].tolist())
            for p in range(k):
                total += 1
                hits += (int(oi[r, p]) in got) or (od[r, p] >= kth * (1 - 2.4e-7))
            m = gi[r] == oi[r]
            if m.any():
                max_rel = max(max_rel, float(np.max(np.abs(gd[r][m] - od[r][m]) / np.maximum(od[r][m], 1e-30))))
        ctx.set_option("depth_build", 1)
        bi, bd, _ = ctx.knn_search(which, k, qb, qe)
        ctx.set_option("depth_build", 0)
        bit = bool(np.array_equal(bi, idx[qb:qe]) and np.array_equal(bd, dst[qb:qe]))
        g = {"recall": hits / max(total, 1), "positions_equal": float((gi == oi).mean()), "max_rel_dist_err": max_rel,
             "bit_identical_to_batch_kernel": bit}
        ok &= g["recall"] == 1.0 and bit and max_rel <= 2.4e-7
        out["graphs"][name] = g
    out["ok"] = bool(ok)
    out["oracle_threads"] = threads
    out["seconds"] = time.perf_counter() - t0
    out["recall"] = min(g["recall"] for g in out["graphs"].values())
    return out


def whole_graph_full(ctx, tnf, cov, k, rank=0, world=1, parity=True):
    """SURVEY.md §8(d)'s metric taken literally: N / wall time of the whole build of both graphs at the step's shape,
    inputs on the host (rank 0), H2D + NVLink broadcast + transform + kernels + exchange + complete host lists on every
    rank inside the timed region.  One pass (the kernels are warm from the steps).  Collective: every rank calls it.
    Phases are CUDA-event times on the library's stream, max over ranks; `parity_sample` is checked afterwards."""
    import torch
    from rosella_b200 import gpu
    n, s = tnf.shape[0], cov.shape[1] // 2
    lists = {name: (np.zeros((n, k), np.uint32), np.zeros((n, k), np.float32)) for name in ("depth", "tnf")}
    for a, b in lists.values():
        a.fill(0); b.fill(0)                                 # touch the pages outside the timed region
    if world > 1:
        # NCCL sets its rings and peer-to-peer connections up on first use (~1 s for the first broadcast, ~1 s for the
        # first all-to-all on 8 GPUs): one small build before the clock starts, as any long-lived host process has
        wt, wc, _ = __import__("rosella_b200.synth", fromlist=["knn_tables"]).knn_tables(8192, 2, seed=2)
        ctx.knn_load_dist(wt if rank == 0 else None, wc if rank == 0 else None, 8192, 136, 2, gpu.VARIANT_RUST, 0)
        for which in (gpu.GRAPH_DEPTH, gpu.GRAPH_TNF):
            ctx.knn_build_dist(which, 8, gather=True)
    torch.cuda.synchronize()
    if world > 1:
        torch.distributed.barrier()
    t0 = time.perf_counter()
    ctx.knn_load_dist(tnf if rank == 0 else None, cov if rank == 0 else None, n, 136, s, gpu.VARIANT_RUST, 0)
    t_load = time.perf_counter() - t0
    ctx.knn_build_dist(gpu.GRAPH_DEPTH, k, gather=True, out=lists["depth"])
    ph_depth = ctx.dist_timings()
    t_depth = time.perf_counter() - t0 - t_load
    ctx.knn_build_dist(gpu.GRAPH_TNF, k, gather=True, out=lists["tnf"])
    ph_tnf = ctx.dist_timings()
    wall = time.perf_counter() - t0
    tm = ctx.timings()
    loc = [wall, t_load, t_depth, wall - t_load - t_depth, ph_depth["load_h2d_ms"], ph_depth["load_bcast_ms"],
           ph_depth["load_transform_ms"], ph_depth["partial_ms"], ph_depth["exchange_ms"], ph_depth["finish_ms"],
           ph_depth["gather_ms"], ph_tnf["finish_ms"], ph_tnf["gather_ms"]]
    t = torch.tensor(loc, dtype=torch.float64, device="cuda")
    if world > 1:
        torch.distributed.all_reduce(t, op=torch.distributed.ReduceOp.MAX)
    v = [float(x) for x in t]
    out = {"workload": f"both graphs of all {n} rows x {s} samples, k={k}, {world} GPU(s): rg_knn_load_dist + rg_knn_build_dist "
                       "(host tables on rank 0 in, complete host lists on every rank out)",
           "seconds": v[0], "contigs_per_s": n / v[0],
           "phases": {"unit": "ms, max over ranks; CUDA events on the library stream except *_wall_s",
                      "load_wall_s": v[1], "depth_wall_s": v[2], "tnf_wall_s": v[3],
                      "load_h2d": v[4], "load_broadcast": v[5], "load_transform_and_collision_scan": v[6],
                      "depth_partial": v[7], "depth_all_to_all": v[8], "depth_merge_rescore": v[9],
                      "depth_gather_d2h": v[10], "tnf_select_rescore": v[11], "tnf_gather_d2h": v[12]},
           "rows_flagged": tm["rows_flagged"], "rows_exact": tm["rows_exact"], "extra_pairs": tm["extra_pairs"]}
    if parity and rank == 0:
        out["parity_sample"] = parity_sample(ctx, tnf, cov, k, lists)
    if world > 1:
        torch.distributed.barrier()
    return out


def main_ours(args):
    from benchmarks import cpu_arm, knn_arm, tnf_arm
    helpers = sys.modules[__name__]
    r = knn_arm.run(args, helpers)
    if r["rank"] != 0:
        return
    world, q_rows = r["world"], r["q_rows"]
    step_s = r["total_ms"] / args.steps * 1e-3
    value = world * q_rows / step_s
    e2e_value = world * q_rows / (r["e2e_total_ms"] / args.steps * 1e-3)
    pk = peaks()
    fp32, micro = fp32_peak_tflops()
    ach = r["flops_depth"] / (r["depth_sel_ms"] * 1e-3) / 1e12
    grid = f"({(q_rows + 31) // 32}, 1, 1)"
    traffic, traffic_src = ncu_traffic("depth_select2_kernel", grid)
    tc_traffic, tc_traffic_src = ncu_traffic("tnf_select_tc_kernel", f"({(q_rows + 127) // 128}, 1, 1)",
                                             min_bytes=0.9 * args.n * 136 * 8)     # hi + lo tf32 planes read at least once
    line = {
        "metric": "kNN-graph contigs/s", "value": value, "unit": "contigs/s", "n_gpus": world, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": step_s * 1e3, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f32 select + f64 re-evaluation", "data": "synthetic",
        "config": workload_config(args, q_rows, world),
        "e2e": {"value": e2e_value, "unit": "contigs/s", "h2d_bytes_per_step": r["h2d_per_step"],
                "d2h_bytes_per_step": r["d2h_per_step"],
                "note": "every step re-loads both f64 tables from pinned host memory (rg_knn_load) and returns host lists"},
        "gpu_launches": r["launches"], "clocks": r["clocks"],
        "roofline": {"kernel": "depth_select2_kernel", "bound": "fp32", "achieved": ach, "peak": fp32,
                     "unit": "TFLOP/s", "frac": (ach / fp32) if fp32 else None, "traffic": traffic,
                     "traffic_unit": "bytes per launch (dram read + write)", "traffic_source": traffic_src,
                     "peak_source": "live FFMA micro-benchmark (rosella_b200/csrc/microbench.cu); MEASURED_PEAKS.json has no FP32 CUDA-core entry",
                     "algorithmic_flops_per_launch": r["flops_depth"], "launch_ms": r["depth_sel_ms"]},
        "kernels_ms": {"depth_select": r["depth_sel_ms"], "depth_rescore": r["depth_res_ms"],
                       "tnf_select": r["tnf_sel_ms"], "tnf_rescore": r["tnf_res_ms"],
                       "tnf_select_tflops": r["flops_tnf"] / (r["tnf_sel_ms"] * 1e-3) / 1e12,
                       "load_h2d": r["load_h2d_ms"], "load_transform": r["load_transform_ms"]},
        # the TNF select runs on the tensor cores as three TF32 passes (hi.hi + hi.lo + lo.hi); TF32 dense peak = half
        # of the measured bf16 figure.  `frac` counts ALGORITHMIC flops (2*136 per pair), as the depth roofline does;
        # the work the three passes execute is reported apart
        "roofline_tnf_select": {"kernel": "tnf_select_tc_kernel", "bound": "tensor",
                                "achieved": r["flops_tnf"] / (r["tnf_sel_ms"] * 1e-3) / 1e12,
                                "peak": pk["bf16_tflops"] / 2.0, "unit": "TFLOP/s",
                                "frac": r["flops_tnf"] / (r["tnf_sel_ms"] * 1e-3) / 1e12 / (pk["bf16_tflops"] / 2.0),
                                "executed_tflops": 3.0 * r["flops_tnf"] / (r["tnf_sel_ms"] * 1e-3) / 1e12,
                                "executed_frac": 3.0 * r["flops_tnf"] / (r["tnf_sel_ms"] * 1e-3) / 1e12 / (pk["bf16_tflops"] / 2.0),
                                "traffic": tc_traffic, "traffic_source": tc_traffic_src,
                                "peak_source": pk["source"] + ", bf16 / 2 for TF32",
                                "algorithmic_flops_per_launch": r["flops_tnf"], "launch_ms": r["tnf_sel_ms"],
                                "note": "tcgen05.mma kind::tf32, 3 passes for FP32 accuracy; algorithmic = 2*136 flop per pair"},
        "full_build_seconds_est": args.n / value,
        "microbench": micro,
    }
    tnf, cov = r["tables"]
    if r.get("whole_graph"):
        line["whole_graph"] = r["whole_graph"]
    if r.get("whole_graph_full"):
        line["whole_graph_full"] = r["whole_graph_full"]
    if r.get("tnf_sharded"):
        line["tnf"] = r["tnf_sharded"]
    if world == 1:
        if not args.no_cpu_baseline:
            line["cpu_baseline"] = {k: v for k, v in cpu_arm.knn_sample(tnf, cov, args.k, args.cpu_seconds).items()
                                    if k in ("value", "unit", "cores", "kind", "sample")}
            line["cpu_baseline"]["reference_code"] = reference_code_c1()
        gbp = args.tnf_gbp
        if gbp < 0:
            import psutil
            gbp = 10.0 if psutil.virtual_memory().available > 96e9 else 2.0
        if gbp > 0:
            del tnf, cov
            t = tnf_arm.run(r["ctx"], int(gbp * 1e9), args.steps, args.warmup, pk["hbm_gbs"], min(32, os.cpu_count() or 8))
            seq, offs = t.pop("sample")
            t["roofline"]["peak_source"] = pk["source"]
            t["roofline"]["traffic"], t["roofline"]["traffic_source"] = ncu_traffic(
                "tnf_count_packed2", min_bytes=0.9 * t["roofline"]["algorithmic_bytes"])
            if not args.no_cpu_baseline:
                t["cpu_baseline"] = cpu_arm.tnf_sample(seq, offs, args.cpu_seconds)
            line["tnf"] = t
    print(json.dumps(line), flush=True)


if __name__ == "__main__":
    a = parse_args()
    if a.impl == "reference":
        main_reference(a)
    else:
        main_ours(a)
