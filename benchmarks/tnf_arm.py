"""TNF counting leg of bench.py (BASELINE.json configs[3]): synthetic contig set, 1 GPU."""
from __future__ import annotations

import time

import numpy as np


def run(ctx, total_bp: int, steps: int, warmup: int, hbm_gbs: float, threads: int):
    import torch

    from rosella_b200 import gpu, synth

    t0 = time.time()
    seq, offs = synth.contigs(total_bp, threads=threads)
    gen_s = time.time() - t0
    n = len(offs) - 1
    bases = int(offs[-1])
    t0 = time.time()
    codes, valid, woff = gpu.pack_host(seq, offs, threads=threads)
    pack_s = time.time() - t0

    # ---- kernel-only, packed input resident in HBM (input >> 126 MB L2, so no flush is needed)
    ctx.tnf_load_packed(codes, valid, woff)
    ms = []
    for i in range(warmup + steps):
        ctx.tnf_count_resident()
        t = ctx.timings()["tnf_count_ms"]           # CUDA events around the counting kernel, on its stream
        if i >= warmup:
            ms.append(t)
    counts, nk, _ = ctx.tnf_fetch()
    kern_ms = float(np.mean(ms))
    algo_bytes = len(codes) * 12 + n * (136 * 4 + 8 + 8)      # 0.375 B/base + per-contig output / offsets
    achieved = algo_bytes / (kern_ms * 1e-3) / 1e9

    # ---- same kernel family straight from ASCII resident in HBM
    ctx.tnf_load_ascii(seq, offs)
    ms_a = []
    for i in range(warmup + steps):
        ctx.tnf_count_resident()
        t = ctx.timings()["tnf_count_ms"]
        if i >= warmup:
            ms_a.append(t)
    c2, nk2, _ = ctx.tnf_fetch()
    assert np.array_equal(counts, c2) and np.array_equal(nk, nk2), "packed and ASCII kernels disagree"
    assert int(nk.sum()) > 0.98 * (bases - 3 * n)

    # ---- end to end through the C ABI: host ASCII (pinned) -> counts / n_kmers / freqs in (pinned) host arrays.
    # rg_tnf_count_ascii pipelines slices on two streams: kernels and device->host copies run under the upload.
    pin = torch.empty(len(seq), dtype=torch.uint8, pin_memory=True)
    pin.numpy()[:] = seq
    seq_pin = pin.numpy()
    outs = (torch.empty((n, 136), dtype=torch.int32, pin_memory=True).numpy().view(np.uint32),
            torch.empty(n, dtype=torch.int64, pin_memory=True).numpy().view(np.uint64),
            torch.empty((n, 136), dtype=torch.float64, pin_memory=True).numpy())
    e2e = []
    b0 = ctx.timings()
    for i in range(2 + max(steps, 2)):
        t = time.perf_counter()
        ctx.tnf_count_ascii(seq_pin, offs, out=outs)
        if i >= 2:
            e2e.append(time.perf_counter() - t)
    b1 = ctx.timings()
    assert np.array_equal(outs[0], counts) and np.array_equal(outs[1], nk), "pipelined call disagrees with the resident kernels"
    runs = 2 + max(steps, 2)
    # PCIe roofline of the call: bytes that must cross the link / e2e time against a measured pinned host->device copy
    dev = torch.device("cuda", ctx.device)
    probe = torch.empty(1 << 30, dtype=torch.uint8, pin_memory=True)
    dst = torch.empty(1 << 30, dtype=torch.uint8, device=dev)
    best = 0.0
    for _ in range(4):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); dst.copy_(probe, non_blocking=True); e1.record(); torch.cuda.synchronize()
        best = max(best, (1 << 30) / (e0.elapsed_time(e1) * 1e-3) / 1e9)
    link_bytes = bases + n * (136 * 12 + 8 + 16)
    pcie = {"bound": "pcie", "achieved": link_bytes / float(np.mean(e2e)) / 1e9, "peak": best, "unit": "GB/s",
            "frac": link_bytes / float(np.mean(e2e)) / 1e9 / best, "bytes_over_the_link_per_call": link_bytes,
            "peak_source": "pinned 1 GiB host->device copy (one cudaMemcpyAsync), best of 4, measured in this run"}
    del probe, dst
    ascii_ms = float(np.mean(ms_a))
    ascii_bytes = bases + n * (136 * 4 + 8 + 8)
    # ---- MinHash sketching (sketch_contigs, contig_sketcher.rs:47-92) of a <= 2 Gbp prefix of the same pinned records:
    # rg_sketch_contigs + rg_sketch_fetch through the C ABI; device time and its phases from the library's CUDA events
    sketch = None
    try:
        ks = int(np.searchsorted(offs, np.uint64(min(bases, 2_000_000_000)), side="right")) - 1
        ks = max(ks, 1)
        sb = int(offs[ks])
        ctx.sketch_contigs(seq_pin[: int(offs[min(ks, 20)])], offs[: min(ks, 20) + 1])
        best_w, best_tm, n_h = None, None, 0
        for _ in range(2):
            t = time.perf_counter()
            h, so, _ = ctx.sketch_contigs(seq_pin[:sb], offs[: ks + 1])
            w = time.perf_counter() - t
            if best_w is None or w < best_w:
                best_w, best_tm, n_h = w, ctx.timings(), int(so[-1])
            del h
        sketch = {"bases": sb, "contigs": ks, "hashes": n_h, "kmer_length": 31, "scale": 0.1, "seed": 42,
                  "device_ms": best_tm["select_ms"], "upload_ms": best_tm["h2d_ms"], "hash_ms": best_tm["transform_ms"],
                  "sort_ms": best_tm["rescore_ms"], "gbp_per_s_device": sb / (best_tm["select_ms"] * 1e-3) / 1e9,
                  "e2e_s": best_w, "gbp_per_s_e2e": sb / best_w / 1e9,
                  "what": "host ASCII (pinned) -> sketches in pageable host memory, rg_sketch_contigs + rg_sketch_fetch"}
    except Exception as exc:  # the bench line must survive a failure of this side leg
        sketch = {"error": repr(exc)}
    return {
        "value": bases / (kern_ms * 1e-3) / 1e9, "unit": "Gbp/s", "bases": bases, "contigs": n,
        "kernel_ms": kern_ms, "ascii_kernel_ms": ascii_ms,
        "ascii_kernel_gbps": bases / (ascii_ms * 1e-3) / 1e9,
        "e2e": {"value": bases / float(np.mean(e2e)) / 1e9, "unit": "Gbp/s",
                "h2d_bytes_per_step": int((b1["h2d_bytes"] - b0["h2d_bytes"]) // runs),
                "d2h_bytes_per_step": int((b1["d2h_bytes"] - b0["d2h_bytes"]) // runs)},
        "roofline_pcie": pcie,
        "roofline_ascii": {"bound": "hbm", "achieved": ascii_bytes / (ascii_ms * 1e-3) / 1e9, "peak": hbm_gbs, "unit": "GB/s",
                           "frac": ascii_bytes / (ascii_ms * 1e-3) / 1e9 / hbm_gbs, "algorithmic_bytes": ascii_bytes,
                           "what": "tnf_count_ascii2_kernel, the kernel rg_tnf_count_ascii runs (1 B per base)"},
        "sketch": sketch,
        "roofline": {"bound": "hbm", "achieved": achieved, "peak": hbm_gbs, "unit": "GB/s",
                     "frac": achieved / hbm_gbs, "traffic": None, "algorithmic_bytes": algo_bytes},
        "host_generate_s": gen_s, "host_pack_s": pack_s,
        "sample": (seq, offs),
    }


def run_sharded(ctx, total_bp: int, steps: int, rank: int, world: int, threads: int):
    """TNF counting on `world` ranks: the records are split into contiguous ranges of equal bases, one
    per rank; every rank uploads its own range over its own PCIe link and counts it — no exchange, the outputs are
    disjoint row ranges.  Each rank synthesises its own range (same generator, rank-dependent seed).  Collective (the
    barrier).  -> dict on every rank; `value` is the aggregate end-to-end Gbp/s (all bases / slowest rank)."""
    import torch
    import torch.distributed as dist

    import oracle
    from rosella_b200 import synth

    seq, offs = synth.contigs(total_bp // world, seed=synth.SEED_TNF + 1000 * world + rank, threads=threads)
    bases = int(offs[-1])
    pin = torch.empty(len(seq), dtype=torch.uint8, pin_memory=True)
    pin.numpy()[:] = seq
    seq_pin = pin.numpy()
    nrec = len(offs) - 1
    outs = (torch.empty((nrec, 136), dtype=torch.int32, pin_memory=True).numpy().view(np.uint32),
            torch.empty(nrec, dtype=torch.int64, pin_memory=True).numpy().view(np.uint64),
            torch.empty((nrec, 136), dtype=torch.float64, pin_memory=True).numpy())
    counts, nk, _ = ctx.tnf_count_ascii(seq_pin, offs, out=outs)  # warm-up + parity on a prefix of this rank's range
    cut = min(len(offs) - 1, 400)
    wc, wnk, _ = oracle.tnf_count(seq, offs[:cut + 1], fast=True, threads=threads)
    ok = bool(np.array_equal(counts[:cut], wc) and np.array_equal(nk[:cut], wnk))
    times = []
    for _ in range(max(steps, 2)):
        torch.cuda.synchronize()
        dist.barrier()
        t = time.perf_counter()
        ctx.tnf_count_ascii(seq_pin, offs, out=outs)
        times.append(time.perf_counter() - t)
    # what the host can feed: every rank copies 1 GiB of pinned memory to its GPU at the same time
    probe = torch.empty(1 << 30, dtype=torch.uint8, pin_memory=True)
    dst = torch.empty(1 << 30, dtype=torch.uint8, device="cuda")
    probe_s = 1e9
    for _ in range(3):
        torch.cuda.synchronize()
        dist.barrier()
        t = time.perf_counter()
        dst.copy_(probe, non_blocking=True)
        torch.cuda.synchronize()
        probe_s = min(probe_s, time.perf_counter() - t)
    del probe, dst
    loc = torch.tensor([float(np.mean(times)), float(bases), probe_s, 0.0 if ok else 1.0], dtype=torch.float64, device="cuda")
    mx = loc.clone()
    dist.all_reduce(mx, op=dist.ReduceOp.MAX)
    sm = loc.clone()
    dist.all_reduce(sm, op=dist.ReduceOp.SUM)
    return {"value": float(sm[1]) / float(mx[0]) / 1e9, "unit": "Gbp/s", "bases": int(sm[1]), "ranks": world,
            "seconds_slowest_rank": float(mx[0]),
            "roofline_pcie": {"bound": "pcie (all ranks uploading at once)", "achieved": float(sm[1]) / float(mx[0]) / 1e9,
                              "peak": world * (1 << 30) / float(mx[2]) / 1e9, "unit": "GB/s",
                              "frac": (float(sm[1]) / float(mx[0])) / (world * (1 << 30) / float(mx[2])),
                              "peak_source": "concurrent pinned 1 GiB host->device copies on all ranks, slowest rank, best of 3"},
            "bit_exact_vs_oracle_prefix": bool(float(mx[3]) == 0.0),
            "what": "end to end through rg_tnf_count_ascii on every rank: pinned host ASCII of the rank's own contiguous "
                    "range -> counts / n_kmers / freqs in pinned host arrays; no exchange"}
