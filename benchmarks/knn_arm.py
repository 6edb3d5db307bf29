"""The GPU arm of bench.py: kNN-graph steps through the C ABI (see bench.py for the contract)."""
from __future__ import annotations

import os
import time

import numpy as np


def run(args, helpers):
    import torch
    import torch.distributed as dist

    from rosella_b200 import gpu, synth

    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    torch.cuda.set_device(local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    n, s, k = args.n, args.samples, args.k
    ctx = gpu.Context(local_rank)
    if world > 1:
        from rosella_b200 import dist as rdist
        rdist.init_comm(ctx, rank, world)           # NCCL communicator inside librosella_gpu.so (rg_comm_init)
    sms = torch.cuda.get_device_properties(local_rank).multi_processor_count
    # depth select: 32 query rows per CTA, 2 CTAs per SM -> a multiple of 2*SMs*32 fills whole waves
    q_rows = args.query_rows or 2 * (2 * sms * 32)
    q_rows = min(q_rows, n // world)

    t0 = time.time()
    tnf, cov, _ = synth.knn_tables(n, s)
    gen_s = time.time() - t0
    # pinned host copies: the e2e leg copies from these every step
    tnf_pin = torch.empty((n, 136), dtype=torch.float64, pin_memory=True)
    cov_pin = torch.empty((n, 2 * s), dtype=torch.float64, pin_memory=True)
    tnf_pin.numpy()[:] = tnf
    cov_pin.numpy()[:] = cov
    tnf, cov = tnf_pin.numpy(), cov_pin.numpy()

    ctx.knn_load(tnf, cov, gpu.VARIANT_RUST)
    load_t = ctx.timings()
    ext = torch.cuda.ExternalStream(ctx.stream, device=local_rank)

    # all-gather buffers: every rank's slab is written in place by rg_knn_search_dev
    dev = torch.device("cuda", local_rank)
    g_idx = [torch.empty((world * q_rows, k), dtype=torch.int32, device=dev) for _ in range(2)]
    g_dst = [torch.empty((world * q_rows, k), dtype=torch.float32, device=dev) for _ in range(2)]
    nv = torch.empty(q_rows, dtype=torch.int32, device=dev)
    flush = torch.empty(256 * 1024 * 1024, dtype=torch.uint8, device=dev)   # > 126 MB L2

    def q_range(step):
        # a different batch of query rows every step; rank r takes the r-th slab of the batch
        nb = (n // (world * q_rows))
        b = (step % max(nb, 1)) * world * q_rows + rank * q_rows
        return b, b + q_rows

    def step_device(step):
        qb, qe = q_range(step)
        for gi, which in enumerate((gpu.GRAPH_DEPTH, gpu.GRAPH_TNF)):
            si = g_idx[gi][rank * q_rows:(rank + 1) * q_rows]
            sd = g_dst[gi][rank * q_rows:(rank + 1) * q_rows]
            ctx.knn_search_dev(which, k, qb, qe, si.data_ptr(), sd.data_ptr(), nv.data_ptr())
            if world > 1:
                with torch.cuda.stream(ext):
                    dist.all_gather_into_tensor(g_idx[gi], si)
                    dist.all_gather_into_tensor(g_dst[gi], sd)

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # ------------------------------------------------------------------ device-timed value
    for w in range(args.warmup):
        step_device(w)
    barrier()
    launches0 = ctx.timings()["kernel_launches"]
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(2 * args.steps)]
    sel_ms, res_ms = [], []
    with helpers.ClockSampler(local_rank) as clk:
        for i in range(args.steps):
            with torch.cuda.stream(ext):
                flush.zero_()                      # L2 flush between timed iterations (outside the events)
                ev[2 * i].record(ext)
            step_device(args.warmup + i)
            with torch.cuda.stream(ext):
                ev[2 * i + 1].record(ext)
        barrier()
    launches = ctx.timings()["kernel_launches"] - launches0
    step_ms = [ev[2 * i].elapsed_time(ev[2 * i + 1]) for i in range(args.steps)]
    total_ms = float(sum(step_ms))
    clocks = clk.summary()

    # dominant kernel (depth select) duration, measured live with the library's CUDA events
    qb, qe = q_range(0)
    ctx.knn_search_dev(gpu.GRAPH_DEPTH, k, qb, qe, g_idx[0].data_ptr(), g_dst[0].data_ptr(), nv.data_ptr())
    tm = ctx.timings()
    depth_sel_ms, depth_res_ms = tm["select_ms"], tm["rescore_ms"]
    ctx.knn_search_dev(gpu.GRAPH_TNF, k, qb, qe, g_idx[1].data_ptr(), g_dst[1].data_ptr(), nv.data_ptr())
    tm = ctx.timings()
    tnf_sel_ms, tnf_res_ms = tm["select_ms"], tm["rescore_ms"]

    # ------------------------------------------------------------------ end-to-end through host buffers
    idx_h = np.zeros((q_rows, k), np.uint32)
    e2e_ms = []
    b0 = ctx.timings()
    for i in range(args.warmup + args.steps):
        barrier()
        t = time.perf_counter()
        ctx.knn_load(tnf, cov, gpu.VARIANT_RUST)             # H2D of both tables from pinned memory + transform
        qb, qe = q_range(i)
        r1 = ctx.knn_search(gpu.GRAPH_DEPTH, k, qb, qe)      # host outputs (D2H inside)
        r2 = ctx.knn_search(gpu.GRAPH_TNF, k, qb, qe)
        idx_h[:] = r1[0]
        barrier()
        if i >= args.warmup:
            e2e_ms.append((time.perf_counter() - t) * 1e3)
    b1 = ctx.timings()
    n_e2e = args.warmup + args.steps
    h2d_per_step = (b1["h2d_bytes"] - b0["h2d_bytes"]) // n_e2e
    d2h_per_step = (b1["d2h_bytes"] - b0["d2h_bytes"]) // n_e2e

    # max over ranks
    t_loc = torch.tensor([total_ms, float(sum(e2e_ms))], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t_loc, op=dist.ReduceOp.MAX)
    total_ms, e2e_total_ms = float(t_loc[0]), float(t_loc[1])

    wgf = None
    if not args.no_whole_graph_full:
        wgf = helpers.whole_graph_full(ctx, tnf, cov, k, rank, world)
    wg = None
    if args.whole_graph_rows > 0:
        wg = helpers.whole_graph(ctx, cov, min(args.whole_graph_rows, n), k, rank, world)

    tnf_sharded = None
    if world > 1 and args.tnf_gbp != 0:
        from benchmarks import tnf_arm
        gbp = args.tnf_gbp if args.tnf_gbp > 0 else 8.0
        tnf_sharded = tnf_arm.run_sharded(ctx, int(gbp * 1e9), args.steps, rank, world, max(2, min(8, (os.cpu_count() or 8) // world)))

    absent_all = (cov[:, 0::2] <= 0).sum(axis=0).astype(np.float64)
    flops_depth = helpers.algorithmic_flops(cov[qb:qe], absent_all, n, graphs_tnf=0)
    flops_tnf = float(q_rows) * n * 272.0
    result = {
        "rank": rank, "world": world, "q_rows": q_rows, "n": n, "s": s, "k": k, "sms": sms,
        "total_ms": total_ms, "step_ms": step_ms, "e2e_total_ms": e2e_total_ms, "launches": int(launches),
        "clocks": clocks, "depth_sel_ms": depth_sel_ms, "depth_res_ms": depth_res_ms, "tnf_sel_ms": tnf_sel_ms,
        "tnf_res_ms": tnf_res_ms, "flops_depth": flops_depth, "flops_tnf": flops_tnf,
        "h2d_per_step": int(h2d_per_step), "d2h_per_step": int(d2h_per_step), "gen_s": gen_s,
        "load_h2d_ms": load_t["h2d_ms"], "load_transform_ms": load_t["transform_ms"],
        "tables": (tnf, cov), "ctx": ctx, "whole_graph": wg, "whole_graph_full": wgf, "tnf_sharded": tnf_sharded,
    }
    return result
