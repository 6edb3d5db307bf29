#!/usr/bin/env python
"""Whole-graph build of BASELINE.json configs[2] on N GPUs (torchrun, NCCL; N = 1 runs without a process group):
coverage graph by the symmetric block build (all-to-all of partial lists) and by query-row shards, TNF graph by
query-row shards.  Times are device-synchronised wall clock, max over ranks.  One JSON line on rank 0.
  python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29513 \
      benchmarks/full_build.py [--rows 1000000] [--samples 50] [--k 30] [--skip-row-shards]"""
import argparse
import json
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

from rosella_b200 import dist as rdist  # noqa: E402
from rosella_b200 import gpu, synth  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--rows", dest="n", type=int, default=1_000_000)
ap.add_argument("--samples", type=int, default=50)
ap.add_argument("--k", type=int, default=30)
ap.add_argument("--skip-row-shards", action="store_true")
a = ap.parse_args()
rank, local, world = int(os.environ.get("RANK", 0)), int(os.environ.get("LOCAL_RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
torch.cuda.set_device(local)
if world > 1:
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
tnf, cov, _ = synth.knn_tables(a.n, a.samples)
ctx = gpu.Context(local)
ctx.knn_load(tnf, cov)


def timed(fn):
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    t = time.perf_counter()
    out = fn()
    torch.cuda.synchronize()
    dt = torch.tensor([time.perf_counter() - t], device="cuda")
    if world > 1:
        dist.all_reduce(dt, op=dist.ReduceOp.MAX)
    return out, float(dt)


report = {"world": world, "n": a.n, "samples": a.samples, "k": a.k}
if world > 1:       # warm the communicators (all-gather ring, all-to-all peer connections) outside the timed regions
    kp = ctx.knn_kprime(gpu.GRAPH_DEPTH, a.k)
    timed(lambda: rdist.exchange_partial_lists(torch.zeros((a.n, kp), device="cuda"), torch.zeros((a.n, kp), dtype=torch.int32, device="cuda"), a.n, rank, world))
    timed(lambda: rdist.all_gather_rows(torch.zeros((rdist.shard_range(a.n, rank, world)[1] - rdist.shard_range(a.n, rank, world)[0], 2), device="cuda"), a.n, world))
(si, sd), t_sym = timed(lambda: rdist.build_depth_graph_symmetric(ctx, a.k, a.n, rank, world))
report["depth_symmetric_s"] = round(t_sym, 3)
if not a.skip_row_shards:
    os.environ["RG_DEPTH_BUILD"] = "batch"
    (bi, bd), t_rows = timed(lambda: rdist.build_graph_distributed(ctx, gpu.GRAPH_DEPTH, a.k, a.n, rank, world))
    os.environ.pop("RG_DEPTH_BUILD")
    report["depth_row_shards_s"] = round(t_rows, 3)
    report["depth_lists_identical"] = bool(torch.equal(si, bi) and torch.equal(sd, bd))
(ti, td), t_tnf = timed(lambda: rdist.build_graph_distributed(ctx, gpu.GRAPH_TNF, a.k, a.n, rank, world))
report["tnf_row_shards_s"] = round(t_tnf, 3)
report["graph_s"] = round(t_sym + t_tnf, 3)
report["contigs_per_s"] = round(a.n / (t_sym + t_tnf), 1)
if rank == 0:
    print(json.dumps(report), flush=True)
if world > 1:
    dist.barrier()
ctx.close()
if world > 1:
    dist.destroy_process_group()
sys.exit(0)
