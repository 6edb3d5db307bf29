#!/usr/bin/env python
"""N-GPU parity of the sharded build (run under torchrun, NCCL): every rank searches its block of query rows with
rg_knn_search_dev, the slabs are all-gathered over NVLink, and the gathered lists must equal, bit for bit, the
lists one GPU computes for all rows.
  python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 \
      benchmarks/multi_gpu_check.py [--rows 20000] [--samples 10] [--k 30]"""
import argparse
import json
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

import numpy as np  # noqa: E402
import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

from rosella_b200 import dist as rdist  # noqa: E402
from rosella_b200 import gpu, synth  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--rows", dest="n", type=int, default=20000)
ap.add_argument("--samples", type=int, default=10)
ap.add_argument("--k", type=int, default=30)
a = ap.parse_args()
rank, local, world = int(os.environ["RANK"]), int(os.environ["LOCAL_RANK"]), int(os.environ["WORLD_SIZE"])
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
tnf, cov, _ = synth.knn_tables(a.n, a.samples)
ctx = gpu.Context(local)
ctx.knn_load(tnf, cov)
report = {"world": world, "n": a.n, "samples": a.samples, "k": a.k, "graphs": {}}
for name, which in (("depth", gpu.GRAPH_DEPTH), ("depth_symmetric", gpu.GRAPH_DEPTH), ("tnf", gpu.GRAPH_TNF),
                    ("tnf_euclid", gpu.GRAPH_TNF_EUCLID)):
    torch.cuda.synchronize(); dist.barrier()
    t = time.perf_counter()
    if name == "depth_symmetric":      # block shares + all-to-all of partial lists + merge + all-gather
        gi, gd = rdist.build_depth_graph_symmetric(ctx, a.k, a.n, rank, world)
    else:                              # query-row shards + all-gather
        gi, gd = rdist.build_graph_distributed(ctx, which, a.k, a.n, rank, world)
    torch.cuda.synchronize(); dist.barrier()
    dt = time.perf_counter() - t
    # every rank must hold the same gathered graph
    chk = torch.stack([gi.to(torch.int64).sum(), (gi.to(torch.int64) * torch.arange(gi.numel(), device=gi.device).view_as(gi) % 1000003).sum()])
    lo, hi = chk.clone(), chk.clone()
    dist.all_reduce(lo, op=dist.ReduceOp.MIN); dist.all_reduce(hi, op=dist.ReduceOp.MAX)
    same_on_all_ranks = bool((lo == hi).all())
    if rank == 0:
        if name == "depth_symmetric":
            os.environ["RG_DEPTH_BUILD"] = "batch"          # the reference lists come from the batch kernel
        fi, fd, _ = ctx.knn_search(which, a.k)              # one GPU, all rows
        os.environ.pop("RG_DEPTH_BUILD", None)
        report["graphs"][name] = {
            "seconds": dt, "same_on_all_ranks": same_on_all_ranks,
            "idx_equal_single_gpu": bool(np.array_equal(gi.cpu().numpy().astype(np.uint32), fi)),
            "dist_equal_single_gpu": bool(np.array_equal(gd.cpu().numpy(), fd))}
if rank == 0:
    ok = all(g["same_on_all_ranks"] and g["idx_equal_single_gpu"] and g["dist_equal_single_gpu"] for g in report["graphs"].values())
    report["ok"] = ok
    print(json.dumps(report), flush=True)
dist.barrier()
ctx.close()
dist.destroy_process_group()
sys.exit(0)
