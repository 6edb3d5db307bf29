#!/bin/bash
# 1-GPU bench pass: BASELINE.json configs[2] (default) and configs[1], both with whole_graph_full + parity_sample
O=gpurun_out; TAG=${TAG:-r2d}
timeout 1500 python bench.py --steps 3 --warmup 3 > $O/${TAG}_bench_1m_s50.json 2> $O/${TAG}_bench_1m_s50.err; echo "rc $?"; tail -c 3000 $O/${TAG}_bench_1m_s50.json; tail -5 $O/${TAG}_bench_1m_s50.err
timeout 900 python bench.py --n 100000 --samples 10 --k 30 --steps 3 --warmup 3 --tnf-gbp 0 --whole-graph-rows 0 > $O/${TAG}_bench_100k_s10.json 2> $O/${TAG}_bench_100k_s10.err; echo "rc $?"; tail -c 1500 $O/${TAG}_bench_100k_s10.json; tail -5 $O/${TAG}_bench_100k_s10.err
