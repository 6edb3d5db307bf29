#!/bin/bash
# Round-end 1-GPU pass: the default bench line, the reference arm, the ncu launch list of the bench command, full-set
# captures of the hot kernels at the bench shape, and the rank statistic timed at 50k contigs.
#   gpurun --timeout 2700 -- 'bash benchmarks/gpu_final.sh r2j'
TAG=${1:-r2j}
O=gpurun_out
mkdir -p $O
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > $O/${TAG}_smi.txt 2>&1
timeout 900 python bench.py > $O/${TAG}_bench_1m_s50.json 2> $O/${TAG}_bench_1m_s50.err; echo "bench rc=$?"
tail -c 1500 $O/${TAG}_bench_1m_s50.json
timeout 300 python bench.py --impl reference --steps 1 --warmup 1 > $O/${TAG}_bench_reference_1m_s50.json 2> $O/${TAG}_bench_reference.err; echo "ref rc=$?"
# ncu: launch list of the bench command (values printed under ncu are never bench values)
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file $O/${TAG}_ncu_launches.csv \
    python bench.py --steps 2 --warmup 1 --no-cpu-baseline --tnf-gbp 1 --whole-graph-rows 40000 --no-whole-graph-full > $O/${TAG}_ncu_bench.log 2>&1; echo "ncu list rc=$?"
# ncu: full set of the two select kernels at the bench shape
timeout 900 ncu --set full --clock-control none --import-source on -o $O/${TAG}_full_select -f \
    --kernel-name 'regex:depth_select2' --launch-skip 1 --launch-count 1 \
    python bench.py --steps 1 --warmup 1 --no-cpu-baseline --tnf-gbp 0 --whole-graph-rows 0 --no-whole-graph-full > $O/${TAG}_ncu_full_select.log 2>&1; echo "ncu select rc=$?"
timeout 600 ncu --set full --clock-control none --import-source on -o $O/${TAG}_full_tc -f \
    --kernel-name 'regex:tnf_select_tc' --launch-skip 1 --launch-count 1 \
    python bench.py --steps 1 --warmup 1 --no-cpu-baseline --tnf-gbp 0 --whole-graph-rows 0 --no-whole-graph-full > $O/${TAG}_ncu_full_tc.log 2>&1; echo "ncu tc rc=$?"
# ncu: the remaining own kernels on the short target (f64 re-evaluation + certificate, exhaustive pass, collision scan,
# TNF counting, symmetric-build kernels)
timeout 600 ncu --set full --clock-control none --import-source on -o $O/${TAG}_full_other -f \
    --kernel-name 'regex:rescore_kernel|exact_rows|collide_kernel|tnf_count|depth_tile|topk_reduce' --launch-count 14 \
    python benchmarks/ncu_target.py --n 100000 --samples 50 --rows 9472 --gbp 0.3 --passes 1 --sym-rows 60000 > $O/${TAG}_ncu_full_other.log 2>&1; echo "ncu other rc=$?"
# flight's rank statistic at 50k contigs x 10 samples (1.25e9 pairs, everything f64)
timeout 600 python - > $O/${TAG}_rank_stat_50k.json 2> $O/${TAG}_rank_stat_50k.err <<'PY'
import json, time
import numpy as np
from rosella_b200 import gpu, synth
n, s = 50_000, 10
tnf, cov, _ = synth.knn_tables(n, s, seed=33, flight=True)
lengths = np.rint(np.exp(np.random.default_rng(1).uniform(np.log(2000.0), np.log(500000.0), n)))
ctx = gpu.Context(0)
ctx.rank_stat(tnf[:2000], cov[:2000], lengths[:2000], k=30, condensed=False)       # kernel images
t = time.perf_counter(); _, idx, sk = ctx.rank_stat(tnf, cov, lengths, k=30, condensed=False); wall = time.perf_counter() - t
tm = ctx.timings()
pairs = n * (n - 1) // 2
print(json.dumps({"n": n, "samples": s, "pairs": pairs, "wall_s": wall, "distances_ms": tm["select_ms"], "ranks_and_statistic_ms": tm["rescore_ms"],
                  "us_per_pair": wall / pairs * 1e6, "output": "30 smallest statistics per row (sparse), f64",
                  "reference": "flight pdist + numba: 10.6 + 0.9 us per pair single-threaded at S = 10 (profiles/r2_c1_flight_cpu.json)"}))
PY
cat $O/${TAG}_rank_stat_50k.json; tail -2 $O/${TAG}_rank_stat_50k.err
ls -la $O | grep ${TAG}
