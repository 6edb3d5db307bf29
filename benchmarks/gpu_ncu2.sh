#!/bin/bash
# ncu pass 2: launch list of the bench STEP kernels (the load's collision scan alone is > 1 000 launches and filled the
# first list), and full-set captures of the kernels the first pass missed.
TAG=${1:-r2j}
O=gpurun_out
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file $O/${TAG}_ncu_launches_step.csv \
    --kernel-name 'regex:depth_select2|rescore_kernel|exact_rows|tnf_select|reset_flag|depth_tile|topk_reduce|list_init|tnf_count|tnf_freq|transform' -c 300 \
    python bench.py --steps 2 --warmup 1 --no-cpu-baseline --tnf-gbp 1 --whole-graph-rows 40000 --no-whole-graph-full > $O/${TAG}_ncu_bench_step.log 2>&1; echo "ncu step list rc=$?"
timeout 600 ncu --set full --clock-control none --import-source on -o $O/${TAG}_full_other2 -f \
    --kernel-name 'regex:rescore_kernel|exact_rows|tnf_count|depth_tile|topk_reduce' --launch-count 12 \
    python benchmarks/ncu_target.py --n 100000 --samples 50 --rows 9472 --gbp 0.3 --passes 1 --sym-rows 60000 > $O/${TAG}_ncu_full_other2.log 2>&1; echo "ncu other rc=$?"
ls -la $O | grep ${TAG}_ncu_launches_step; ls -la $O | grep other2
