#!/bin/bash
# Round-end GPU pass: parity tests, smoke, both bench arms, ncu launch list of the bench command and full-set captures
# of the hot kernels at the bench shape.  Usage: gpurun --timeout 2400 -- 'bash benchmarks/gpu_final.sh TAG'
TAG=${1:-r1z}
O=gpurun_out
mkdir -p $O
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > $O/${TAG}_smi.txt 2>&1
timeout 600 python -m pytest tests -m gpu -x -q > $O/${TAG}_pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a $O/${TAG}_pytest_gpu.log
timeout 300 python __graft_entry__.py --smoke > $O/${TAG}_smoke.log 2>&1; echo "smoke rc=$?" | tee -a $O/${TAG}_smoke.log
timeout 900 python bench.py > $O/${TAG}_bench.json 2> $O/${TAG}_bench.err; echo "bench rc=$?"
timeout 300 python bench.py --impl reference --steps 1 --warmup 1 > $O/${TAG}_bench_reference.json 2> $O/${TAG}_bench_reference.err; echo "ref rc=$?"
# ncu: launch list of the bench command (values printed under ncu are never bench values)
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/${TAG}_ncu_launches.csv \
    python bench.py --steps 2 --warmup 1 --no-cpu-baseline --tnf-gbp 1 --whole-graph-rows 40000 --no-whole-graph-full > $O/${TAG}_ncu_bench.log 2>&1; echo "ncu list rc=$?"
# ncu: full set of the two select kernels at the bench shape (first timed step)
timeout 1200 ncu --set full --clock-control none --import-source on -o $O/${TAG}_full_select -f \
    --kernel-name 'regex:depth_select2' --launch-skip 1 --launch-count 1 \
    python bench.py --steps 1 --warmup 1 --no-cpu-baseline --tnf-gbp 0 --whole-graph-rows 0 --no-whole-graph-full > $O/${TAG}_ncu_full_select.log 2>&1; echo "ncu select rc=$?"
timeout 600 ncu --set full --clock-control none --import-source on -o $O/${TAG}_full_tc -f \
    --kernel-name 'regex:tnf_select_tc' --launch-skip 1 --launch-count 1 \
    python bench.py --steps 1 --warmup 1 --no-cpu-baseline --tnf-gbp 0 --whole-graph-rows 0 --no-whole-graph-full > $O/${TAG}_ncu_full_tc.log 2>&1; echo "ncu tc rc=$?"
# ncu: full set of the remaining kernels on the short target (second pass)
timeout 600 ncu --set full --clock-control none --import-source on -o $O/${TAG}_full_other -f \
    --kernel-name 'regex:rescore|tnf_count|tnf_freq|transform' --launch-skip 8 --launch-count 6 \
    python benchmarks/ncu_target.py --n 100000 --samples 50 --rows 9472 --gbp 0.3 --passes 2 > $O/${TAG}_ncu_full_other.log 2>&1; echo "ncu other rc=$?"
# ncu: full set of the symmetric whole-graph kernels (evaluation of one block launch, forward and transposed reducers)
timeout 600 ncu --set full --clock-control none --import-source on -o $O/${TAG}_full_sym -f \
    --kernel-name 'regex:depth_tile|topk_reduce' --launch-skip 3 --launch-count 3 \
    python benchmarks/ncu_target.py --n 60000 --samples 50 --rows 2048 --gbp 0.05 --passes 1 --sym-rows 60000 > $O/${TAG}_ncu_full_sym.log 2>&1; echo "ncu sym rc=$?"
ls -la $O | tail -20
