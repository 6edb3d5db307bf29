#!/bin/bash
# One GPU-box pass: parity tests, smoke, bench (ours + reference arm), ncu launch list and one full-set capture.
# Usage (from the repo root): gpurun --timeout 1500 -- 'bash benchmarks/gpu_check.sh TAG'
TAG=${1:-r1c}
O=gpurun_out
mkdir -p $O
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > $O/${TAG}_smi.txt 2>&1
timeout 600 python -m pytest tests -m gpu -x -q > $O/${TAG}_pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a $O/${TAG}_pytest_gpu.log
timeout 300 python __graft_entry__.py --smoke > $O/${TAG}_smoke.log 2>&1; echo "smoke rc=$?" | tee -a $O/${TAG}_smoke.log
timeout 900 python bench.py > $O/${TAG}_bench.json 2> $O/${TAG}_bench.err; echo "bench rc=$?"
tail -c 600 $O/${TAG}_bench.json
timeout 300 python bench.py --impl reference --steps 1 --warmup 1 > $O/${TAG}_bench_reference.json 2> $O/${TAG}_bench_reference.err; echo "ref rc=$?"
# ncu: launch list of the bench command (values printed under ncu are never bench values)
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/${TAG}_ncu_launches.csv \
    python bench.py --steps 2 --warmup 1 --no-cpu-baseline --tnf-gbp 1 > $O/${TAG}_ncu_bench.log 2>&1; echo "ncu list rc=$?"
# ncu: full set of every hot kernel once (short target, second pass only)
timeout 900 ncu --set full --clock-control none --import-source on -o $O/${TAG}_full -f \
    --kernel-name 'regex:select|rescore|tnf_count' --launch-skip 6 --launch-count 6 \
    python benchmarks/ncu_target.py --n 100000 --samples 50 --rows 9472 --gbp 0.3 --passes 2 > $O/${TAG}_ncu_full.log 2>&1; echo "ncu full rc=$?"
ls -la $O | tail -20
