#!/bin/bash
# 8-GPU pass: configs[2] and configs[4] through rg_multi (one process), then the bench contract under torchrun
O=gpurun_out; TAG=${TAG:-r2g}
nvidia-smi --query-gpu=index,name,memory.total --format=csv > $O/${TAG}_gpus.csv
timeout 900 python benchmarks/multi_build.py --gpus 8 --n 1000000 --samples 50 --k 30 --tnf-gbp 10 --tag ${TAG}_cfg2_n8 > $O/${TAG}_cfg2_n8.json 2> $O/${TAG}_cfg2_n8.err; echo "cfg2 rc $?"; tail -c 2500 $O/${TAG}_cfg2_n8.json; tail -3 $O/${TAG}_cfg2_n8.err
timeout 1500 python benchmarks/multi_build.py --gpus 8 --n 5000000 --samples 100 --k 50 --tag ${TAG}_cfg4_n8 > $O/${TAG}_cfg4_n8.json 2> $O/${TAG}_cfg4_n8.err; echo "cfg4 rc $?"; tail -c 2500 $O/${TAG}_cfg4_n8.json; tail -3 $O/${TAG}_cfg4_n8.err
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29535 \
    bench.py --gpus 8 --steps 3 --warmup 3 > $O/${TAG}_bench_1m_s50_n8.json 2> $O/${TAG}_bench_1m_s50_n8.err; echo "bench rc $?"
tail -c 3000 $O/${TAG}_bench_1m_s50_n8.json; tail -5 $O/${TAG}_bench_1m_s50_n8.err
