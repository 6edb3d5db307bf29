#!/bin/bash
# N-GPU pass (N = $NGPU): rg_multi whole-graph build (one process) + the driver's bench contract under torchrun
O=gpurun_out; TAG=${TAG:-r2f}; N=${NGPU:-2}
MB_N=${MB_N:-200000}; MB_S=${MB_S:-50}; MB_K=${MB_K:-30}; MB_GBP=${MB_GBP:-1}
timeout 1500 python benchmarks/multi_build.py --gpus $N --n $MB_N --samples $MB_S --k $MB_K --tnf-gbp $MB_GBP --tag ${TAG}_multi_n$N \
    > $O/${TAG}_multi_n$N.json 2> $O/${TAG}_multi_n$N.err; echo "multi_build rc $?"; tail -c 2500 $O/${TAG}_multi_n$N.json; tail -3 $O/${TAG}_multi_n$N.err
if [ "${SKIP_BENCH:-0}" != "1" ]; then
timeout 1500 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29533 \
    bench.py --gpus $N --steps 3 --warmup 3 > $O/${TAG}_bench_1m_s50_n$N.json 2> $O/${TAG}_bench_1m_s50_n$N.err; echo "bench rc $?"
tail -c 3500 $O/${TAG}_bench_1m_s50_n$N.json; tail -5 $O/${TAG}_bench_1m_s50_n$N.err
OMP_NUM_THREADS=1 timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29534 \
    bench.py --impl reference --gpus $N --steps 1 --warmup 1 > $O/${TAG}_bench_reference_n$N.json 2> $O/${TAG}_bench_reference_n$N.err; echo "ref rc $?"
tail -c 600 $O/${TAG}_bench_reference_n$N.json
fi
