# N-GPU round-end pass: parity of all graphs (incl. the symmetric coverage build), whole 1M graph timing, bench.py
N=${1:-8}; TAG=${2:-r1s}
O=gpurun_out; mkdir -p $O
nvidia-smi topo -m > $O/${TAG}_topo_n$N.txt 2>&1
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 \
    benchmarks/multi_gpu_check.py --rows 40000 > $O/${TAG}_multi_check_n$N.json 2> $O/${TAG}_multi_check_n$N.err; echo "check rc=$?"; cut -c1-600 $O/${TAG}_multi_check_n$N.json
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29513 \
    benchmarks/full_build.py --rows 1000000 > $O/${TAG}_full_build_1m_n$N.json 2> $O/${TAG}_full_build_1m_n$N.err; echo "build rc=$?"; cat $O/${TAG}_full_build_1m_n$N.json; tail -3 $O/${TAG}_full_build_1m_n$N.err
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 \
    bench.py --gpus $N --steps 3 --warmup 3 > $O/${TAG}_bench_n$N.out 2> $O/${TAG}_bench_n$N.err; echo "bench rc=$?"
grep '^{' $O/${TAG}_bench_n$N.out > $O/${TAG}_bench_n$N.json; cut -c1-700 $O/${TAG}_bench_n$N.json; tail -3 $O/${TAG}_bench_n$N.err
