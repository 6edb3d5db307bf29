# N-GPU pass (gpurun --gpus N): sharded-build parity over NCCL, then bench.py under torchrun
N=${1:-2}; TAG=${2:-r1e}
O=gpurun_out; mkdir -p $O
nvidia-smi topo -m > $O/${TAG}_topo.txt 2>&1
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 \
    benchmarks/multi_gpu_check.py > $O/${TAG}_multi_check_n$N.json 2> $O/${TAG}_multi_check_n$N.err; echo "check rc=$?"; cat $O/${TAG}_multi_check_n$N.json; tail -3 $O/${TAG}_multi_check_n$N.err
NCCL_DEBUG=INFO NCCL_DEBUG_SUBSYS=INIT timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 \
    bench.py --gpus $N --steps 3 --warmup 3 > $O/${TAG}_bench_n$N.out 2> $O/${TAG}_bench_n$N.err; echo "bench rc=$?"
grep '^{' $O/${TAG}_bench_n$N.out > $O/${TAG}_bench_n$N.json; cut -c1-700 $O/${TAG}_bench_n$N.json; grep -i "nvls\|via P2P\|NVLS" $O/${TAG}_bench_n$N.out $O/${TAG}_bench_n$N.err | head -5; tail -3 $O/${TAG}_bench_n$N.err
