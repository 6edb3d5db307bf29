# N-GPU pass for the symmetric coverage build: parity check, then whole-graph timing
N=${1:-2}; TAG=${2:-r1n}; NB=${3:-200000}
O=gpurun_out; mkdir -p $O
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 \
    benchmarks/multi_gpu_check.py --rows 40000 > $O/${TAG}_multi_check_n$N.json 2> $O/${TAG}_multi_check_n$N.err; echo "check rc=$?"; cat $O/${TAG}_multi_check_n$N.json; tail -3 $O/${TAG}_multi_check_n$N.err
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29513 \
    benchmarks/full_build.py --rows $NB > $O/${TAG}_full_build_n$N.json 2> $O/${TAG}_full_build_n$N.err; echo "build rc=$?"; cat $O/${TAG}_full_build_n$N.json; tail -3 $O/${TAG}_full_build_n$N.err
