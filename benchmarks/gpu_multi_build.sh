# N-GPU whole-graph timing only (symmetric coverage build + TNF row shards)
N=${1:-4}; TAG=${2:-r1w}
O=gpurun_out; mkdir -p $O
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29513 \
    benchmarks/full_build.py --rows 1000000 --skip-row-shards > $O/${TAG}_full_build_1m_n$N.json 2> $O/${TAG}_full_build_1m_n$N.err; echo "build rc=$?"; grep '^{' $O/${TAG}_full_build_1m_n$N.json; tail -2 $O/${TAG}_full_build_1m_n$N.err
