# default bench line on one GPU
TAG=${1:-r1z}
O=gpurun_out; mkdir -p $O
timeout 900 python bench.py > $O/${TAG}_bench.json 2> $O/${TAG}_bench.err; echo "bench rc=$?"; cut -c1-300 $O/${TAG}_bench.json; tail -3 $O/${TAG}_bench.err
