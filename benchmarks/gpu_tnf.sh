# TNF leg on the GPU box: parity tests, then the 4 Gbp counting benchmark
TAG=${1:-r1r}
O=gpurun_out; mkdir -p $O
timeout 600 python -m pytest tests/test_gpu_tnf.py -m gpu -x -q > $O/${TAG}_pytest_tnf.log 2>&1; echo "pytest rc=$?"; tail -5 $O/${TAG}_pytest_tnf.log
timeout 300 python benchmarks/tnf_only.py --gbp 4 > $O/${TAG}_tnf.json 2> $O/${TAG}_tnf.err; echo "rc=$?"; cut -c1-600 $O/${TAG}_tnf.json; tail -3 $O/${TAG}_tnf.err
