# all GPU parity tests on the box: gpurun --timeout 900 -- 'bash benchmarks/gpu_tests.sh TAG'
TAG=${1:-r1}; O=gpurun_out; mkdir -p $O
timeout 800 python -m pytest tests -m gpu -x -q > $O/${TAG}_pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -25 $O/${TAG}_pytest_gpu.log
