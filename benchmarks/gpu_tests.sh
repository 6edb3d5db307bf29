# all GPU parity tests on the box: gpurun --timeout 900 -- 'bash benchmarks/gpu_tests.sh TAG'
TAG=${1:-r1}; O=gpurun_out; mkdir -p $O
timeout 800 python -m pytest tests -m gpu -x -q > $O/${TAG}_pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -25 $O/${TAG}_pytest_gpu.log
# kernel timings on the ncu target shape (no profiler): S=50 k=30 and the configs[4] shape S=100 k=50
timeout 300 python benchmarks/ncu_target.py --n 100000 --samples 50 --rows 9472 --gbp 0.1 --passes 2 2>&1 | tail -1
timeout 300 python benchmarks/ncu_target.py --n 100000 --samples 100 --k 50 --rows 9472 --gbp 0.1 --passes 2 2>&1 | tail -1
