#!/bin/bash
# symmetric whole-graph depth build: parity tests, then timing against the batch kernel
tag=${1:-sym}
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_knn.py -m gpu -x -q > gpurun_out/${tag}_pytest.log 2>&1
echo "pytest exit $?" >> gpurun_out/${tag}_pytest.log
tail -5 gpurun_out/${tag}_pytest.log
timeout 600 python benchmarks/sym_build.py 200000 50 30 > gpurun_out/${tag}_build_200k.json 2> gpurun_out/${tag}_build_200k.err
cat gpurun_out/${tag}_build_200k.json; tail -3 gpurun_out/${tag}_build_200k.err
