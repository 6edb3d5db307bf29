#!/bin/bash
# one GPU: the GPU suite, then the whole 1M x 50 graph (symmetric coverage build + TNF)
tag=${1:-r1q}
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/${tag}_pytest_gpu.log 2>&1
echo "pytest exit $?" >> gpurun_out/${tag}_pytest_gpu.log
tail -4 gpurun_out/${tag}_pytest_gpu.log
timeout 900 python benchmarks/full_build.py --rows 1000000 --skip-row-shards > gpurun_out/${tag}_full_build_1m_n1.json 2> gpurun_out/${tag}_full_build_1m_n1.err
cat gpurun_out/${tag}_full_build_1m_n1.json; tail -3 gpurun_out/${tag}_full_build_1m_n1.err
