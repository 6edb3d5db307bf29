#!/bin/bash
# 2-GPU pass: every GPU test (the 2-GPU cases of tests/test_multi_gpu.py do not skip here)
O=gpurun_out; TAG=${TAG:-r2c}
nvidia-smi -L > $O/${TAG}_gpus.txt
timeout 1700 python -m pytest tests -m gpu -q -x > $O/${TAG}_pytest_gpu_n2.log 2>&1; echo "pytest exit $?" >> $O/${TAG}_pytest_gpu_n2.log
tail -25 $O/${TAG}_pytest_gpu_n2.log
