#!/bin/bash
# K1 ASCII kernel: parity tests, then the kernel alone on a resident set
O=gpurun_out; TAG=${TAG:-r2q}
timeout 400 python -m pytest tests/test_gpu_tnf.py -m gpu -x -q > $O/${TAG}_pytest_tnf.log 2>&1; echo "pytest exit $?" >> $O/${TAG}_pytest_tnf.log
tail -4 $O/${TAG}_pytest_tnf.log
timeout 400 python benchmarks/tnf_ascii_kernel.py ${GBP:-4} $O/${TAG}_tnf_ascii_kernel.json 2>&1 | tail -3
