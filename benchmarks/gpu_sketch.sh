#!/bin/bash
# MinHash sketching: parity tests, then throughput
O=gpurun_out; TAG=${TAG:-r2s}
timeout 600 python -m pytest tests/test_gpu_sketch.py -m gpu -x -q > $O/${TAG}_pytest_sketch.log 2>&1; echo "pytest exit $?" >> $O/${TAG}_pytest_sketch.log
tail -15 $O/${TAG}_pytest_sketch.log
timeout 400 python benchmarks/sketch_arm.py ${GBP:-2} $O/${TAG}_sketch_arm.json 2>&1 | tail -3
