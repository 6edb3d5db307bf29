#!/bin/bash
# 1-GPU pass: the allocation-free entry point alone in a fresh process, then all GPU tests (it is collected last) + smoke
O=gpurun_out; TAG=${TAG:-r2p}
timeout 300 python -m pytest tests/test_zz_no_alloc.py -m gpu -q > $O/${TAG}_noalloc_alone.log 2>&1; echo "pytest exit $?" >> $O/${TAG}_noalloc_alone.log
tail -5 $O/${TAG}_noalloc_alone.log
timeout 900 python -m pytest tests -m gpu -x -q > $O/${TAG}_pytest_gpu.log 2>&1; echo "pytest exit $?" >> $O/${TAG}_pytest_gpu.log
tail -8 $O/${TAG}_pytest_gpu.log
timeout 200 python -c "import __graft_entry__ as g; g.smoke()" > $O/${TAG}_smoke.log 2>&1; echo "smoke exit $?" >> $O/${TAG}_smoke.log; tail -3 $O/${TAG}_smoke.log
