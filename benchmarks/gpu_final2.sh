#!/bin/bash
# Last 1-GPU pass of the round on the final code: all GPU tests (+ smoke), then the default bench line
O=gpurun_out; TAG=${TAG:-r2v}
timeout 900 python -m pytest tests -m gpu -x -q > $O/${TAG}_pytest_gpu.log 2>&1; echo "pytest exit $?" >> $O/${TAG}_pytest_gpu.log
tail -6 $O/${TAG}_pytest_gpu.log
timeout 200 python -c "import __graft_entry__ as g; g.smoke()" > $O/${TAG}_smoke.log 2>&1; echo "smoke exit $?" >> $O/${TAG}_smoke.log; tail -2 $O/${TAG}_smoke.log
timeout 900 python bench.py > $O/${TAG}_bench_1m_s50.json 2> $O/${TAG}_bench_1m_s50.err; echo "bench rc=$?"
tail -c 2500 $O/${TAG}_bench_1m_s50.json
