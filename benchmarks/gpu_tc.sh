# tensor-core TNF select bring-up: one small parity test first (short timeout: a pipeline deadlock must not hang the box)
O=gpurun_out; mkdir -p $O
timeout 90 python -m pytest tests/test_gpu_knn.py -m gpu -x -q -k "synthetic_3000" > $O/r1h_tc_small.log 2>&1; rc=$?; echo "small rc=$rc"; tail -30 $O/r1h_tc_small.log
if [ $rc -ne 0 ]; then exit 0; fi
timeout 300 python -m pytest tests/test_gpu_knn.py tests/test_distributed.py tests/test_host_mirror.py -m gpu -x -q > $O/r1h_tc_knn.log 2>&1; echo "knn rc=$?"; tail -30 $O/r1h_tc_knn.log
timeout 300 python benchmarks/ncu_target.py --n 200000 --samples 10 --rows 18944 --gbp 0.05 --passes 2 2>&1 | tail -1
RG_TNF_SELECT_IMPL=fp32 timeout 300 python benchmarks/ncu_target.py --n 200000 --samples 10 --rows 18944 --gbp 0.05 --passes 2 2>&1 | tail -1
