O=gpurun_out; mkdir -p $O
timeout 600 ncu --set full --clock-control none --import-source on -o $O/r1d_tnf_packed_full -f --kernel-name 'regex:tnf_count_packed' --launch-skip 1 --launch-count 1 python benchmarks/tnf_only.py --gbp 0.5 --steps 1 --warmup 1 > $O/r1d_ncu_tnf.log 2>&1; echo "ncu rc=$?"
timeout 600 python -m pytest tests/test_gpu_knn.py -m gpu -x -q > $O/r1d_pytest_knn.log 2>&1; echo "pytest rc=$?"; tail -5 $O/r1d_pytest_knn.log
