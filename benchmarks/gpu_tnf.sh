# TNF leg on the GPU box: parity tests, then the 4 Gbp counting benchmark
O=gpurun_out; mkdir -p $O
timeout 600 python -m pytest tests/test_gpu_tnf.py tests/test_gpu_knn.py -m gpu -x -q > $O/r1d_pytest.log 2>&1; echo "pytest rc=$?"; tail -15 $O/r1d_pytest.log
timeout 300 python benchmarks/tnf_only.py --gbp 4 > $O/r1d_tnf_v2.json 2> $O/r1d_tnf_v2.err; echo "v2 rc=$?"; cut -c1-400 $O/r1d_tnf_v2.json; tail -3 $O/r1d_tnf_v2.err
