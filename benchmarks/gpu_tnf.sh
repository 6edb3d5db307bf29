O=gpurun_out; mkdir -p $O
timeout 600 python -m pytest tests/test_gpu_tnf.py -m gpu -x -q > $O/r1d_pytest_tnf.log 2>&1; echo "pytest rc=$?"; tail -15 $O/r1d_pytest_tnf.log
timeout 300 python benchmarks/tnf_only.py --gbp 4 > $O/r1d_tnf_v2.json 2> $O/r1d_tnf_v2.err; echo "v2 rc=$?"; cat $O/r1d_tnf_v2.json; tail -3 $O/r1d_tnf_v2.err
RG_TNF_IMPL=1 timeout 300 python benchmarks/tnf_only.py --gbp 4 > $O/r1d_tnf_v1.json 2> $O/r1d_tnf_v1.err; echo "v1 rc=$?"; cat $O/r1d_tnf_v1.json
