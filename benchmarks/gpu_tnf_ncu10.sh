# ncu full set of the TNF counting kernels at the bench's 10 Gbp shape (for roofline.traffic of the tnf key)
TAG=${1:-r1z}
O=gpurun_out; mkdir -p $O
timeout 900 ncu --set full --clock-control none --import-source on -o $O/${TAG}_full_tnf10g -f \
    --kernel-name 'regex:tnf_count' --launch-count 2 \
    python benchmarks/ncu_target.py --n 20000 --samples 4 --rows 64 --gbp 10 --passes 1 > $O/${TAG}_ncu_full_tnf10g.log 2>&1; echo "ncu rc=$?"; tail -3 $O/${TAG}_ncu_full_tnf10g.log
