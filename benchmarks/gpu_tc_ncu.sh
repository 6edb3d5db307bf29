O=gpurun_out; mkdir -p $O
timeout 600 ncu --set full --clock-control none --import-source on -o $O/r1h_tc_full -f --kernel-name 'regex:tnf_select_tc' --launch-skip 1 --launch-count 1 python benchmarks/ncu_target.py --n 200000 --samples 4 --rows 18944 --gbp 0.02 --passes 2 > $O/r1h_ncu_tc.log 2>&1; echo "ncu rc=$?"; tail -3 $O/r1h_ncu_tc.log
