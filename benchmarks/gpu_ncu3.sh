#!/bin/bash
# one full-set ncu capture of the ASCII TNF kernel (second launch) and the two sketch hash passes
O=gpurun_out; TAG=${TAG:-r2y}
timeout 150 ncu --set full --clock-control none --import-source on -o $O/${TAG}_full_k1_sketch -f \
    --kernel-name 'regex:tnf_count_ascii2|sketch_hash' --launch-skip 1 --launch-count 3 \
    python benchmarks/ncu_k1_sketch.py 0.5 > $O/${TAG}_ncu_k1_sketch.log 2>&1; echo "ncu rc=$?"
tail -3 $O/${TAG}_ncu_k1_sketch.log; ls -la $O | grep ${TAG}
