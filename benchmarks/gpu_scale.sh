#!/bin/bash
# rg_multi strong-scaling sweep at BASELINE.json configs[2] on the GPUs of this box: N = 1, 2, 4 (8: benchmarks/gpu_n8.sh).
# At N = 4 the sample saved for the offline oracle check holds 2 048 rows.
O=gpurun_out; TAG=${TAG:-r2k}
for N in ${NS:-1 2 4}; do
  SR=64; if [ "$N" = "4" ]; then SR=2048; fi
  timeout 900 python benchmarks/multi_build.py --gpus $N --n 1000000 --samples 50 --k 30 --sample-rows $SR --tnf-gbp ${GBP:-4} --tag ${TAG}_cfg2_n$N \
      > $O/${TAG}_cfg2_n$N.json 2> $O/${TAG}_cfg2_n$N.err; echo "N=$N rc $?"; python -c "
import json,sys
d=json.load(open('$O/${TAG}_cfg2_n$N.json'))
print({k:d[k] for k in ('gpus','seconds','contigs_per_s','rows_flagged')}, d['phases_ms_max_over_gpus']['depth'], d.get('tnf',{}).get('value'), d['parity_bit_rows'])"
done
