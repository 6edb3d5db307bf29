# bench.py (ours + reference arm) on the box: gpurun --timeout 1200 -- 'bash benchmarks/gpu_bench.sh TAG'
TAG=${1:-r1}; O=gpurun_out; mkdir -p $O
timeout 900 python bench.py > $O/${TAG}_bench.json 2> $O/${TAG}_bench.err; echo "bench rc=$?"; tail -3 $O/${TAG}_bench.err
python - <<PY
import json
d=json.load(open("$O/${TAG}_bench.json"))
print({k:d[k] for k in ("value","ms_per_step","gpu_launches","clocks")}); print(d["e2e"]); print(d["roofline"]); print(d["kernels_ms"]); print(d.get("cpu_baseline"))
t=d.get("tnf",{}); print({k:t.get(k) for k in ("value","kernel_ms","ascii_kernel_gbps","e2e","roofline")})
PY
