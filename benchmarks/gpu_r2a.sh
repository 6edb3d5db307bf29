#!/bin/bash
# round-2 pass A: GPU parity tests + smoke + load-time cost of the collision patches at 1M x 50
O=gpurun_out; TAG=r2a
timeout 1500 python -m pytest tests -m gpu -x -q > $O/${TAG}_pytest_gpu.log 2>&1; echo "pytest exit $?" >> $O/${TAG}_pytest_gpu.log
tail -15 $O/${TAG}_pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > $O/${TAG}_smoke.log 2>&1; tail -3 $O/${TAG}_smoke.log
timeout 600 python - > $O/${TAG}_load_1m.json 2>$O/${TAG}_load_1m.err <<'PY'
import json, time
from rosella_b200 import gpu, synth
tnf, cov, _ = synth.knn_tables(1_000_000, 50)
ctx = gpu.Context(0)
out = {}
for certify in (1, 0, 1):
    ctx.set_option("certify", certify)
    t = time.perf_counter(); ctx.knn_load(tnf, cov); dt = time.perf_counter() - t
    tm = ctx.timings()
    out[f"load_certify{certify}"] = {"wall_s": dt, "h2d_ms": tm["h2d_ms"], "transform_ms": tm["transform_ms"], "extra_pairs": tm["extra_pairs"]}
for which, name in ((gpu.GRAPH_DEPTH, "depth"), (gpu.GRAPH_TNF, "tnf")):
    t = time.perf_counter(); ctx.knn_search(which, 30, 500_000, 500_000 + 18944); dt = time.perf_counter() - t
    tm = ctx.timings()
    out[name] = {"wall_s": dt, "select_ms": tm["select_ms"], "rescore_ms": tm["rescore_ms"], "rows_flagged": tm["rows_flagged"], "rows_exact": tm["rows_exact"]}
print(json.dumps(out))
PY
cat $O/${TAG}_load_1m.json; tail -3 $O/${TAG}_load_1m.err
