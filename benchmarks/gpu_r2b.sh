#!/bin/bash
O=gpurun_out; TAG=r2b
timeout 900 python - > $O/${TAG}_flags_1m.json 2>$O/${TAG}_flags_1m.err <<'PY'
import json, time
import numpy as np
from rosella_b200 import gpu, synth
tnf, cov, _ = synth.knn_tables(1_000_000, 50)
ctx = gpu.Context(0)
ctx.knn_load(tnf, cov)
out = {}
qb = 500_000
t = time.perf_counter(); di, dd, _ = ctx.knn_search(gpu.GRAPH_DEPTH, 30, qb, qb + 18944); dt = time.perf_counter() - t
tm = ctx.timings()
rows, forced, dk, eps = ctx.knn_flagged_rows()
order = np.argsort(rows)
out["depth"] = {"wall_s": dt, "select_ms": tm["select_ms"], "rescore_ms": tm["rescore_ms"], "flagged": int(len(rows)), "forced": int(forced.sum()),
                "rows": rows[order][:40].tolist(), "dk": dk[order][:40].tolist(), "eps": eps[order][:40].tolist(),
                "eps_quantiles": np.quantile(eps, [0, .1, .5, .9, 1]).tolist() if len(eps) else [],
                "dk_quantiles": np.quantile(dk, [0, .1, .5, .9, 1]).tolist() if len(dk) else []}
ctx.set_option("certify", 0)
ctx.knn_load(tnf, cov)
ui, ud, _ = ctx.knn_search(gpu.GRAPH_DEPTH, 30, qb, qb + 18944)
out["rows_changed_by_exact_pass"] = int((ui != di).any(axis=1).sum())
out["flagged_rows_changed"] = int((ui[rows] != di[rows]).any(axis=1).sum()) if len(rows) else 0
print(json.dumps(out))
PY
cat $O/${TAG}_flags_1m.json; tail -3 $O/${TAG}_flags_1m.err
timeout 1500 python -m pytest tests -m gpu -q > $O/${TAG}_pytest_gpu.log 2>&1; echo "pytest exit $?" >> $O/${TAG}_pytest_gpu.log
tail -15 $O/${TAG}_pytest_gpu.log
