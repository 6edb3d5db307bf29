TAG=${1:-r1v}
O=gpurun_out; mkdir -p $O
cat > /tmp/k50.py <<'PY'
import json, os, sys
sys.path.insert(0, os.getcwd())
import numpy as np
from rosella_b200 import gpu, synth
import oracle
n, s, k = 400000, 4, 50
tnf, cov, _ = synth.knn_tables(n, s)
ctx = gpu.Context(0)
ctx.knn_load(tnf, None)
r = ctx.knn_search(gpu.GRAPH_TNF, k, 0, 18944)
r = ctx.knn_search(gpu.GRAPH_TNF, k, 0, 18944)
oi, od, _ = oracle.knn(tnf, oracle.TNF, k, qrows=np.arange(64, dtype=np.uint64))
np.save(sys.argv[1], r[0])
print(json.dumps({"impl": os.environ.get("RG_TNF_SELECT_IMPL", "tc"), "n": n, "k": k, "query_rows": 18944, "select_ms": ctx.timings()["select_ms"],
                  "kprime": ctx.knn_kprime(gpu.GRAPH_TNF, k), "oracle_idx_equal_64_rows": float((r[0][:64] == oi).mean())}))
PY
timeout 300 python /tmp/k50.py /tmp/tc.npy > $O/${TAG}_k50_tc.json 2> $O/${TAG}_k50.err; cat $O/${TAG}_k50_tc.json
RG_TNF_SELECT_IMPL=fp32 timeout 300 python /tmp/k50.py /tmp/fp32.npy > $O/${TAG}_k50_fp32.json 2>> $O/${TAG}_k50.err; cat $O/${TAG}_k50_fp32.json
python -c "import numpy as np; print('identical', bool(np.array_equal(np.load('/tmp/tc.npy'), np.load('/tmp/fp32.npy'))))" | tee $O/${TAG}_k50_identical.txt
tail -3 $O/${TAG}_k50.err
