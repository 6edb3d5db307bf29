# BASELINE.json configs[4] shape on one GPU: 5M contigs x (136 + 100 samples), k = 50, one bench step per leg
TAG=${1:-r1t}
O=gpurun_out; mkdir -p $O
free -g | head -2
timeout 1500 python bench.py --n 5000000 --samples 100 --k 50 --steps 1 --warmup 1 --no-cpu-baseline --tnf-gbp 0 --whole-graph-rows 0 --no-whole-graph-full \
    > $O/${TAG}_bench_5m_s100_k50.json 2> $O/${TAG}_bench_5m_s100_k50.err; echo "rc=$?"
cut -c1-900 $O/${TAG}_bench_5m_s100_k50.json; tail -3 $O/${TAG}_bench_5m_s100_k50.err
