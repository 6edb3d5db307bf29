# symmetric vs batch whole-graph coverage build at small sample counts
TAG=${1:-r1u}
O=gpurun_out; mkdir -p $O
for S in 1 3 10 25; do
  timeout 300 python benchmarks/sym_build.py 100000 $S 30 > $O/${TAG}_sym_100k_s$S.json 2> $O/${TAG}_sym_100k_s$S.err; cat $O/${TAG}_sym_100k_s$S.json; tail -2 $O/${TAG}_sym_100k_s$S.err
done
