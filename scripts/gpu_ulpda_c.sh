set -x
O=gpurun_out/ulpda_c; mkdir -p $O
for c in 256 259 148 296; do
timeout 300 python scripts/prof_image.py ulpda $c 2 band 0 > $O/bench_$c.log 2>&1
done
cat $O/bench_*.log
