set -x
O=gpurun_out/myula_band; mkdir -p $O
for e in band stream; do
timeout 300 python scripts/prof_image.py myula 256 10 $e 512 > $O/bench_512_$e.log 2>&1
done
cat $O/bench_512_*.log
