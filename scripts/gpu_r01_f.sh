# round-1 final GPU pass (r01g): all GPU tests, smoke, full bench (both arms), launch lists + full captures of the two band kernels
set -x
O=gpurun_out/r01g; mkdir -p $O
nvidia-smi --query-gpu=name,clocks.max.sm,memory.total --format=csv > $O/gpu.txt
timeout 1500 python -m pytest tests -m gpu -x -q > $O/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> $O/pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > $O/smoke.log 2>&1; echo "smoke rc=$?" >> $O/smoke.log
timeout 900 python bench.py > $O/bench_full.log 2> $O/bench_full.err; echo "bench rc=$?" >> $O/bench_full.err
timeout 600 python bench.py --impl reference --steps 1 --warmup 0 > $O/bench_ref.log 2>&1
for k in ulpda myula; do
  if [ $k = ulpda ]; then CMD="python scripts/prof_image.py ulpda 256 2 band 0"; else CMD="python scripts/prof_image.py myula 256 10 band 512"; fi
  timeout 300 $CMD > $O/${k}_plain.log 2>&1 &&
  timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 60 --csv --log-file $O/launches_${k}_band.csv $CMD > $O/ncu_l_$k.log 2>&1
  timeout 1200 ncu --set full --clock-control none --import-source on -k regex:img_${k}_band -s 2 -c 1 -o $O/prof_${k}_band $CMD > $O/ncu_f_$k.log 2>&1
done
ls -la $O
