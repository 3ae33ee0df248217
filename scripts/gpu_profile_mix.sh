set -x
python bench.py > gpurun_out/bench_full.log 2>gpurun_out/bench_full.err
CMD="python bench.py --steps 1 --warmup 3 --K 2000 --no-image --no-cpu"
$CMD > gpurun_out/plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 80 --csv --log-file gpurun_out/launches_mix.csv $CMD > gpurun_out/ncu1.log 2>&1
$CMD > gpurun_out/plain2.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:mix_chain -s 12 -c 4 -o gpurun_out/prof_mix $CMD > gpurun_out/ncu2.log 2>&1
ls -la gpurun_out
