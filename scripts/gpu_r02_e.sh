#!/bin/bash
mkdir -p gpurun_out/r02e
timeout 600 python scripts/debug_map.py 400 > gpurun_out/r02e/debug_map.log 2>&1
CMD="python bench.py --steps 1 --warmup 3 --K 2000 --no-image --no-cpu --no-f64"
$CMD > gpurun_out/r02e/plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 80 --csv --log-file gpurun_out/r02e/launches_mix.csv $CMD > gpurun_out/r02e/ncu1.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:mix_chain -s 12 -c 4 -o gpurun_out/r02e/prof_mix $CMD > gpurun_out/r02e/ncu2.log 2>&1
ls -la gpurun_out/r02e; tail -n 20 gpurun_out/r02e/debug_map.log
