#!/bin/bash
# r02f: ncu capture of the packed mixture kernels (summaries made on the box: the .ncu-rep is too large to travel),
# MAP pinned test with the early/tail gates, W2 exact check at K=5000
mkdir -p gpurun_out/r02f
CMD="python bench.py --steps 1 --warmup 3 --K 2000 --no-image --no-cpu --no-f64"
$CMD > gpurun_out/r02f/plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 80 --csv --log-file gpurun_out/r02f/launches_mix.csv $CMD > gpurun_out/r02f/ncu1.log 2>&1
timeout 900 ncu --set full --clock-control none -k regex:mix_chain -s 12 -c 4 -o /tmp/prof_mix $CMD > gpurun_out/r02f/ncu2.log 2>&1
python scripts/ncu_summary.py /tmp/prof_mix.ncu-rep gpurun_out/r02f/mix_ncu_full_summary.csv
ncu -i /tmp/prof_mix.ncu-rep --page raw --csv > gpurun_out/r02f/mix_raw.csv 2>/dev/null
timeout 900 python -m pytest tests/test_pinned_gpu.py -q -k map > gpurun_out/r02f/pinned_map.log 2>&1
echo "rc=$?" >> gpurun_out/r02f/pinned_map.log
cp gpurun_out/pinned_MAP.json gpurun_out/r02f/
timeout 900 python -m pytest tests/test_w2_gpu.py -q -s -k driver_scale > gpurun_out/r02f/w2.log 2>&1
echo "rc=$?" >> gpurun_out/r02f/w2.log
tail -n 5 gpurun_out/r02f/pinned_map.log gpurun_out/r02f/w2.log
