#!/bin/bash
# r02i: final mixture kernels (no-shift, 2 CTAs/SM): tests, full bench (both arms), fp32 divergence table, ncu capture
mkdir -p gpurun_out/r02i
timeout 900 python -m pytest tests/test_mix_production_gpu.py tests/test_mix_gpu.py tests/test_stats_gpu.py -q > gpurun_out/r02i/mix_tests.log 2>&1
echo "rc=$?" >> gpurun_out/r02i/mix_tests.log
timeout 1500 python bench.py > gpurun_out/r02i/bench.json 2> gpurun_out/r02i/bench.err
echo "bench rc=$?" >> gpurun_out/r02i/bench.err
timeout 600 python bench.py --impl reference --steps 1 --warmup 0 > gpurun_out/r02i/bench_reference.json 2>> gpurun_out/r02i/bench.err
timeout 900 python scripts/fp32_divergence.py > gpurun_out/r02i/fp32_divergence.log 2>&1
cp gpurun_out/fp32_divergence.json gpurun_out/r02i/ 2>/dev/null
CMD="python bench.py --steps 1 --warmup 3 --K 2000 --no-image --no-cpu --no-f64"
$CMD > gpurun_out/r02i/plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 80 --csv --log-file gpurun_out/r02i/launches_mix.csv $CMD > gpurun_out/r02i/ncu1.log 2>&1
timeout 900 ncu --set full --clock-control none -k regex:mix_chain -s 12 -c 4 -o /tmp/prof_mix $CMD > gpurun_out/r02i/ncu2.log 2>&1
python scripts/ncu_summary.py /tmp/prof_mix.ncu-rep gpurun_out/r02i/mix_ncu_full_summary.csv
tail -n 3 gpurun_out/r02i/mix_tests.log; tail -n 5 gpurun_out/r02i/bench.err; head -c 1500 gpurun_out/r02i/bench.json; tail -n 5 gpurun_out/r02i/fp32_divergence.log
