#!/bin/bash
# r02b: A/B of the TV loop bound (niter vs niter+1 passes) against the reference figures + new parity tests
mkdir -p gpurun_out/r02b
cp gpurun_out/pinned_*.json gpurun_out/r02b/ 2>/dev/null
LMC_B200_LIB=$PWD/lmc_atomi_b200/_lmc_b200_tv0.so timeout 900 python -m pytest tests/test_pinned_gpu.py -q > gpurun_out/r02b/pinned_tv0.log 2>&1
for a in MYULA ULPDA; do mv gpurun_out/pinned_$a.json gpurun_out/r02b/pinned_${a}_niter_passes.json; done
timeout 900 python -m pytest tests/test_pinned_gpu.py -q > gpurun_out/r02b/pinned_tv1.log 2>&1
for a in MYULA ULPDA; do cp gpurun_out/pinned_$a.json gpurun_out/r02b/pinned_${a}_niter_plus_1_passes.json; done
timeout 1200 python -m pytest tests/test_bench_shapes_gpu.py tests/test_chunked_stats_gpu.py tests/test_mix_production_gpu.py -q -s > gpurun_out/r02b/new_tests.log 2>&1
echo "rc=$?" >> gpurun_out/r02b/new_tests.log
tail -3 gpurun_out/r02b/*.log
