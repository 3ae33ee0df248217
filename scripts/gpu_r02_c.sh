#!/bin/bash
mkdir -p gpurun_out/r02c
./build_tmp/ubench_issue > gpurun_out/r02c/ubench_issue.txt 2>&1
timeout 900 python -m pytest tests/test_bench_shapes_gpu.py tests/test_mix_production_gpu.py tests/test_pinned_gpu.py -q > gpurun_out/r02c/tests.log 2>&1
echo "rc=$?" >> gpurun_out/r02c/tests.log
cat gpurun_out/r02c/ubench_issue.txt; tail -5 gpurun_out/r02c/tests.log
