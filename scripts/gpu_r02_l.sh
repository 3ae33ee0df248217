#!/bin/bash
mkdir -p gpurun_out/r02l
timeout 1200 python -m pytest tests/test_ncvx_gpu.py tests/test_img_gpu.py tests/test_chunked_stats_gpu.py tests/test_algs_gpu.py tests/test_stats_gpu.py tests/test_bench_shapes_gpu.py -q > gpurun_out/r02l/tests.log 2>&1
echo "rc=$?" >> gpurun_out/r02l/tests.log
timeout 900 python -m pytest tests/test_pinned_gpu.py -q > gpurun_out/r02l/pinned.log 2>&1
echo "rc=$?" >> gpurun_out/r02l/pinned.log
tail -n 25 gpurun_out/r02l/tests.log | cut -c1-250; tail -n 5 gpurun_out/r02l/pinned.log
