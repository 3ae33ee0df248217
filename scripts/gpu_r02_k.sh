#!/bin/bash
mkdir -p gpurun_out/r02k
timeout 1200 python -m pytest tests/test_ncvx_gpu.py tests/test_img_gpu.py tests/test_img_onchip_gpu.py tests/test_chunked_stats_gpu.py tests/test_algs_gpu.py tests/test_stats_gpu.py -q -x > gpurun_out/r02k/tests.log 2>&1
echo "rc=$?" >> gpurun_out/r02k/tests.log
timeout 900 python -m pytest tests/test_pinned_gpu.py -q > gpurun_out/r02k/pinned.log 2>&1
echo "rc=$?" >> gpurun_out/r02k/pinned.log
tail -n 15 gpurun_out/r02k/tests.log; tail -n 5 gpurun_out/r02k/pinned.log
