#!/bin/bash
# r02a: pinned-oracle run (einstein, 9 models x 2 samplers, N=1000) + full GPU suite after the TV / stat fixes
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm --format=csv > gpurun_out/r02a_smi.txt
timeout 1500 python -m pytest tests/test_pinned_gpu.py -x -q -s > gpurun_out/r02a_pinned.log 2>&1
echo "pinned rc=$?" >> gpurun_out/r02a_pinned.log
timeout 1200 python -m pytest tests -m gpu -q --deselect tests/test_pinned_gpu.py > gpurun_out/r02a_tests.log 2>&1
echo "tests rc=$?" >> gpurun_out/r02a_tests.log
tail -5 gpurun_out/r02a_pinned.log gpurun_out/r02a_tests.log
