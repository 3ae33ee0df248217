#!/bin/bash
O=gpurun_out/r02ak; mkdir -p $O
timeout 900 python -m pytest tests/test_bench_shapes_gpu.py -q -x -k "wide" > $O/tests.log 2>&1; echo "rc=$?" >> $O/tests.log
tail -n 12 $O/tests.log
