#!/bin/bash
O=gpurun_out/r02ag; mkdir -p $O
timeout 1200 python -m pytest tests/test_algs_gpu.py tests/test_pinned_gpu.py tests/test_ncvx_gpu.py -q -x > $O/tests.log 2>&1; echo "rc=$?" >> $O/tests.log
tail -n 12 $O/tests.log
