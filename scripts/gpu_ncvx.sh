set -x
O=gpurun_out/ncvx; mkdir -p $O
timeout 900 python -m pytest tests/test_ncvx_gpu.py tests/test_stats_gpu.py -x -q > $O/tests.log 2>&1; echo "rc=$?" >> $O/tests.log
tail -n 40 $O/tests.log
