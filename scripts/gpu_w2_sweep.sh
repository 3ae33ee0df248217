set -x
O=gpurun_out/w2; mkdir -p $O; rm -f $O/sweep.log
timeout 300 python scripts/prof_w2.py 5000 5000 >> $O/sweep.log 2>&1
timeout 300 python scripts/prof_w2.py 10000 10000 >> $O/sweep.log 2>&1
timeout 300 python scripts/prof_w2.py 5000 1000 >> $O/sweep.log 2>&1
timeout 600 python -m pytest tests/test_w2_gpu.py -x -q > $O/pytest2.log 2>&1; tail -n 3 $O/pytest2.log
cat $O/sweep.log
