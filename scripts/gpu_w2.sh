set -x
O=gpurun_out/w2; mkdir -p $O
timeout 900 python -m pytest tests/test_w2_gpu.py -x -q -s > $O/pytest.log 2>&1; echo "rc=$?" >> $O/pytest.log
tail -n 30 $O/pytest.log
