set -x
O=gpurun_out/ulpda_b; mkdir -p $O
timeout 900 python -m pytest tests/test_img_gpu.py tests/test_algs_gpu.py -x -q > $O/tests.log 2>&1; echo "rc=$?" >> $O/tests.log
timeout 300 python scripts/prof_image.py ulpda 256 2 band 0 > $O/bench_pairs.log 2>&1
tail -n 3 $O/tests.log; cat $O/bench_pairs.log
