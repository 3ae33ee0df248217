set -x
O=gpurun_out/myula_band; mkdir -p $O
timeout 900 python -m pytest tests/test_img_gpu.py tests/test_algs_gpu.py tests/test_ncvx_gpu.py tests/test_stats_gpu.py -x -q > $O/tests.log 2>&1; echo "rc=$?" >> $O/tests.log
timeout 300 python scripts/prof_image.py myula 256 10 band 512 > $O/bench_512_band.log 2>&1
tail -n 4 $O/tests.log; cat $O/bench_512_band.log
