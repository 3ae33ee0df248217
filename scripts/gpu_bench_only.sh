set -x
O=gpurun_out/r01f; mkdir -p $O
timeout 900 python -m pytest tests/test_mix_gpu.py tests/test_mix_production_gpu.py -x -q > $O/mix_tests.log 2>&1; echo "rc=$?" >> $O/mix_tests.log
timeout 900 python bench.py > $O/bench_full2.log 2> $O/bench_full2.err; echo "bench rc=$?" >> $O/bench_full2.err
tail -n 2 $O/mix_tests.log
