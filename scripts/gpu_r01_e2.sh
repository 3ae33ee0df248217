# r01e: band-kernel regression tests + fresh ncu captures of the CURRENT mixture kernels and the on-chip MYULA kernel
set -x
O=gpurun_out/r01e2; mkdir -p $O
timeout 900 python -m pytest tests/test_img_gpu.py -x -q > $O/tests.log 2>&1; echo "rc=$?" >> $O/tests.log
CMD="python bench.py --steps 1 --warmup 3 --K 2000 --no-image --no-cpu"
timeout 300 $CMD > $O/mix_plain.log 2>&1 &&
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 80 --csv --log-file $O/launches_mix.csv $CMD > $O/ncu_l_mix.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:mix_chain -s 12 -c 4 -o $O/prof_mix $CMD > $O/ncu_f_mix.log 2>&1
CMD="python scripts/prof_image.py myula 256 10"
timeout 300 $CMD > $O/myula_plain.log 2>&1 &&
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 60 --csv --log-file $O/launches_myula.csv $CMD > $O/ncu_l_myula.log 2>&1
tail -n 3 $O/tests.log
