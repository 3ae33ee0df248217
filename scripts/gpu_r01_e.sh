# round-1 fifth GPU pass: all GPU tests, smoke, full bench (both arms), launch lists + full capture of the ULPDA band kernel
set -x
mkdir -p gpurun_out/r01e
O=gpurun_out/r01e
nvidia-smi --query-gpu=name,clocks.max.sm,memory.total --format=csv > $O/gpu.txt
timeout 1500 python -m pytest tests -m gpu -x -q > $O/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> $O/pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > $O/smoke.log 2>&1; echo "smoke rc=$?" >> $O/smoke.log
timeout 900 python bench.py > $O/bench_full.log 2> $O/bench_full.err; echo "bench rc=$?" >> $O/bench_full.err
timeout 600 python bench.py --impl reference --steps 1 --warmup 0 > $O/bench_ref.log 2>&1
CMD="python scripts/prof_image.py ulpda 256 2 band 0"
timeout 300 $CMD > $O/ulpda_plain.log 2>&1 &&
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 60 --csv --log-file $O/launches_ulpda_band.csv $CMD > $O/ncu_l_ulpda.log 2>&1
timeout 1200 ncu --set full --clock-control none --import-source on -k regex:img_ulpda_band -s 2 -c 1 -o $O/prof_ulpda_band $CMD > $O/ncu_ulpda_band.log 2>&1
ls -la $O
