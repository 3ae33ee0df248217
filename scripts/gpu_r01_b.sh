# round-1 second GPU pass: all GPU tests, smoke, full bench, launch lists + full captures of the image kernels
set -x
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.max.sm,memory.total --format=csv > gpurun_out/gpu.txt
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?" >> gpurun_out/smoke.log
python bench.py > gpurun_out/bench_full.log 2> gpurun_out/bench_full.err; echo "bench rc=$?" >> gpurun_out/bench_full.err
python bench.py --impl reference --steps 1 --warmup 0 > gpurun_out/bench_ref.log 2>&1
for w in myula ulpda; do
  CMD="python scripts/prof_image.py $w 256 2"
  $CMD > gpurun_out/prof_${w}_plain.log 2>&1 &&
  timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 60 --csv --log-file gpurun_out/launches_$w.csv $CMD > gpurun_out/ncu_l_$w.log 2>&1
  timeout 900 ncu --set full --clock-control none --import-source on -k regex:img_${w}_kernel -s 3 -c 1 -o gpurun_out/prof_$w $CMD > gpurun_out/ncu_f_$w.log 2>&1
done
ls -la gpurun_out
