set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_img_gpu.py tests/test_algs_gpu.py -x -q > gpurun_out/ulpda_tests.log 2>&1; echo "rc=$?" >> gpurun_out/ulpda_tests.log
for g in 0 8 16 48 256; do
  timeout 300 python scripts/prof_image.py ulpda 256 2 band $g > gpurun_out/ulpda_bench_g$g.log 2>&1; echo "rc=$?" >> gpurun_out/ulpda_bench_g$g.log
done
timeout 300 python scripts/prof_image.py ulpda 256 2 tile 0 > gpurun_out/ulpda_bench_tile.log 2>&1
