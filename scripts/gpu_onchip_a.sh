set -x
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_img_onchip_gpu.py tests/test_img_gpu.py -x -q -s -k "onchip or fp32" > gpurun_out/onchip_tests.log 2>&1; echo "rc=$?" >> gpurun_out/onchip_tests.log
timeout 600 compute-sanitizer --tool memcheck python -m pytest tests/test_img_onchip_gpu.py -x -q -k "niter_edge" > gpurun_out/onchip_memcheck.log 2>&1; echo "rc=$?" >> gpurun_out/onchip_memcheck.log
timeout 300 python scripts/prof_image.py myula 1024 20 > gpurun_out/onchip_bench.log 2>&1; echo "rc=$?" >> gpurun_out/onchip_bench.log
