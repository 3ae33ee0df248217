set -x
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_img_onchip_gpu.py tests/test_img_gpu.py -x -q -s -k "onchip or fp32" > gpurun_out/onchip_tests.log 2>&1; echo "rc=$?" >> gpurun_out/onchip_tests.log
timeout 300 python scripts/prof_image.py myula 1024 20 > gpurun_out/onchip_bench.log 2>&1; echo "rc=$?" >> gpurun_out/onchip_bench.log
CMD="python scripts/prof_image.py myula 256 10"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:img_myula_onchip -s 2 -c 1 -o gpurun_out/prof_onchip2 $CMD > gpurun_out/ncu_onchip2.log 2>&1
