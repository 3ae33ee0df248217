set -x
mkdir -p gpurun_out
for v in "" "_split4"; do
  if [ -n "$v" ]; then export LMC_B200_LIB=$PWD/lmc_atomi_b200/_lmc_b200$v.so; fi
  timeout 600 python -m pytest tests/test_img_onchip_gpu.py -x -q > gpurun_out/onchip_tests$v.log 2>&1; echo "rc=$?" >> gpurun_out/onchip_tests$v.log
  timeout 300 python scripts/prof_image.py myula 1024 20 > gpurun_out/onchip_bench$v.log 2>&1; echo "rc=$?" >> gpurun_out/onchip_bench$v.log
done
