set -x
mkdir -p gpurun_out
CMD="python scripts/prof_image.py myula 256 10"
$CMD > gpurun_out/onchip_plain.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:img_myula_onchip -s 2 -c 1 -o gpurun_out/prof_onchip3 $CMD > gpurun_out/ncu_onchip3.log 2>&1
