set -x
mkdir -p gpurun_out
CMD="python scripts/prof_image.py ulpda 128 2 band 0"
$CMD > gpurun_out/ulpda_plain.log 2>&1 &&
timeout 1200 ncu --set full --clock-control none --import-source on -k regex:img_ulpda_band -s 2 -c 1 -o gpurun_out/prof_ulpda_band $CMD > gpurun_out/ncu_ulpda_band.log 2>&1
