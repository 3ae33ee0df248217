set -x
O=gpurun_out/ulpda_prof; mkdir -p $O
CMD="python scripts/prof_image.py ulpda 256 2 band 0"
$CMD > $O/plain.log 2>&1 &&
timeout 1200 ncu --set full --clock-control none --import-source on -k regex:img_ulpda_band -s 2 -c 1 -o $O/prof_ulpda_band $CMD > $O/ncu.log 2>&1
cat $O/plain.log
