set -x
O=gpurun_out/myula_band; mkdir -p $O
CMD="python scripts/prof_image.py myula 256 10 band 512"
$CMD > $O/plain.log 2>&1 &&
timeout 1200 ncu --set full --clock-control none --import-source on -k regex:img_myula_band -s 2 -c 1 -o $O/prof_myula_band $CMD > $O/ncu.log 2>&1
cat $O/plain.log
