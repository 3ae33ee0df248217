#!/bin/bash
# r02r: fresh ncu --set full captures (with source) of the two band kernels at the bench shape 512x512
O=gpurun_out/r02r; mkdir -p $O
for k in ulpda myula; do
  if [ $k = ulpda ]; then CMD="python scripts/prof_image.py ulpda 296 2 band 0"; else CMD="python scripts/prof_image.py myula 296 6 band 512"; fi
  timeout 300 $CMD > $O/${k}_plain.log 2>&1 || { tail $O/${k}_plain.log; continue; }
  timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 60 --csv --log-file $O/${k}_band_launches.csv $CMD > $O/ncu_l_$k.log 2>&1
  timeout 1500 ncu --set full --clock-control none --import-source on -k regex:img_${k}_band -s 2 -c 1 -o /tmp/prof_${k}_band $CMD > $O/ncu_f_$k.log 2>&1
  python scripts/ncu_summary.py /tmp/prof_${k}_band.ncu-rep $O/${k}_band_ncu_full_summary.csv
  ncu -i /tmp/prof_${k}_band.ncu-rep --page source --csv 2>/dev/null | gzip > $O/${k}_band_source.csv.gz
  ls -la /tmp/prof_${k}_band.ncu-rep $O/
done
