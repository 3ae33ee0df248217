set -x
O=gpurun_out/w2; mkdir -p $O
CMD="python scripts/prof_w2.py 5000 5000"
timeout 300 $CMD > $O/plain.log 2>&1 &&
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -s 400 -c 60 --csv --log-file $O/launches_w2.csv $CMD > $O/ncu_l.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:w2_half_step -s 2000 -c 1 -o $O/prof_w2 $CMD > $O/ncu_f.log 2>&1
cat $O/plain.log; tail -n 3 $O/ncu_f.log; ls -la $O
