#!/bin/bash
# r02ac: ncu capture of the final mixture kernels (mirror MLA, 3-word MALA stream, folded ULA / PULA scalars)
O=gpurun_out/r02ac; mkdir -p $O
CMD="python bench.py --steps 1 --warmup 3 --K 2000 --no-image --no-cpu --no-f64"
$CMD > $O/plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 80 --csv --log-file $O/mix_launches.csv $CMD > $O/ncu1.log 2>&1
timeout 900 ncu --set full --clock-control none -k regex:mix_chain -s 12 -c 4 -o /tmp/prof_mix $CMD > $O/ncu2.log 2>&1
python scripts/ncu_summary.py /tmp/prof_mix.ncu-rep $O/mix_ncu_full_summary.csv
tail -n 2 $O/plain.log | head -c 600; ls -la $O
