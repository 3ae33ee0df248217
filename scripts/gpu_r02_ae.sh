#!/bin/bash
# on-chip MYULA with 4 row strips per CTA (1024 threads, 64 registers): parity + rate against the shipped 2-strip kernel
O=gpurun_out/r02ae; mkdir -p $O
LMC_B200_LIB=$PWD/lmc_atomi_b200/_lmc_b200_s4.so timeout 900 python -m pytest tests -m gpu -q -x -k "onchip" > $O/tests_s4.log 2>&1; echo "rc=$?" >> $O/tests_s4.log
tail -n 4 $O/tests_s4.log
cat > /tmp/ab.py <<'PY'
import torch, sys
from lmc_atomi_b200 import deconv_bench as B
dev = torch.device("cuda", 0)
for rep in range(2):
    r = B.run_myula(dev, steps=2, warmup=1)
    print(sys.argv[1], "myula256", round(r["value"]), r["engine"], r["tv_sweeps_per_iter"], flush=True)
PY
for v in base s4 base s4; do
  if [ $v = base ]; then unset LMC_B200_LIB; else export LMC_B200_LIB=$PWD/lmc_atomi_b200/_lmc_b200_$v.so; fi
  PYTHONPATH=$PWD timeout 600 python /tmp/ab.py $v >> $O/ab.log 2>&1
done
grep myula256 $O/ab.log || tail $O/ab.log
