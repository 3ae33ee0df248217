#!/bin/bash
# ULPDA band L2-prefetch A/B (variant library _lmc_b200_pf.so built with -DLMC_BAND_L2_PREFETCH=1)
O=gpurun_out/r02p; mkdir -p $O
cat > /tmp/ab.py <<'PY'
import torch, sys
from lmc_atomi_b200 import deconv_bench as B
dev = torch.device("cuda", 0)
for rep in range(2):
    r = B.run_ulpda(dev, steps=2, warmup=1)
    print(sys.argv[1], "ulpda512", r["value"], r["roofline"].get("frac"), r.get("lsqr_iters_per_iter"), flush=True)
PY
for v in base pf base pf; do
  if [ $v = pf ]; then export LMC_B200_LIB=$PWD/lmc_atomi_b200/_lmc_b200_pf.so; else unset LMC_B200_LIB; fi
  PYTHONPATH=$PWD timeout 600 python /tmp/ab.py $v >> $O/ab.log 2>&1
done
grep ulpda512 $O/ab.log || tail $O/ab.log
