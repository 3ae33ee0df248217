#!/bin/bash
# ULPDA band variants: x1 RB8/3CTA, x2 RB8/2CTA, x3 baseline (bench-only build), x4 PFD2, x5 PFD2/RB8/3CTA
O=gpurun_out/r02u; mkdir -p $O
cat > /tmp/ab.py <<'PY'
import torch, sys
from lmc_atomi_b200 import deconv_bench as B
dev = torch.device("cuda", 0)
for rep in range(2):
    r = B.run_ulpda(dev, steps=2, warmup=1, iters=50)
    print(sys.argv[1], "ulpda512", round(r["value"]), flush=True)
PY
for v in x3 x1 x2 x4 x5 x3 x4; do
  LMC_B200_LIB=$PWD/lmc_atomi_b200/_lmc_b200_$v.so PYTHONPATH=$PWD timeout 600 python /tmp/ab.py $v >> $O/ab.log 2>&1
done
timeout 60 ./scripts/cluster_probe > $O/cluster_probe.txt 2>&1; cat $O/cluster_probe.txt
grep ulpda512 $O/ab.log || tail $O/ab.log
