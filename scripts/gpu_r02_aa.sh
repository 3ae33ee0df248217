#!/bin/bash
# ULPDA band variants: y1 forward select, y2 row sums in fp32, y3 both (H5 kernels only; main lib = baseline)
O=gpurun_out/r02aa; mkdir -p $O
cat > /tmp/ab.py <<'PY'
import torch, sys
from lmc_atomi_b200 import deconv_bench as B
dev = torch.device("cuda", 0)
for rep in range(2):
    r = B.run_ulpda(dev, steps=2, warmup=1, iters=50)
    print(sys.argv[1], "ulpda512", round(r["value"]), flush=True)
PY
for v in base y1 y2 y3 base y1 y2 y3; do
  if [ $v = base ]; then unset LMC_B200_LIB; else export LMC_B200_LIB=$PWD/lmc_atomi_b200/_lmc_b200_$v.so; fi
  PYTHONPATH=$PWD timeout 600 python /tmp/ab.py $v >> $O/ab.log 2>&1
done
grep ulpda512 $O/ab.log || tail $O/ab.log
