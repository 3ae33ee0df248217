#!/bin/bash
# ULPDA band: chain-group sweep (L2-resident LSQR working set?) with and without the L2 prefetch variant
O=gpurun_out/r02q; mkdir -p $O
cat > /tmp/grp.py <<'PY'
import torch, sys
from lmc_atomi_b200 import deconv_bench as B
dev = torch.device("cuda", 0)
for g in (0, 148, 74, 37, 24, 16):
    r = B.run_ulpda(dev, steps=1, warmup=1, iters=40, group=g)
    print(sys.argv[1], "group", g, "value", round(r["value"]), flush=True)
PY
PYTHONPATH=$PWD timeout 900 python /tmp/grp.py base >> $O/grp.log 2>&1
LMC_B200_LIB=$PWD/lmc_atomi_b200/_lmc_b200_pf.so PYTHONPATH=$PWD timeout 900 python /tmp/grp.py pf >> $O/grp.log 2>&1
grep group $O/grp.log || tail $O/grp.log
