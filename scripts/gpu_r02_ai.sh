#!/bin/bash
O=gpurun_out/r02ai; mkdir -p $O
cat > /tmp/e2e.py <<'PY'
import torch
from lmc_atomi_b200 import deconv_bench as B
dev = torch.device("cuda", 0)
for rep in range(3):
    r = B.run_ulpda(dev, steps=3, warmup=2)
    print("ulpda512", round(r["value"]), "e2e", round(r["e2e"]["value"]), flush=True)
PY
PYTHONPATH=$PWD timeout 900 python /tmp/e2e.py > $O/e2e.log 2>&1; cat $O/e2e.log | tail -4; nproc; lscpu | grep -i "model name\|numa" | head -5
