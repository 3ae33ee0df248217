#!/bin/bash
# ULPDA band after the translation-unit split (+ experiments): ULPDA parity tests + rate
O=gpurun_out/r02z; mkdir -p $O
timeout 1500 python -m pytest tests -m gpu -q -x -k "ulpda or ncvx or chunked or bench_shapes or img_gpu" > $O/tests.log 2>&1; echo "rc=$?" >> $O/tests.log
tail -n 4 $O/tests.log
cat > /tmp/ab.py <<'PY'
import torch, sys
from lmc_atomi_b200 import deconv_bench as B
dev = torch.device("cuda", 0)
for rep in range(3):
    r = B.run_ulpda(dev, steps=2, warmup=1, iters=50)
    print("ulpda512", round(r["value"]), flush=True)
PY
PYTHONPATH=$PWD timeout 900 python /tmp/ab.py > $O/ab.log 2>&1
grep 512 $O/ab.log || tail $O/ab.log
