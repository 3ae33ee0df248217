#!/bin/bash
# pipelined TMA TV sweep (MYULA band): parity + rate; ULPDA prefetch depth variants x4 (PFD 2), x6 (PFD 1)
O=gpurun_out/r02v; mkdir -p $O
timeout 1500 python -m pytest tests -m gpu -q -x -k "img or band or ncvx or chunked or bench_shapes or pinned" > $O/tests.log 2>&1; echo "rc=$?" >> $O/tests.log
tail -n 6 $O/tests.log
cat > /tmp/ab.py <<'PY'
import torch, sys
from lmc_atomi_b200 import deconv_bench as B
dev = torch.device("cuda", 0)
for rep in range(2):
    if sys.argv[1] == "main":
        r = B.run_myula(dev, steps=2, warmup=1, chains=256, dims=(512, 512), iters=50, engine="band")
        print("myula512", round(r["value"]), r["tv_sweeps_per_iter"], flush=True)
    r = B.run_ulpda(dev, steps=2, warmup=1, iters=50)
    print(sys.argv[1], "ulpda512", round(r["value"]), flush=True)
PY
PYTHONPATH=$PWD timeout 900 python /tmp/ab.py main > $O/ab.log 2>&1
for v in x4 x6 x4 x6; do
  LMC_B200_LIB=$PWD/lmc_atomi_b200/_lmc_b200_$v.so PYTHONPATH=$PWD timeout 600 python /tmp/ab.py $v >> $O/ab.log 2>&1
done
grep 512 $O/ab.log || tail $O/ab.log
