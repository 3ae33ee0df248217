#!/bin/bash
# register/shuffle TV sweep of the MYULA band kernel: parity + rate
O=gpurun_out/r02al; mkdir -p $O
timeout 1500 python -m pytest tests -m gpu -q -x -k "img or band or ncvx or chunked or bench_shapes or pinned or algs" > $O/tests.log 2>&1; echo "rc=$?" >> $O/tests.log
tail -n 6 $O/tests.log
cat > /tmp/ab.py <<'PY'
import torch, sys
from lmc_atomi_b200 import deconv_bench as B, imaging
dev = torch.device("cuda", 0)
for rep in range(2):
    r = B.run_myula(dev, steps=2, warmup=1, chains=256, dims=(512, 512), iters=50, engine="band")
    print("myula512", round(r["value"]), r["tv_sweeps_per_iter"], flush=True)
mc = imaging.MoreauTerm("mc", 0.3, 15.0, True)
r = B.run_myula(dev, steps=1, warmup=1, chains=256, dims=(512, 512), iters=20, engine="band", ncvx=mc)
print("myula512 mc", round(r["value"]))
PY
PYTHONPATH=$PWD timeout 900 python /tmp/ab.py > $O/ab.log 2>&1
grep 512 $O/ab.log || tail $O/ab.log
