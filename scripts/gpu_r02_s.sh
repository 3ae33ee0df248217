#!/bin/bash
# MYULA band L2-prefetch A/B (variant _lmc_b200_nopf.so: H5 MYULA band without the prefetch) + band parity tests
O=gpurun_out/r02s; mkdir -p $O
cat > /tmp/ab.py <<'PY'
import torch, sys
from lmc_atomi_b200 import deconv_bench as B
dev = torch.device("cuda", 0)
for rep in range(2):
    r = B.run_myula(dev, steps=2, warmup=1, chains=256, dims=(512, 512), iters=50, engine="band")
    print(sys.argv[1], "myula512", round(r["value"]), r["tv_sweeps_per_iter"], flush=True)
PY
for v in nopf pf nopf pf; do
  if [ $v = nopf ]; then export LMC_B200_LIB=$PWD/lmc_atomi_b200/_lmc_b200_nopf.so; else unset LMC_B200_LIB; fi
  PYTHONPATH=$PWD timeout 600 python /tmp/ab.py $v >> $O/ab.log 2>&1
done
unset LMC_B200_LIB
grep myula512 $O/ab.log || tail $O/ab.log
timeout 1500 python -m pytest tests -m gpu -q -k "img or band or ncvx or chunked or bench_shapes" > $O/tests.log 2>&1; echo "rc=$?" >> $O/tests.log
tail -n 4 $O/tests.log
