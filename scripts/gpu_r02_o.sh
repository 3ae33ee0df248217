#!/bin/bash
# no-alloc entry point test (alone, fresh process) + ULPDA band L2-prefetch A/B
O=gpurun_out/r02o; mkdir -p $O
timeout 2400 python -m pytest tests -m gpu -x -q > $O/all_tests.log 2>&1; echo "rc=$?" >> $O/all_tests.log
tail -n 5 $O/all_tests.log
cat > /tmp/ab.py <<'PY'
import torch, sys
from lmc_atomi_b200 import deconv_bench as B
dev = torch.device("cuda", 0)
for rep in range(2):
    r = B.run_ulpda(dev, steps=2, warmup=1)
    print(sys.argv[1], "ulpda512", r["value"], r["ms_per_step"], r["roofline"]["frac"], flush=True)
PY
for v in base pf base pf; do
  if [ $v = pf ]; then export LMC_B200_LIB=$PWD/lmc_atomi_b200/_lmc_b200_pf.so; else unset LMC_B200_LIB; fi
  PYTHONPATH=$PWD timeout 600 python /tmp/ab.py $v >> $O/ab.log 2>&1
done
unset LMC_B200_LIB
cat $O/ab.log | grep ulpda512
LMC_B200_LIB=$PWD/lmc_atomi_b200/_lmc_b200_pf.so timeout 900 python -m pytest tests/test_img_gpu.py tests/test_bench_shapes_gpu.py -q -k ulpda > $O/pf_tests.log 2>&1; echo "rc=$?" >> $O/pf_tests.log
tail -n 4 $O/pf_tests.log
