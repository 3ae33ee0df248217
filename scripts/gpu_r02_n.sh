#!/bin/bash
O=gpurun_out/r02n; mkdir -p $O
timeout 1200 python -m pytest tests/test_img_gpu.py tests/test_ncvx_gpu.py tests/test_bench_shapes_gpu.py tests/test_chunked_stats_gpu.py tests/test_mix_production_gpu.py -q > $O/tests.log 2>&1; echo "rc=$?" >> $O/tests.log
timeout 600 python - > $O/band.log 2>&1 <<'PY'
import torch, json
from lmc_atomi_b200 import deconv_bench as B, imaging
dev = torch.device("cuda", 0)
r = B.run_myula(dev, steps=2, warmup=1, chains=256, dims=(512, 512), iters=50, engine="band")
print("band512", r["value"], r["tv_sweeps_per_iter"], r["roofline"]["frac"])
mc = imaging.MoreauTerm("mc", 0.3, 15.0, True)
r = B.run_myula(dev, steps=1, warmup=1, chains=256, dims=(512, 512), iters=20, engine="band", ncvx=mc)
print("band512 mc", r["value"])
PY
tail -n 4 $O/tests.log; cat $O/band.log | tail -4
