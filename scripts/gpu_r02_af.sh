#!/bin/bash
# H5 / H6 / H7 at the benchmark shapes on the round's final band kernels
O=gpurun_out/r02af; mkdir -p $O
python - <<'PY' > $O/h567.log 2>&1
import json, torch
from lmc_atomi_b200 import deconv_bench as B
dev = torch.device("cuda", 0)
orig = B.make_engine
for k in (5, 6, 7):
    B.make_engine = lambda dims, dev, dtype="float32", kernel=5, seed=0, k=k: orig(dims, dev, dtype, kernel=k, seed=seed)
    u = B.run_ulpda(dev, steps=2, warmup=1, iters=50)
    m = B.run_myula(dev, steps=2, warmup=1, chains=256, dims=(512, 512), iters=50, engine="band")
    print(json.dumps({"k": k, "ulpda": round(u["value"]), "ulpda_frac": round(u["roofline"]["frac"], 3), "lsqr": u["lsqr_iters_per_solve"],
                      "myula512": round(m["value"]), "myula_frac": round(m["roofline"]["frac"], 3), "tv": m["tv_sweeps_per_iter"]}), flush=True)
PY
cat $O/h567.log
