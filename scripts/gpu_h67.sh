set -x
O=gpurun_out/h67; mkdir -p $O
timeout 900 python -m pytest tests/test_img_gpu.py tests/test_algs_gpu.py tests/test_ncvx_gpu.py tests/test_stats_gpu.py -x -q > $O/tests.log 2>&1; echo "rc=$?" >> $O/tests.log
tail -n 4 $O/tests.log
python - <<'PY' > $O/bench.log 2>&1
import json, torch
from lmc_atomi_b200 import deconv_bench as B
dev = torch.device("cuda", 0)
orig = B.make_engine
for k in (5, 6, 7):
    B.make_engine = lambda dims, dev, dtype="float32", kernel=5, seed=0, k=k: orig(dims, dev, dtype, kernel=k, seed=seed)
    u = B.run_ulpda(dev, steps=2, warmup=1)
    m = B.run_myula(dev, steps=2, warmup=1, chains=256, dims=(512, 512), iters=10, engine="band")
    print(json.dumps({"k": k, "ulpda": u["value"], "ulpda_frac": u["roofline"]["frac"], "lsqr": u["lsqr_iters_per_solve"],
                      "myula512": m["value"], "myula_frac": m["roofline"]["frac"]}))
PY
cat $O/bench.log
