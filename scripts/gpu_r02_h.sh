#!/bin/bash
mkdir -p gpurun_out/r02h
for rep in 1 2; do
for v in v0 v3 v5 v6 v2; do
  LMC_B200_LIB=$PWD/build_tmp/lmc_$v.so timeout 300 python bench.py --no-image --no-cpu --no-f64 --steps 3 --warmup 3 > gpurun_out/r02h/bench_${v}_$rep.json 2> gpurun_out/r02h/bench_${v}_$rep.err
done
done
python - <<'PY' | tee gpurun_out/r02h/variants.txt
import json
for rep in (1,2):
  for n in ("v0","v3","v5","v6","v2"):
    try:
        d=json.load(open(f"gpurun_out/r02h/bench_{n}_{rep}.json"))
        print(rep, n, round(d["value"]/1e9,1), {a:round(v["chain_steps_per_sec_per_gpu"]/1e9,1) for a,v in d["roofline"]["per_sampler"].items()})
    except Exception as e: print(n,"failed",e)
PY
