#!/bin/bash
# r02g: A/B of mixture-kernel build variants (scripts/build_mix_variants.sh) + full GPU suite on the default build
mkdir -p gpurun_out/r02g
for v in v0 v1 v2 v3 v4; do
  LMC_B200_LIB=$PWD/build_tmp/lmc_$v.so timeout 300 python bench.py --no-image --no-cpu --no-f64 --steps 2 --warmup 3 --K 20000 > gpurun_out/r02g/bench_$v.json 2> gpurun_out/r02g/bench_$v.err
done
LMC_B200_LIB=$PWD/build_tmp/lmc_v2.so timeout 600 python -m pytest tests/test_mix_production_gpu.py tests/test_mix_gpu.py -q -k "laplace and (ula or mala or pula or mla) and not ihpula" > gpurun_out/r02g/tests_v2.log 2>&1
echo "rc=$?" >> gpurun_out/r02g/tests_v2.log
timeout 1500 python -m pytest tests -m gpu -q > gpurun_out/r02g/tests_all.log 2>&1
echo "rc=$?" >> gpurun_out/r02g/tests_all.log
python - <<'PY' | tee gpurun_out/r02g/variants.txt
import json
for n in ("v0","v1","v2","v3","v4"):
    try:
        d=json.load(open(f"gpurun_out/r02g/bench_{n}.json"))
        print(n, round(d["value"]/1e9,1), {a:round(v["chain_steps_per_sec_per_gpu"]/1e9,1) for a,v in d["roofline"]["per_sampler"].items()})
    except Exception as e: print(n,"failed",e)
PY
tail -n 6 gpurun_out/r02g/tests_v2.log gpurun_out/r02g/tests_all.log
