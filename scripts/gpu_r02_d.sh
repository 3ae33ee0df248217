#!/bin/bash
# r02d: packed (two chains per thread, FFMA2) mixture kernels: parity + A/B throughput; MAP branch pinned test
mkdir -p gpurun_out/r02d
timeout 1200 python -m pytest tests/test_mix_production_gpu.py tests/test_mix_gpu.py tests/test_stats_gpu.py -q -x > gpurun_out/r02d/mix_tests.log 2>&1
echo "rc=$?" >> gpurun_out/r02d/mix_tests.log
timeout 600 python bench.py --no-image --no-cpu --no-f64 --steps 2 --warmup 3 --K 20000 > gpurun_out/r02d/bench_packed.json 2> gpurun_out/r02d/bench_packed.err
LMC_MIX_SCALAR=1 timeout 600 python bench.py --no-image --no-cpu --no-f64 --steps 2 --warmup 3 --K 20000 > gpurun_out/r02d/bench_scalar.json 2> gpurun_out/r02d/bench_scalar.err
timeout 900 python -m pytest tests/test_pinned_gpu.py -q -k map > gpurun_out/r02d/pinned_map.log 2>&1
echo "rc=$?" >> gpurun_out/r02d/pinned_map.log
cp gpurun_out/pinned_MAP.json gpurun_out/r02d/ 2>/dev/null
tail -3 gpurun_out/r02d/*.log; python - <<'PY'
import json
for n in ("packed","scalar"):
    try:
        d=json.load(open(f"gpurun_out/r02d/bench_{n}.json"))
        print(n, d["value"], {a:round(v["chain_steps_per_sec_per_gpu"]/1e9,1) for a,v in d["roofline"]["per_sampler"].items()})
    except Exception as e: print(n,"failed",e)
PY
