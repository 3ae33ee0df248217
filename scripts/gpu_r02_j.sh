#!/bin/bash
# r02j (2 GPUs): mixture tests after the TU split / NOSHIFT fix, then the N=2 bench line (both arms)
mkdir -p gpurun_out/r02j
timeout 900 python -m pytest tests/test_mix_production_gpu.py tests/test_mix_gpu.py tests/test_stats_gpu.py tests/test_parallel_cpu.py -q > gpurun_out/r02j/mix_tests.log 2>&1
echo "rc=$?" >> gpurun_out/r02j/mix_tests.log
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 2 --warmup 3 > gpurun_out/r02j/bench_n2.json 2> gpurun_out/r02j/bench_n2.err
echo "bench rc=$?" >> gpurun_out/r02j/bench_n2.err
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --impl reference --gpus 2 --steps 1 --warmup 0 > gpurun_out/r02j/bench_reference_n2.json 2>> gpurun_out/r02j/bench_n2.err
tail -n 4 gpurun_out/r02j/mix_tests.log; tail -n 6 gpurun_out/r02j/bench_n2.err; python - <<'PY'
import json
d=json.load(open("gpurun_out/r02j/bench_n2.json"))
print(d["n_gpus"], d["value"], d["e2e"]["value"], d["myula_deconv"]["value"], d["myula_deconv"]["e2e"]["value"], d["myula_deconv"]["ulpda"]["value"], d["myula_deconv"]["ulpda"]["e2e"]["value"])
PY
