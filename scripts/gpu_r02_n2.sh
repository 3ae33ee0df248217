#!/bin/bash
O=gpurun_out/r02n2; mkdir -p $O
timeout 1200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29531 bench.py --gpus 2 > $O/bench_n2.json 2> $O/bench_n2.err
echo "bench rc=$?" >> $O/bench_n2.err
timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29532 bench.py --impl reference --gpus 2 --steps 1 --warmup 0 > $O/bench_reference_n2.json 2>> $O/bench_n2.err
tail -n 3 $O/bench_n2.err; python - <<'PY'
import json
d=json.loads([l for l in open("gpurun_out/r02n2/bench_n2.json").read().splitlines() if l.startswith("{")][-1])
m=d["myula_deconv"]
print(d["n_gpus"], d["value"], d["e2e"]["value"], m["value"], m["e2e"]["value"], m["band_kernel_512"]["value"], m["ulpda"]["value"], m["ulpda"]["e2e"]["value"])
PY
