#!/bin/bash
# r02 (8 GPUs): the N=8 bench line (default steps, as the driver runs it) and the reference arm
O=gpurun_out/r02n8b; mkdir -p $O
nvidia-smi -L > $O/gpus.txt
timeout 1200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29521 bench.py --gpus 8 > $O/bench_n8.json 2> $O/bench_n8.err
echo "bench rc=$?" >> $O/bench_n8.err
timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29522 bench.py --impl reference --gpus 8 --steps 1 --warmup 0 > $O/bench_reference_n8.json 2>> $O/bench_n8.err
tail -n 5 $O/bench_n8.err; python - <<'PY'
import json
d=json.loads([l for l in open("gpurun_out/r02n8b/bench_n8.json").read().splitlines() if l.startswith("{")][-1])
m=d["myula_deconv"]
print(d["n_gpus"], d["value"], d["e2e"]["value"], m["value"], m["e2e"]["value"], m["band_kernel_512"]["value"], m["ulpda"]["value"], m["ulpda"]["e2e"]["value"])
PY
