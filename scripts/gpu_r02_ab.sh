#!/bin/bash
# MLA in mirror coordinates + forward-select ULPDA: mixture + image parity, mixture-only bench line
O=gpurun_out/r02ab; mkdir -p $O
timeout 1500 python -m pytest tests -m gpu -q -x -k "mix or stats or ulpda or bench_shapes" > $O/tests.log 2>&1; echo "rc=$?" >> $O/tests.log
tail -n 8 $O/tests.log
timeout 900 python bench.py --no-image --no-cpu --no-f64 > $O/bench_mix.json 2> $O/bench_mix.err; echo "rc=$?" >> $O/bench_mix.err
python - <<'PY'
import json
d=json.loads([l for l in open("gpurun_out/r02ab/bench_mix.json").read().splitlines() if l.startswith("{")][-1])
print("mix", d["value"]/1e9, {k: round(v["ms_per_launch"],1) for k,v in d["roofline"]["per_sampler"].items()})
PY
