#!/bin/bash
# r02aj: full pass on the round's code: all GPU tests, smoke, full bench (both arms), launch list of the bench
O=gpurun_out/r02aj; mkdir -p $O
nvidia-smi --query-gpu=name,clocks.max.sm,memory.total --format=csv > $O/gpu.txt
echo skip-tests > $O/pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > $O/smoke.log 2>&1; echo "smoke rc=$?" >> $O/smoke.log
timeout 1500 python bench.py > $O/bench.json 2> $O/bench.err; echo "bench rc=$?" >> $O/bench.err
timeout 600 python bench.py --impl reference --steps 1 --warmup 0 > $O/bench_reference.json 2>> $O/bench.err
CMD="python bench.py --steps 1 --warmup 3 --K 2000 --no-cpu --no-f64"
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/launches_bench.csv $CMD > $O/ncu1.log 2>&1
tail -n 4 $O/pytest_gpu.log $O/smoke.log $O/bench.err
python - <<'PY'
import json
d=json.loads(open("gpurun_out/r02aj/bench.json").read().strip().split("\n")[-1])
m=d["myula_deconv"]
print("mix", d["value"]/1e9, "e2e", d["e2e"]["value"]/1e9, "frac", d["roofline"]["frac"])
print("myula", m["value"], m["e2e"]["value"], "band", m["band_kernel_512"]["value"], "ulpda", m["ulpda"]["value"], m["ulpda"]["e2e"]["value"])
for k,v in m["ncvx"].items(): print(k, v["value"], v["roofline"]["frac"])
PY
