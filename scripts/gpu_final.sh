# last check of the round: all GPU tests, smoke, the default bench of both arms
set -x
O=gpurun_out/r01h; mkdir -p $O
timeout 1500 python -m pytest tests -m gpu -x -q > $O/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> $O/pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > $O/smoke.log 2>&1; echo "smoke rc=$?" >> $O/smoke.log
timeout 900 python bench.py > $O/bench_full.log 2> $O/bench_full.err; echo "bench rc=$?" >> $O/bench_full.err
timeout 600 python bench.py --impl reference --steps 1 --warmup 0 > $O/bench_ref.log 2>&1; echo "ref rc=$?" >> $O/bench_ref.log
for f in pytest_gpu.log smoke.log bench_full.err bench_ref.log; do tail -n 2 $O/$f | cut -c1-300; done
