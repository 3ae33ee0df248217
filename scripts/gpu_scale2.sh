set -x
O=gpurun_out/r01g_scale; mkdir -p $O
N=${1:-2}
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 3 --warmup 3 > $O/bench_n$N.log 2> $O/bench_n$N.err; echo "rc=$?" >> $O/bench_n$N.err
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 bench.py --impl reference --gpus $N --steps 1 --warmup 0 > $O/ref_n$N.log 2> $O/ref_n$N.err; echo "rc=$?" >> $O/ref_n$N.err
for f in $O/*.err; do tail -n 3 $f; done
