O=gpurun_out/dropin; mkdir -p $O
timeout 900 python -m pytest tests/test_algs_gpu.py tests/test_ncvx_gpu.py tests/test_stats_gpu.py -x -q 2>&1 | tail -n 2
bash scripts/gpu_dropin.sh
