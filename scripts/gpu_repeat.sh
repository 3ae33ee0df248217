# flakiness check: the image-kernel parity tests five times over (work-queue / barrier races would show as intermittent failures)
O=gpurun_out/repeat; mkdir -p $O
for r in 1 2 3 4 5; do
  timeout 600 python -m pytest tests/test_img_gpu.py tests/test_img_onchip_gpu.py tests/test_ncvx_gpu.py -x -q -p no:cacheprovider 2>&1 | tail -n 1
done | tee $O/repeat.log
