set -x
mkdir -p gpurun_out
python -c "
import sys; sys.path.insert(0,'.')
from lmc_atomi_b200 import _lib
print('clusters', _lib.load().lmc_img_myula_onchip_clusters())
" > gpurun_out/onchip_clusters.log 2>&1
CMD="python scripts/prof_image.py myula 256 10"
$CMD > gpurun_out/onchip_plain.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:img_myula_onchip -s 2 -c 1 -o gpurun_out/prof_onchip $CMD > gpurun_out/ncu_onchip.log 2>&1
ls -la gpurun_out
