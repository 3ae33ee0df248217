"""Small image-chain run for ncu:  python scripts/prof_image.py [myula|ulpda] [chains] [iters]"""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from lmc_atomi_b200 import deconv_bench

which = sys.argv[1] if len(sys.argv) > 1 else "myula"
chains = int(sys.argv[2]) if len(sys.argv) > 2 else 512
iters = int(sys.argv[3]) if len(sys.argv) > 3 else 4
dev = torch.device("cuda", 0)
torch.cuda.set_device(dev)
if which == "myula":
    engine = sys.argv[4] if len(sys.argv) > 4 else "auto"
    side = int(sys.argv[5]) if len(sys.argv) > 5 else 256
    r = deconv_bench.run_myula(dev, steps=2, warmup=1, chains=chains, iters=iters, engine=engine, dims=(side, side))
else:
    engine = sys.argv[4] if len(sys.argv) > 4 else "auto"
    group = int(sys.argv[5]) if len(sys.argv) > 5 else 0
    r = deconv_bench.run_ulpda(dev, steps=2, warmup=1, chains=chains, iters=iters, engine=engine, group=group)
print(json.dumps(r))
