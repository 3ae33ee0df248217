"""Timing of the GPU W2 evaluation:  python scripts/prof_w2.py [n] [m]"""
import sys, time
import numpy as np
import torch
sys.path.insert(0, ".")
from lmc_atomi_b200 import w2

n = int(sys.argv[1]) if len(sys.argv) > 1 else 5000
m = int(sys.argv[2]) if len(sys.argv) > 2 else n
rng = np.random.default_rng(1)
x = rng.normal(size=(n, 2)) * 1.5
y = rng.normal(size=(m, 2)) * 1.35 + 0.2
ev = w2.SinkhornW2(x)
import warnings
with warnings.catch_warnings():
    warnings.simplefilter('ignore')
    w2.SinkhornW2(x[:64], max_iters=1).distance(y[:64])       # loads the library; ~40 tiny launches
torch.cuda.synchronize()
t0 = time.time()
d = ev.distance(y)
dt = time.time() - t0
it = ev.last["iters"]
print("W2 %dx%d = %.6f  %.3f s  %d iterations  %.1f us/iteration  %.3g pairs/s" % (n, m, d, dt, it, dt / it * 1e6, 2.0 * n * m * it / dt))
if len(sys.argv) > 3:
    t0 = time.time()
    c = ev.curve(y, interval=100)
    print("curve of %d prefixes: %.2f s, last %.6f" % (len(c), time.time() - t0, c[-1]))
    print("iterations per prefix:", [h["iters"] for h in ev.history[-len(c):]], "all converged:", all(h["converged"] for h in ev.history))
    cold = w2.SinkhornW2(x).distance(y[:4902])
    print("cold value of the last prefix: %.6f" % cold)
