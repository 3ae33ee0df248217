O=gpurun_out/dropin; mkdir -p $O
python - <<'PY' > $O/dropin.log 2>&1
import time, json, numpy as np, torch
from lmc_atomi_b200 import algs, compat, deconv_bench as B
dims = (512, 512); P = dims[0] * dims[1]; sigma = 0.75
img = B.synthetic_image(*dims)
H = compat.Convolve2D(dims, h=np.ones((5, 5)) / 25, offset=(2, 2))
y = H * img.ravel() + np.random.default_rng(0).normal(0, sigma, P)
G = compat.Gradient(dims)
for dtype in ("float64", "float32"):
    for philox in (False, True):
        l2 = compat.L2(Op=H, b=y, sigma=1 / sigma ** 2, niter=50, warm=True)
        for name, fn in (("ulpda", lambda n: algs.UnadjustedLangevinPrimalDual(l2, compat.L21(ndim=2, sigma=0.3), G, np.zeros(P), tau=0.95 * sigma ** 2, mu=1., theta=1., niter=n, gfirst=False, seed=0, dtype=dtype, philox=philox)),
                         ("myula", lambda n: algs.MoreauYosidaUnadjustedLangevin(l2, compat.TV(dims, sigma=0.3, niter=10), np.zeros(P), tau=0.2 * sigma ** 2, gamma=sigma ** 2, niter=n, seed=0, dtype=dtype, philox=philox))):
            fn(5); torch.cuda.synchronize()
            t0 = time.perf_counter(); s = fn(100); torch.cuda.synchronize(); dt = time.perf_counter() - t0
            print(json.dumps({"alg": name, "dtype": dtype, "philox": philox, "iters_per_s": 100 / dt, "shape": list(s.shape)}))
PY
cat $O/dropin.log
