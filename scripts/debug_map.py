"""GPU MAP (primaldual.AdaptivePrimalDual) vs the oracle on a small image: first diverging iteration + LSQR counts."""
import sys
import numpy as np
sys.path.insert(0, ".")
from oracle import deconv_ref as D
from lmc_atomi_b200 import compat, primaldual

dims, N = (64, 80), int(sys.argv[1]) if len(sys.argv) > 1 else 300
pb = D.build_problem(D.synthetic_image(*dims))
l2o = pb["l2"]()
xs_o = []
D.AdaptivePrimalDual(l2o, pb["l1iso"](), pb["Gop"], pb["x0"], pb["tau0"], pb["mu0"], niter=N,
                     callback=lambda x: xs_o.append(x.copy()))
H = compat.Convolve2D(dims, h=np.ones((5, 5)) / 25.0, offset=(2, 2))
l2 = compat.L2(Op=H, b=pb["y"].ravel(), sigma=1 / 0.75 ** 2, niter=50, warm=True)
xs_g, trace = [], []
primaldual.AdaptivePrimalDual(l2, compat.L21(ndim=2, sigma=0.3), compat.Gradient(dims), pb["x0"], pb["tau0"],
                              pb["mu0"], niter=N, callback=lambda x: xs_g.append(x.copy()), trace=trace)
its_g = np.diff([0] + [t["lsqr_iterations"] for t in trace])
its_o = np.array(l2o.lsqr_itn)
err = np.array([np.abs(a - b).max() / max(1, np.abs(b).max()) for a, b in zip(xs_g, xs_o)])
bad = np.nonzero(err > 1e-9)[0]
print("max err", err.max(), "first bad", bad[:1], "lsqr its equal:", np.array_equal(its_g, its_o))
k = int(bad[0]) if len(bad) else 0
lo = max(0, k - 3)
for i in range(lo, min(N, lo + 12)):
    print(i, "err %.2e" % err[i], "lsqr gpu/oracle", its_g[i], its_o[i], "p %.6e d %.6e tau %.6e" % (trace[i]["p"], trace[i]["d"], trace[i]["tau"]))
print("mismatching lsqr iterations at", np.nonzero(its_g != its_o)[0][:20])
