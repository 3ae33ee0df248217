"""Small image-chain check used by ``__graft_entry__.smoke()``: MYULA and ULPDA through the
drop-in ``algs`` functions on a 48x40 problem, compared with the oracle (the only place the
package touches oracle/, and only as the checker)."""
from __future__ import annotations

import numpy as np


def run(N=6):
    from oracle import deconv_ref as D
    from . import algs, compat
    dims = (48, 40)
    pb = D.build_problem(D.synthetic_image(*dims))
    H = compat.Convolve2D(dims, h=np.ones((5, 5)) / 25.0, offset=(2, 2))
    l2 = compat.L2(Op=H, b=pb["y"].ravel(), sigma=1 / 0.75 ** 2, niter=50, warm=True)
    want = D.MoreauYosidaUnadjustedLangevin(pb["l2"](), pb["tv"](), pb["x0"], tau=pb["tau_myula"],
                                            gamma=pb["gamma_myula"], niter=N, seed=0)
    got = algs.MoreauYosidaUnadjustedLangevin(l2, compat.TV(dims, sigma=0.3, niter=10), pb["x0"],
                                              tau=pb["tau_myula"], gamma=pb["gamma_myula"], niter=N, seed=0)
    e1 = np.abs(got - want).max() / np.abs(want).max()
    assert e1 < 1e-11, f"MYULA image chain differs from the oracle: {e1}"
    want = D.UnadjustedLangevinPrimalDual(pb["l2"](), pb["l1iso"](), pb["Gop"], pb["x0"], pb["tau0"], pb["mu0"],
                                          theta=1., niter=N, seed=0, gfirst=False)
    got = algs.UnadjustedLangevinPrimalDual(l2, compat.L21(ndim=2, sigma=0.3), compat.Gradient(dims), pb["x0"],
                                            tau=pb["tau0"], mu=pb["mu0"], theta=1., niter=N, seed=0, gfirst=False)
    e2 = np.abs(got - want).max() / np.abs(want).max()
    assert e2 < 1e-10, f"ULPDA image chain differs from the oracle: {e2}"
    print(f"deconv smoke ok: MYULA rel err {e1:.2e}, ULPDA rel err {e2:.2e}")
