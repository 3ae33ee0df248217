"""``pyproximal.optimization.primaldual.AdaptivePrimalDual`` on the CUDA operators -- the MAP branch of the reference
driver (prox_lmc_deconv.py:138-445), SURVEY 8(f).4.

pyproximal is not installed here and the reference pins no version; the iteration below is the adaptive PDHG of
Goldstein, Li, Yuan, Esser, Baraniuk ("Adaptive primal-dual hybrid gradient methods for saddle-point problems") as
pyproximal carries it, PINNED against the reference's own MAP figure
(fig/fig_prox_lmc_deconv_einstein_MAP_1000_snr_psnr_mse.pdf -> tests/golden/deconv_curves.npz, gates in
tests/test_oracle_pinned.py and tests/test_pinned_gpu.py): l2-norm residuals, alpha = 0.5, eta = 0.95, s = 1,
delta = 1.5 reproduce the per-iteration SNR / PSNR / MSE curves to the resolution of the plot; l1-norm residuals
diverge from it at the first step-size adaptation (iteration ~8).

    x     = proxf.prox(x - tau (A^T y + z), tau)
    y     = proxg.proxdual(y + mu A(2 x - x_old), mu)
    p, d  = ||(x_old - x)/tau - A^T(y_old - y)||_2,  ||(y_old - y)/mu - A(x_old - x)||_2
    p > s d delta:  tau /= (1 - alpha); mu *= (1 - alpha); alpha *= eta
    p < s d / delta: tau *= (1 - alpha); mu /= (1 - alpha); alpha *= eta

Supported operator combinations = those of ``algs.UnadjustedLangevinPrimalDual``: proxf = pyproximal.L2 or
algs.L2_ncvx_tv over a Convolve2D blur, proxg = L21 / L1, A = forward-difference Gradient.  No CPU fallback.
"""
from __future__ import annotations

import ctypes as C
import time

import numpy as np

from . import _lib, algs, imaging

__all__ = ["AdaptivePrimalDual"]


def AdaptivePrimalDual(proxf, proxg, A, x0, tau, mu, alpha=0.5, eta=0.95, s=1., delta=1.5, z=None, niter=10,
                       tol=1e-10, callback=None, show=False, dtype="float64", device_metrics=None, trace=None):
    """Returns ``(x, steps)`` like pyproximal: the final iterate (P,) float64 and the (niter, 3) array of
    (tau, mu, alpha) after each iteration.  ``device_metrics``: an ``algs.DeviceMetrics`` collecting the driver's
    callback quantities on the GPU instead of a per-iteration host ``callback``."""
    eng = algs._engine(proxf, dtype)
    torch, lib = eng.torch, eng.lib
    P, dims = eng.P, eng.dims
    ncvx = algs._moreau_term(proxf, dims)
    algs._require(algs._kind(proxg) in ("L21", "L1"), f"proxg of type {algs._kind(proxg)}; supported: L21, L1")
    iso = algs._kind(proxg) == "L21"
    algs._require(algs._is_gradient(A, dims), "A must be pylops.Gradient(dims, kind='forward', edge=False)")
    g_sigma = float(proxg.sigma)
    dt = imaging._dt(eng.dtype)
    x = eng._buf(np.asarray(x0, dtype=np.float64), 1)
    y = torch.zeros((1, 2, P), dtype=eng.tdt, device=eng.dev)
    Ax = torch.zeros((1, 2, P), dtype=eng.tdt, device=eng.dev)
    ATy = torch.zeros((1, P), dtype=eng.tdt, device=eng.dev)
    zt = torch.as_tensor(z).to(eng.dev, eng.tdt).reshape(P).contiguous() if z is not None else None
    warm_on = bool(getattr(proxf, "warm", True))
    warm = eng._buf(getattr(proxf, "x0", None) if warm_on else None, 1)
    kw = getattr(proxf, "kwargs_solver", None) or {}
    nb = lib.lmc_img_pdhg_workspace_bytes()
    ws = torch.empty(nb, dtype=torch.uint8, device=eng.dev)
    resid = torch.zeros(2, dtype=torch.float64, device=eng.dev)
    stream = imaging._stream(torch, eng.dev)
    steps = np.zeros((niter, 3))
    tau, mu, alpha = float(tau), float(mu), float(alpha)
    if show:
        tstart = time.time()
        print('Adaptive Primal-dual: min_x f(Ax) + x^T z + g(x)\n'
              '---------------------------------------------------------\n'
              'Proximal operator (f): %s\nProximal operator (g): %s\nLinear operator (A): %s\n'
              'tau0 = %10e\tmu0 = %10e\nalpha0 = %10e\teta = %10e\ns = %10e\tdelta = %10e\nniter = %d\t\ttol = %10e\n'
              % (type(proxf), type(proxg), type(A), tau, mu, alpha, eta, s, delta, niter, tol))
        print('   Itn       x[0]          f           g          z^x       J = f + g + z^x')
    iiter = 0
    count0 = int(getattr(proxf, "count", 0))
    aq = algs._linear_term(proxf, P)
    aq = None if aq is None else eng._buf(aq, 1)
    while iiter < niter:
        xold = x.clone()
        # primal step: x - tau (A^T y + z), [L2_ncvx_tv: + tau lamda grad_moreau, algs.py:208-223], LSQR solve
        arg = imaging.pd_arg(dims, x, y, tau, zt)
        if ncvx is not None:
            arg = imaging.moreau_apply(dims, arg, ncvx, tau=tau)
        if aq is not None:      # y = x + tau OpTb - tau alpha q (algs.py:226-227), after the Moreau pre-step
            arg = arg - tau * aq
        if not warm_on:
            warm.zero_()
        x = eng.l2_prox(arg, tau, warm, iter_lim=algs._lsqr_limit(proxf, count0 + iiter + 1), atol=kw.get("atol", 1e-6),
                        btol=kw.get("btol", 1e-6), conlim=kw.get("conlim", 1e8))
        with torch.cuda.device(eng.dev):
            rc = lib.lmc_img_pdhg_dual(dt, dims[0], dims[1], mu, g_sigma, 1 if iso else 0, x.data_ptr(), Ax.data_ptr(),
                                       y.data_ptr(), resid[1:].data_ptr(), ws.data_ptr(), nb, stream)
            _lib.check(rc, "lmc_img_pdhg_dual")
            rc = lib.lmc_img_pdhg_primal_resid(dt, dims[0], dims[1], tau, xold.data_ptr(), x.data_ptr(), y.data_ptr(),
                                               ATy.data_ptr(), resid.data_ptr(), ws.data_ptr(), nb, stream)
            _lib.check(rc, "lmc_img_pdhg_primal_resid")
        p, d = (float(v) for v in resid.cpu().numpy())        # synchronises: the step-size rule runs on the host
        if p > s * d * delta:
            tau, mu, alpha = tau / (1 - alpha), mu * (1 - alpha), alpha * eta
        elif p < s * d / delta:
            tau, mu, alpha = tau * (1 - alpha), mu / (1 - alpha), alpha * eta
        steps[iiter] = (tau, mu, alpha)
        if trace is not None:       # debugging aid: residuals and the LSQR work of every iteration
            trace.append({"p": p, "d": d, "tau": tau, "mu": mu, **eng.read_counters()})
        if device_metrics is not None:
            device_metrics._extend(eng.sample_metrics(x, device_metrics.xtrue, g_sigma, iso))
        if callback is not None or (show and algs._show_line(iiter, niter)):
            xh = x[0].double().cpu().numpy()
            if callback is not None:
                callback(xh)
            if show and algs._show_line(iiter, niter):
                pf, pg = proxf(xh), proxg(A.matvec(xh))
                zx = 0. if z is None else float(np.dot(z, xh))
                print('%6g  %12.5e  %10.3e  %10.3e  %10.3e      %10.3e' % (iiter + 1, xh[0], pf, pg, zx, pf + pg + zx))
        iiter += 1
        if max(p, d) < tol:      # never reached in the reference's runs (1000 iterates plotted)
            steps = steps[:iiter]
            break
    if warm_on and niter > 0:
        proxf.x0 = warm[0].double().cpu().numpy()
    if hasattr(proxf, "count"):
        proxf.count += iiter
    if show:
        print('\nTotal time (s) = %.2f' % (time.time() - tstart))
        print('---------------------------------------------------------\n')
    return x[0].double().cpu().numpy(), steps
