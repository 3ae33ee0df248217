"""Headless ``prox_lmc_deconv`` (prox_lmc_deconv.py:40-42, same keyword arguments = the CLI flags).

Builds the operators exactly as the driver does (prox_lmc_deconv.py:52-125: H5/H6/H7 box blurs, observation
y = H5 img + N(0, sigma^2) from ``default_rng(seed)``, L2 / MC-TV / ME-TV data terms, TV / L21 / L1 priors) and
runs the nine sampling experiments (:447-703) on the CUDA path: ULPDA or MYULA for blur sizes 5, 6, 7 and data
terms L2, MC-TV, ME-TV.  The per-iteration host callback of the driver (:128-133) is replaced by
``algs.DeviceMetrics`` (the samples never leave the GPU except for the returned posterior mean).
``compute_MAP=True`` runs the driver's other branch (:138-445): adaptive PDHG MAP estimates of the same nine models
(``primaldual.AdaptivePrimalDual``).  Plotting is out of scope."""
from __future__ import annotations

import os

import numpy as np
from numpy.random import default_rng

from . import algs, compat, deconv_bench, primaldual

MODELS = tuple((k, term) for term in ("l2", "mc", "me") for k in (5, 6, 7))   # order of prox_lmc_deconv.py:447-703


def load_image(image):
    """'einstein' reads fig/einstein.png as the driver does (:45-46) when it is there; skimage's 'camera' and
    scipy's 'ascent' are not available in this image: a deterministic 512x512 stand-in is used.
    An ndarray is taken as is."""
    if isinstance(image, np.ndarray):
        return np.asarray(image, dtype=np.float64)
    if image == "einstein" and os.path.isfile("fig/einstein.png"):
        from PIL import Image
        return np.asarray(Image.open("fig/einstein.png").convert("L"), dtype=np.float64)
    if image in ("einstein", "camera", "ascent"):
        return deconv_bench.synthetic_image(512, 512)
    raise ValueError(f"unknown image {image!r}")


def prox_lmc_deconv(gamma_mc=15., gamma_me=15., sigma=0.75, tau=0.3, N=1000, niter_l2=50, niter_tv=10,
                    niter_MAP=1000, image='camera', alg='ULPDA', compute_MAP=False, seed=0, models=MODELS,
                    dtype="float64", noise=None):
    """Keyword arguments of prox_lmc_deconv.py:40-42, plus ``models`` (subset of MODELS), ``dtype`` of the kernels
    and ``noise``: the (N, P) N(0,1) draws shared by all models (every sampler call of the driver re-seeds
    ``default_rng(seed)``, algs.py:431,561, so all nine models see the same stream; passing it once, e.g. as a
    device tensor, avoids drawing it nine times).  Returns a dict: per model the posterior mean, the per-sample
    metric curves of the driver's callback (:128-133) and the PSNR of the mean (:707-735)."""
    if not compute_MAP and alg not in ("ULPDA", "MYULA"):
        raise ValueError("alg must be 'ULPDA' or 'MYULA'")
    img = load_image(image)
    ny, nx = img.shape
    rng = default_rng(seed)
    H = {k: compat.Convolve2D((ny, nx), h=np.ones((k, k)) / k ** 2, offset=(k // 2, k // 2)) for k in (5, 6, 7)}
    y = H[5] * img.ravel() + rng.normal(0, sigma, size=(ny, nx)).ravel()      # :59
    L = 1. / sigma ** 2
    tau0, mu0 = 0.95 / L, 1.                                                   # :88-90
    gamma_myula = 1. / L                                                       # :92-94
    tau_myula = 0.2 * gamma_myula
    Gop = compat.Gradient(dims=(ny, nx), sampling=1., edge=False, kind='forward', dtype='float64')

    def data_term(k, term):
        if term == "l2":
            return compat.L2(Op=H[k], b=y, sigma=1 / sigma ** 2, niter=50, warm=True)                       # :101-103
        if term == "mc":
            return algs.L2_ncvx_tv(dims=(ny, nx), Op=H[k], Op2=Gop, b=y, sigma=1 / sigma ** 2, lamda=tau,
                                   gamma=gamma_mc, isotropic=True, niter=niter_l2, warm=True)                # :106-108
        return algs.L2_ncvx_tv(dims=(ny, nx), Op=H[k], b=y, sigma=1 / sigma ** 2, lamda=tau, gamma=gamma_me,
                               isotropic=True, niter=niter_l2, warm=True)                                    # :111-113

    l1iso, l1 = compat.L21(ndim=2, sigma=tau), compat.L1(sigma=tau)            # :116,119
    tv = compat.TV(dims=(ny, nx), sigma=tau, niter=niter_tv)                   # :122
    x0 = np.zeros(ny * nx)
    out = {"image": img, "observation": y.reshape(ny, nx), "alg": "MAP" if compute_MAP else alg, "models": {}}
    for k, term in models:
        f = data_term(k, term)
        dm = algs.DeviceMetrics(img.ravel())
        if compute_MAP:     # :138-445: adaptive PDHG, the driver's `if compute_MAP: ... else: <samplers>`
            g = l1 if term == "mc" else l1iso
            xmap, _ = primaldual.AdaptivePrimalDual(f, g, Gop, tau=tau0, mu=mu0, x0=x0, niter=niter_MAP, dtype=dtype,
                                                    device_metrics=dm)
            out["models"][f"{term}_{k}"] = {
                "map": xmap.reshape(ny, nx), "cost_l2_plus_g": dm.cost, "err": dm.err, "snr": dm.snr, "psnr": dm.psnr,
                "mse": dm.mse, "map_psnr": float(10 * np.log10(255. ** 2 / np.mean((xmap - img.ravel()) ** 2)))}
            continue
        if alg == "ULPDA":
            g = l1 if term == "mc" else l1iso          # the driver pairs MC-TV with the anisotropic prior (:542)
            s = algs.UnadjustedLangevinPrimalDual(f, g, Gop, tau=tau0, mu=mu0, theta=1., x0=x0, gfirst=False,
                                                  niter=N, seed=seed, dtype=dtype, device_metrics=dm, noise=noise,
                                                  reduce="mean")
        else:
            s = algs.MoreauYosidaUnadjustedLangevin(f, tv, tau=tau_myula, gamma=gamma_myula, x0=x0, niter=N,
                                                    seed=seed, dtype=dtype, device_metrics=dm, noise=noise,
                                                    reduce="mean")
        mean = s                                                               # samples.mean(axis=0) of :474-475
        out["models"][f"{term}_{k}"] = {
            "mean": mean.reshape(ny, nx), "cost_l2_plus_g": dm.cost, "err": dm.err, "snr": dm.snr, "psnr": dm.psnr,
            "mse": dm.mse, "mean_psnr": float(10 * np.log10(255. ** 2 / np.mean((mean - img.ravel()) ** 2)))}
    return out
