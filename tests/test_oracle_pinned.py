"""The image-path oracle against evidence the REFERENCE holds: the per-sample SNR / PSNR / MSE curves of its
committed result figures (fig/fig_prox_lmc_deconv_einstein_{MYULA,ULPDA}_1000_snr_psnr_mse.pdf, written by
prox_lmc_deconv.py:810-853 from the run of expt_deconv.sh:5-6), digitised by oracle/digitise_figs.py into
tests/golden/deconv_curves.npz, on the reference's own input image (tests/golden/einstein.png = fig/einstein.png).

Resolution of the evidence: the PDF quotes vertex coordinates to 1e-6 pt = 2.2e-7 dB on the SNR / PSNR axes and
1.1e-4 on the MSE axis; every kept vertex is an original (sample index, value) pair.  CPU only, ~40 s."""
import os

import numpy as np
import pytest

from oracle import deconv_ref as D

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
DB_TOL = 5e-7          # ~2x the quantisation of the plot
MSE_ABS_TOL = 3e-4


@pytest.fixture(scope="module")
def einstein():
    from PIL import Image
    img = np.array(Image.open(os.path.join(GOLDEN, "einstein.png")))
    assert img.shape == (667, 877) and img.dtype == np.uint8
    return img


@pytest.fixture(scope="module")
def curves():
    return np.load(os.path.join(GOLDEN, "deconv_curves.npz"))


def _compare(curves, fig, model, snr, psnr, mse, n):
    worst = {}
    for name, got, tol in (("snr", snr, DB_TOL), ("psnr", psnr, DB_TOL), ("mse", mse, MSE_ABS_TOL)):
        idx, val = curves[f"{fig}_{name}_M{model}_idx"], curves[f"{fig}_{name}_M{model}_val"]
        keep = idx < n
        assert keep.sum() >= 3
        d = np.abs(np.asarray(got)[idx[keep]] - val[keep])
        worst[name] = float(d.max())
        assert d.max() < tol, (fig, model, name, idx[keep], d)
    return worst


def test_golden_curves_are_self_consistent(einstein, curves):
    """PSNR = 10 log10(255^2 / MSE) and SNR = PSNR + 10 log10(mean(x*^2) / 255^2) hold between the three digitised
    axes (independent calibrations) wherever two curves kept the same sample."""
    shift = 10 * np.log10(np.mean(einstein.astype(np.float64) ** 2) / 255. ** 2)
    n = 0
    for fig in ("einstein_MYULA", "einstein_ULPDA", "einstein_MAP"):
        for m in range(1, 10):
            s_i, s_v = curves[f"{fig}_snr_M{m}_idx"], curves[f"{fig}_snr_M{m}_val"]
            p_i, p_v = curves[f"{fig}_psnr_M{m}_idx"], curves[f"{fig}_psnr_M{m}_val"]
            e_i, e_v = curves[f"{fig}_mse_M{m}_idx"], curves[f"{fig}_mse_M{m}_val"]
            assert s_i[0] == 0 and s_i[-1] == 999 and np.all(np.diff(s_i) > 0)
            common = np.intersect1d(s_i, p_i)
            d = s_v[np.searchsorted(s_i, common)] - p_v[np.searchsorted(p_i, common)]
            assert np.abs(d - shift).max() < 1e-6
            common = np.intersect1d(p_i, e_i)
            pe = 10 * np.log10(255. ** 2 / e_v[np.searchsorted(e_i, common)])
            # the MSE axis resolves 1e-4 absolute: loose in dB where the MSE is small
            tol = 10 / np.log(10) * 2e-4 / e_v[np.searchsorted(e_i, common)] + 1e-6
            assert np.all(np.abs(pe - p_v[np.searchsorted(p_i, common)]) < tol)
            n += 1
    assert n == 27


def _run_myula(img, n, tv_sigma_in_obj=True):
    pb = D.build_problem(img)
    xt = img.ravel()
    snr, psnr, mse = [], [], []
    tv = pb["tv"]()
    if not tv_sigma_in_obj:      # negative control: the pre-pinning recollection (objective without TV's sigma)
        tv.prox = lambda x, tau: D.tv_prox_2d(x.reshape(tv.dims), tv.sigma * tau, tv.niter, tv.rtol)[0].ravel()
    D.MoreauYosidaUnadjustedLangevin(pb["l2"](), tv, pb["x0"], tau=pb["tau_myula"], gamma=pb["gamma_myula"], niter=n,
                                     seed=0, callback=lambda x: (snr.append(D.signal_noise_ratio(xt, x)),
                                                                 psnr.append(D.psnr(xt, x)), mse.append(D.mse(xt, x))))
    return snr, psnr, mse


def test_myula_m1_matches_reference_figure(einstein, curves):
    """Pins Convolve2D (H5, zero boundary), the observation noise, the sampler noise stream, L2.grad and TV.prox
    (loop bound, objective with TV's sigma, FISTA recurrence, returned iterate)."""
    n = 24
    worst = _compare(curves, "einstein_MYULA", 1, *_run_myula(einstein, n), n)
    print("MYULA M1 worst |diff|:", worst)


def test_myula_m1_negative_control(einstein, curves):
    """The evidence discriminates: leaving TV's sigma out of the exit-test objective drifts by ~1e-5 dB."""
    n = 24
    snr, _, _ = _run_myula(einstein, n, tv_sigma_in_obj=False)
    idx, val = curves["einstein_MYULA_snr_M1_idx"], curves["einstein_MYULA_snr_M1_val"]
    keep = idx < n
    d = np.abs(np.asarray(snr)[idx[keep]] - val[keep])
    assert d.max() > 10 * DB_TOL, d


def test_ulpda_m1_matches_reference_figure(einstein, curves):
    """Pins Gradient / its adjoint, L2.prox (LSQR on I + tau sigma H^T H, warm start) and L21.proxdual."""
    n = 9
    pb = D.build_problem(einstein)
    xt = einstein.ravel()
    snr, psnr, mse = [], [], []
    D.UnadjustedLangevinPrimalDual(pb["l2"](), pb["l1iso"](), pb["Gop"], pb["x0"], tau=pb["tau0"], mu=pb["mu0"],
                                   theta=1., gfirst=False, niter=n, seed=0,
                                   callback=lambda x: (snr.append(D.signal_noise_ratio(xt, x)),
                                                       psnr.append(D.psnr(xt, x)), mse.append(D.mse(xt, x))))
    worst = _compare(curves, "einstein_ULPDA", 1, snr, psnr, mse, n)
    print("ULPDA M1 worst |diff|:", worst)


def test_map_m1_matches_reference_figure(einstein, curves):
    """Pins the MAP branch (pyproximal AdaptivePrimalDual, prox_lmc_deconv.py:138-153): l2 residual norms and the
    (alpha, eta, s, delta) = (0.5, 0.95, 1, 1.5) adaptation, through the first step-size changes."""
    n = 13
    pb = D.build_problem(einstein)
    xt = einstein.ravel()
    snr, psnr, mse = [], [], []
    _, steps = D.AdaptivePrimalDual(pb["l2"](), pb["l1iso"](), pb["Gop"], pb["x0"], pb["tau0"], pb["mu0"], niter=n,
                                    callback=lambda x: (snr.append(D.signal_noise_ratio(xt, x)),
                                                        psnr.append(D.psnr(xt, x)), mse.append(D.mse(xt, x))))
    assert len({round(t, 12) for t, _, _ in steps}) > 1          # the step size did adapt inside the window
    worst = _compare(curves, "einstein_MAP", 1, snr, psnr, mse, n)
    print("MAP M1 worst |diff|:", worst)
