"""North-star statistics gates beyond trajectory parity:
  * 2-Wasserstein distances to the true samples (lmc.py:386-416) of the CUDA chains vs the reference's chains:
    within 1e-3 (exact EMD by oracle/w2_ref.py, K small enough for the LP);
  * posterior-mean PSNR of the image samplers within 0.05 dB of the oracle's, and the on-device per-sample
    metrics (csrc/img_metrics.cu) equal to what the reference's host callback computes."""
import numpy as np
import pytest

from helpers import alg_params, make_samplers

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("which,alg", [("gauss", "ula"), ("gauss", "mla"), ("gauss", "ihpula"), ("laplace", "pula"),
                                       ("prox", "myula"), ("prox", "lbmumla")])
def test_wasserstein_to_true_samples_matches_reference(which, alg):
    from oracle import w2_ref as W
    K = 160
    z, s, _ = make_samplers(K=K)[which]
    ref = z[alg][:K]                       # the reference's own chain (golden, same seed stream)
    kw, _ = alg_params(z, alg)
    got = getattr(s, alg)(float(z["gamma"]), **kw)
    assert got.shape == ref.shape
    # the reference's true samples where the class provides them (lmc.py:82-90), else a fixed K-point cloud
    true = s.true_samples() if which == "gauss" else np.random.default_rng(0).normal(size=(K, 2)) * 1.5
    for k in (20, 80, K - 1):
        w_ref, w_got = W.wasserstein2(true, ref[:k + 1]), W.wasserstein2(true, got[:k + 1])
        assert abs(w_ref - w_got) < 1e-3, (k, w_ref, w_got)


def _deconv(dims, k=5):
    from oracle import deconv_ref as D
    from lmc_atomi_b200 import compat
    pb = D.build_problem(D.synthetic_image(*dims), kernel=k)
    hk = np.ones((k, k)) / k ** 2
    H = compat.Convolve2D(dims, h=hk, offset=(k // 2, k // 2))
    l2 = compat.L2(Op=H, b=pb["y"].ravel(), sigma=1 / 0.75 ** 2, niter=50, warm=True)
    return D, compat, pb, l2


def _callback(D, f, g, K, xtrue, store):
    def cb(x):
        store["cost"].append(f(x) + g(K.matvec(x)))
        store["err"].append(np.linalg.norm(x - xtrue))
        store["snr"].append(D.signal_noise_ratio(xtrue, x))
        store["psnr"].append(D.psnr(xtrue, x))
        store["mse"].append(D.mse(xtrue, x))
    return cb


@pytest.mark.parametrize("alg", ["myula", "ulpda_iso", "ulpda_l1"])
def test_device_metrics_equal_reference_callback_and_mean_psnr(alg):
    from lmc_atomi_b200 import algs
    dims, N = (48, 44), 25
    D, compat, pb, l2 = _deconv(dims)
    xtrue = pb["img"].ravel().astype(np.float64)
    store = {k: [] for k in algs.DeviceMetrics.KEYS}
    dm = algs.DeviceMetrics(xtrue)
    if alg == "myula":
        f, g = pb["l2"](), pb["tv"]()
        want = D.MoreauYosidaUnadjustedLangevin(f, g, pb["x0"], tau=pb["tau_myula"], gamma=pb["gamma_myula"], niter=N,
                                                seed=0, callback=_callback(D, f, g, D.Identity(xtrue.size), xtrue, store))
        got = algs.MoreauYosidaUnadjustedLangevin(l2, compat.TV(dims, sigma=0.3, niter=10), pb["x0"],
                                                  tau=pb["tau_myula"], gamma=pb["gamma_myula"], niter=N, seed=0,
                                                  device_metrics=dm)
    else:
        iso = alg == "ulpda_iso"
        f, g = pb["l2"](), (pb["l1iso"]() if iso else pb["l1"]())
        want = D.UnadjustedLangevinPrimalDual(f, g, pb["Gop"], pb["x0"], pb["tau0"], pb["mu0"], theta=1., niter=N,
                                              seed=0, gfirst=False,
                                              callback=_callback(D, f, g, pb["Gop"], xtrue, store))
        gm = compat.L21(ndim=2, sigma=0.3) if iso else compat.L1(sigma=0.3)
        got = algs.UnadjustedLangevinPrimalDual(l2, gm, compat.Gradient(dims), pb["x0"], tau=pb["tau0"], mu=pb["mu0"],
                                                theta=1., niter=N, seed=0, gfirst=False, device_metrics=dm)
    for k in algs.DeviceMetrics.KEYS:
        a, b = np.array(getattr(dm, k)), np.array(store[k])
        assert a.shape == b.shape == (N,)
        assert np.abs(a - b).max() <= 1e-9 * np.abs(b).max(), k
    # posterior mean PSNR (prox_lmc_deconv.py:474-475 + :131) within 0.05 dB
    assert abs(D.psnr(xtrue, got.mean(axis=0)) - D.psnr(xtrue, want.mean(axis=0))) < 0.05


def test_fp32_production_posterior_mean_psnr_within_tolerance():
    """fp32 MYULA (on-chip kernel) with the reference's noise stream: posterior-mean PSNR vs the fp64 oracle."""
    from lmc_atomi_b200 import algs
    dims, N = (64, 64), 60
    D, compat, pb, l2 = _deconv(dims)
    xtrue = pb["img"].ravel().astype(np.float64)
    want = D.MoreauYosidaUnadjustedLangevin(pb["l2"](), pb["tv"](), pb["x0"], tau=pb["tau_myula"],
                                            gamma=pb["gamma_myula"], niter=N, seed=0)
    got = algs.MoreauYosidaUnadjustedLangevin(l2, compat.TV(dims, sigma=0.3, niter=10), pb["x0"], tau=pb["tau_myula"],
                                              gamma=pb["gamma_myula"], niter=N, seed=0, dtype="float32")
    assert abs(D.psnr(xtrue, got.mean(axis=0)) - D.psnr(xtrue, want.mean(axis=0))) < 0.05


@pytest.mark.parametrize("alg", ["ULPDA", "MYULA"])
def test_headless_deconv_main_matches_oracle(alg):
    """prox_lmc_deconv(...) with the driver's keywords on a small image: posterior means and PSNR of three of
    the nine models (L2/H5, MC-TV/H6, ME-TV/H7) against the oracle's run of the same experiment."""
    from oracle import deconv_ref as D
    from lmc_atomi_b200.prox_lmc_deconv import prox_lmc_deconv
    img = D.synthetic_image(40, 36).astype(np.float64)
    N, models = 5, ((5, "l2"), (6, "mc"), (7, "me"))
    out = prox_lmc_deconv(N=N, image=img, alg=alg, seed=0, models=models)
    assert set(out["models"]) == {"l2_5", "mc_6", "me_7"}
    for k, term in models:
        pb = D.build_problem(img, kernel=k)
        assert np.abs(out["observation"] - pb["y"]).max() < 1e-12
        f = pb["l2"]() if term == "l2" else pb["l2_" + term]()
        if alg == "ULPDA":
            g = pb["l1"]() if term == "mc" else pb["l1iso"]()
            want = D.UnadjustedLangevinPrimalDual(f, g, pb["Gop"], pb["x0"], pb["tau0"], pb["mu0"], theta=1.,
                                                  niter=N, seed=0, gfirst=False)
        else:
            want = D.MoreauYosidaUnadjustedLangevin(f, pb["tv"](), pb["x0"], tau=pb["tau_myula"],
                                                    gamma=pb["gamma_myula"], niter=N, seed=0)
        m = out["models"][f"{term}_{k}"]
        assert np.abs(m["mean"].ravel() - want.mean(axis=0)).max() <= 1e-8 * np.abs(want).max()
        assert abs(m["mean_psnr"] - D.psnr(img.ravel(), want.mean(axis=0))) < 0.05
        assert len(m["psnr"]) == N and abs(m["psnr"][-1] - D.psnr(img.ravel(), want[-1])) < 1e-6


def test_headless_mixture_mains_reproduce_golden_chains():
    from helpers import load_golden
    from lmc_atomi_b200.lmc import lmc_gaussian_mixture
    from lmc_atomi_b200.lmc_laplace import lmc_laplacian_mixture
    from lmc_atomi_b200.prox_lmc import prox_lmc_gaussian_mixture
    K = 300
    g = load_golden("mix_gauss.npz")
    out = lmc_gaussian_mixture(0.1, 0.1, 0.1, 0.1, 0.1, n=5, K=K, seed=0)
    for a in ("ula", "pula", "ihpula", "mla"):
        assert np.abs(out[a] - g[a][:K]).max() < 1e-11
    assert out["density"].shape == (300, 300) and out["true_samples"].shape == (K, 2)
    l = load_golden("mix_laplace.npz")
    out = lmc_laplacian_mixture(0.05, 0.05, 0.05, 0.05, lamda=0.1, alpha=0.5, n=5, K=K, seed=0)
    for a in ("ula", "pula", "mla"):
        assert np.abs(out[a] - l[a][:K]).max() < 1e-11
    assert out["mala"].shape == (out["mala_effective"], 2) and out["true_samples"].shape == (K, 2)
    p = load_golden("mix_prox.npz")
    out = prox_lmc_gaussian_mixture(0.05, 0.05, 0.05, 0.05, 0.05, 0.05, lamda=0.25, alpha=0.15, n=5, t=100, K=K, seed=0)
    for a in ("pgld", "myula", "ppula", "fbula", "lbmumla"):
        assert np.abs(out[a] - p[a][:K]).max() < 1e-10
