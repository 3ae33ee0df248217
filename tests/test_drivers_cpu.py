"""Evaluation helpers of the drop-in sampler classes (densities on plot grids, true samples) against values
computed by the reference's own classes (tests/golden/mix_density.npz, oracle/make_golden_density.py), and the
keyword surface of the headless mains.  No GPU: these helpers are host numpy, as in the reference."""
import inspect

import numpy as np

from helpers import load_golden


def _close(a, b):
    return np.abs(np.asarray(a) - np.asarray(b)).max() <= 1e-13 * max(1.0, np.abs(b).max())


def test_density_helpers_match_reference():
    from lmc_atomi_b200 import lmc, lmc_laplace, prox_lmc
    z = load_golden("mix_density.npz")
    pos = z["pos"]
    for n in (1, 4, 5):
        mus, Sg, om = lmc.mixture_parameters(n)
        s = lmc.LangevinMonteCarlo(mus, Sg, om, 48, 3)
        assert _close(s.density_gaussian_mixture(pos), z[f"gauss_n{n}"])
    assert _close(s.density_gaussian_mixture(np.array([0.3, -0.7])), z["gauss_point"])
    assert np.array_equal(s.true_samples(), z["gauss_true_samples"])        # same rng draw order: bit-equal
    mus, _, om = lmc.mixture_parameters(5)
    s = lmc_laplace.LangevinMonteCarloLaplacian(mus, np.ones(5) * 0.5, om, 0.1, 10, 0)
    assert _close(s.density_laplacian_mixture(pos), z["laplace"])
    assert _close(s.smooth_density_laplacian_mixture(pos), z["laplace_smooth"])
    s = lmc_laplace.LangevinMonteCarloLaplacian(mus, np.ones(5) * 0.5, om, 0.1, 40, 2)
    np.random.seed(1234)                                  # the reference's rvs() uses numpy's global state
    assert _close(s.true_samples(), z["laplace_true_samples"])
    mus, Sg, om = lmc.mixture_parameters(5)
    s = prox_lmc.ProximalLangevinMonteCarlo(mus, Sg, om, 0.25, 0.15, np.zeros(2), 10, 0)
    assert _close(s.density_gaussian_mixture(pos) * s.multivariate_laplacian(pos), z["prox"])
    assert _close(s.density_gaussian_mixture(pos) * s.smooth_multivariate_laplacian(pos), z["prox_smooth"])


def test_mains_keep_the_reference_keywords():
    from lmc_atomi_b200 import lmc, lmc_laplace, prox_lmc
    want = {
        lmc.lmc_gaussian_mixture: dict(gamma_ula=5e-2, gamma_mala=5e-2, gamma_pula=5e-2, gamma_ihpula=5e-2,
                                       gamma_mla=5e-2, n=5, K=5000, seed=0),                    # lmc.py:194-196
        lmc_laplace.lmc_laplacian_mixture: dict(gamma_ula=5e-2, gamma_mala=5e-2, gamma_pula=5e-2, gamma_mla=5e-2,
                                                lamda=1e-1, alpha=1e-1, n=5, K=5000, seed=0),   # lmc_laplace.py:220-222
        prox_lmc.prox_lmc_gaussian_mixture: dict(gamma_pgld=5e-2, gamma_myula=5e-2, gamma_mymala=5e-2,
                                                 gamma_ppula=5e-2, gamma_fbula=5e-2, gamma_lbmumla=5e-2, lamda=0.01,
                                                 alpha=.1, n=5, t=100, K=10000, seed=0),        # prox_lmc.py:259-263
    }
    for fn, kw in want.items():
        sig = inspect.signature(fn).parameters
        for k, v in kw.items():
            assert k in sig and sig[k].default == v, (fn.__name__, k)
        assert list(sig)[:len(kw)] == list(kw)
