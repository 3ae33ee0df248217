"""Drop-in for the sampler class of ``lmc_laplace.py`` (lmc_laplace.py:31-216)."""
from __future__ import annotations

import numpy as np

from . import _lib
from .lmc import BETA_MIRROR, M_PRECONDITIONER, mixture_parameters, plot_grid
from .mixtures import MixtureSampler


class LangevinMonteCarloLaplacian(MixtureSampler):
    """Moreau-smoothed Laplacian mixture with ULA / MALA / PULA / IHPULA / MLA."""

    _valid_algs = ("ula", "mala", "pula", "ihpula", "mla")
    _hess_shift = 0.02   # lmc_laplace.py:186

    def __init__(self, mus, alphas, omegas, lamda, K=1000, seed=0, dtype="float64") -> None:
        self.mus = mus
        self.alphas = alphas
        self.omegas = omegas
        self.lamda = lamda
        self.n = K
        self.seed = seed
        self.d = mus[0].shape[0]
        self.Sigmas = [np.eye(self.d) * 2 / self.alphas[i] for i in range(len(self.alphas))]
        self.dtype = dtype

    def _target(self):
        n = len(self.mus)
        if not 1 <= n <= _lib.MAX_COMP:
            raise ValueError(f"1..{_lib.MAX_COMP} mixture components supported, got {n}")
        t = _lib.MixTarget()
        t.kind = _lib.TARGET_LAPLACE
        t.n = n
        t.lamda = float(self.lamda)
        for i in range(n):
            mu = np.asarray(self.mus[i], dtype=np.float64)
            if mu.shape != (2,):
                raise ValueError("the CUDA path covers the reference's d=2 experiments")
            t.mu[i][0], t.mu[i][1] = mu
            t.alpha[i] = float(self.alphas[i])
            t.omega[i] = float(self.omegas[i])
        return t

    def ula(self, gamma):
        return self._reference_call("ula", gamma)

    def mala(self, gamma):
        return self._reference_call("mala", gamma)

    def pula(self, gamma, M):
        return self._reference_call("pula", gamma, M=M)

    def ihpula(self, gamma):
        return self._reference_call("ihpula", gamma)

    def mla(self, gamma, beta):
        return self._reference_call("mla", gamma, beta=beta)

    # ---- evaluation helpers of the driver (plot grids; host numpy, not on the sampler path) -----------
    def multivariate_laplacian(self, theta, mu, alpha):
        """(alpha/2)^d exp(-alpha |theta - mu|_1)   (lmc_laplace.py:43-44)."""
        return (alpha / 2) ** self.d * np.exp(-alpha * np.abs(np.asarray(theta, dtype=np.float64) - mu).sum(axis=-1))

    def density_laplacian_mixture(self, theta):
        """lmc_laplace.py:46-48."""
        return sum(self.omegas[i] * self.multivariate_laplacian(theta, self.mus[i], self.alphas[i])
                   for i in range(len(self.alphas)))

    def prox_uncentered_laplace(self, theta, gamma, mu):
        """mu + soft(theta - mu, gamma)   (lmc_laplace.py:53-54, prox.py:18-19)."""
        d = np.asarray(theta, dtype=np.float64) - mu
        return mu + np.sign(d) * np.maximum(np.abs(d) - gamma, 0)

    def moreau_env_uncentered_laplace(self, theta, mu, alpha):
        """lmc_laplace.py:56-58."""
        p = self.prox_uncentered_laplace(theta, self.lamda * alpha, mu)
        return alpha * np.abs(p - mu).sum(axis=-1) + ((p - theta) ** 2).sum(axis=-1) / (2 * self.lamda)

    def smooth_multivariate_laplacian(self, theta, mu, alpha):
        return (alpha / 2) ** self.d * np.exp(-self.moreau_env_uncentered_laplace(theta, mu, alpha))

    def smooth_density_laplacian_mixture(self, theta):
        """lmc_laplace.py:63-65."""
        return sum(self.omegas[i] * self.smooth_multivariate_laplacian(theta, self.mus[i], self.alphas[i])
                   for i in range(len(self.alphas)))

    def true_samples(self):
        """K draws of the (unsmoothed) Laplacian mixture (lmc_laplace.py:99-107): the component index comes from
        ``default_rng(seed)``; the Laplace variates come, as in the reference, from numpy's GLOBAL state
        (``multivariate_laplace.rvs`` is called without ``random_state``, multivariate_laplace.py:436-443), so a
        caller who wants a reproducible set seeds ``np.random.seed`` first -- the draws then equal the reference's.
        Host numpy: evaluation data, not on the sampler path."""
        rng = np.random.default_rng(self.seed)
        out = np.empty((self.n, self.d))
        for k in range(self.n):
            i = rng.integers(0, len(self.mus))
            _, sv, vt = np.linalg.svd(np.asarray(self.Sigmas[i], dtype=np.float64))   # scipy.linalg.svd: same LAPACK gesdd
            z = np.random.laplace(loc=0.0, scale=1.0, size=[1, self.d])
            out[k] = (z @ (np.sqrt(sv)[:, None] * vt) + np.asarray(self.mus[i], dtype=np.float64)).reshape(-1)
        return out


def lmc_laplacian_mixture(gamma_ula=5e-2, gamma_mala=5e-2, gamma_pula=5e-2, gamma_mla=5e-2, lamda=1e-1, alpha=1e-1,
                          n=5, K=5000, seed=0, dtype="float64", wasserstein=False):
    """Headless ``lmc_laplacian_mixture`` (lmc_laplace.py:220-222, same keyword arguments): ULA, MALA, PULA and
    MLA on the smoothed Laplacian mixture through the CUDA path; returns the arrays the reference plots.
    ``wasserstein=True`` adds the 2-Wasserstein curves of lmc_laplace.py:386-451 (first K1 = 10 000 samples of
    every set), evaluated on the device by ``w2.wasserstein_curves``."""
    mus, _, omegas = mixture_parameters(n)
    X, Y, pos = plot_grid()
    s = LangevinMonteCarloLaplacian(mus, np.ones(n) * alpha, omegas, lamda, K, seed, dtype=dtype)
    out = {"X": X, "Y": Y, "density": s.density_laplacian_mixture(pos),
           "smooth_density": s.smooth_density_laplacian_mixture(pos), "true_samples": s.true_samples()}
    out["ula"] = s.ula(gamma_ula)
    out["mala"], out["mala_effective"] = s.mala(gamma_mala)
    out["pula"] = s.pula(gamma_pula, M_PRECONDITIONER)
    out["mla"] = s.mla(gamma_mla, BETA_MIRROR)
    if wasserstein:
        from . import w2
        K1 = 10000                                                                   # lmc_laplace.py:387-392
        out["wasserstein"] = w2.wasserstein_curves(out["true_samples"][:K1],
                                                   {a: out[a][:K1] for a in ("ula", "mala", "pula", "mla")})
    return out
