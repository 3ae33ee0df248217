"""Drop-in for the sampler class of ``prox_lmc.py`` (prox_lmc.py:29-255)."""
from __future__ import annotations

import numpy as np

from . import _lib
from .lmc import BETA_MIRROR, M_PRECONDITIONER, LangevinMonteCarlo, gaussian_components, mixture_parameters, plot_grid
from .mixtures import MixtureSampler


class ProximalLangevinMonteCarlo(MixtureSampler):
    """Gaussian mixture x Laplacian prior with PGLD / MYULA / MYMALA / PPULA / FBULA /
    LBMUMLA."""

    _valid_algs = ("pgld", "myula", "mymala", "ppula", "fbula", "lbmumla")

    def __init__(self, mus, Sigmas, omegas, lamda, alpha, mu, K=1000, seed=0, dtype="float64") -> None:
        self.mus = mus
        self.Sigmas = Sigmas
        self.omegas = omegas
        self.lamda = lamda
        self.alpha = alpha
        self.mu = mu
        self.n = K
        self.seed = seed
        self.d = mus[0].shape[0]
        self.dtype = dtype

    def _target(self):
        t = _lib.MixTarget()
        t.kind = _lib.TARGET_GAUSS_L1
        gaussian_components(t, self.mus, self.Sigmas, self.omegas)
        t.lamda = float(self.lamda)
        t.prior_alpha = float(self.alpha)
        mu = np.asarray(self.mu, dtype=np.float64)
        t.prior_mu[0], t.prior_mu[1] = mu
        return t

    def pgld(self, gamma):
        return self._reference_call("pgld", gamma)

    def myula(self, gamma):
        return self._reference_call("myula", gamma)

    def mymala(self, gamma):
        return self._reference_call("mymala", gamma)

    def ppula(self, gamma, M, Q, t=100):
        return self._reference_call("ppula", gamma, M=M, Q=Q, t=t)

    def fbula(self, gamma):
        return self._reference_call("fbula", gamma)

    def lbmumla(self, gamma, beta, sigma):
        return self._reference_call("lbmumla", gamma, beta=beta, sigma=sigma)

    # ---- evaluation helpers of the driver (plot grids; host numpy, not on the sampler path) -----------
    multivariate_gaussian = LangevinMonteCarlo.multivariate_gaussian          # prox_lmc.py:42-47
    density_gaussian_mixture = LangevinMonteCarlo.density_gaussian_mixture    # prox_lmc.py:49-51

    def multivariate_laplacian(self, theta):
        """The Laplacian prior (prox_lmc.py:56-57)."""
        return (self.alpha / 2) ** self.d * np.exp(
            -self.alpha * np.abs(np.asarray(theta, dtype=np.float64) - self.mu).sum(axis=-1))

    def moreau_env_uncentered_laplace(self, theta):
        """prox_lmc.py:62-64."""
        d = np.asarray(theta, dtype=np.float64) - self.mu
        p = self.mu + np.sign(d) * np.maximum(np.abs(d) - self.lamda * self.alpha, 0)
        return self.alpha * np.abs(p - self.mu).sum(axis=-1) + ((p - theta) ** 2).sum(axis=-1) / (2 * self.lamda)

    def smooth_multivariate_laplacian(self, theta):
        """prox_lmc.py:66-67."""
        return (self.alpha / 2) ** self.d * np.exp(-self.moreau_env_uncentered_laplace(theta))


Q_PRECONDITIONER = np.array([[1.0, 0.1], [0.1, 1.5]])   # prox_lmc.py:375
SIGMA_BREGMAN = np.array([0.8, 0.2])                    # prox_lmc.py:381


def prox_lmc_gaussian_mixture(gamma_pgld=5e-2, gamma_myula=5e-2, gamma_mymala=5e-2, gamma_ppula=5e-2,
                              gamma_fbula=5e-2, gamma_lbmumla=5e-2, lamda=0.01, alpha=.1, n=5, t=100, K=10000,
                              seed=0, dtype="float64"):
    """Headless ``prox_lmc_gaussian_mixture`` (prox_lmc.py:259-263, same keyword arguments): the six proximal
    samplers through the CUDA path; returns the arrays the reference plots."""
    mus, Sigmas, omegas = mixture_parameters(n)
    X, Y, pos = plot_grid()
    s = ProximalLangevinMonteCarlo(mus, Sigmas, omegas, lamda, alpha, np.zeros_like(mus[0]), K, seed, dtype=dtype)
    gm = s.density_gaussian_mixture(pos)
    out = {"X": X, "Y": Y, "density": gm * s.multivariate_laplacian(pos),
           "smooth_density": gm * s.smooth_multivariate_laplacian(pos)}
    out["pgld"] = s.pgld(gamma_pgld)
    out["myula"] = s.myula(gamma_myula)
    out["mymala"], out["mymala_effective"] = s.mymala(gamma_mymala)
    out["ppula"] = s.ppula(gamma_ppula, M_PRECONDITIONER, Q_PRECONDITIONER, t)
    out["fbula"] = s.fbula(gamma_fbula)
    out["lbmumla"] = s.lbmumla(gamma_lbmumla, BETA_MIRROR, SIGMA_BREGMAN)
    return out
