"""Drop-in for the sampler class of ``prox_lmc.py`` (prox_lmc.py:29-255)."""
from __future__ import annotations

import numpy as np

from . import _lib
from .lmc import gaussian_components
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
