"""Drop-in for the sampler class of ``lmc_laplace.py`` (lmc_laplace.py:31-216)."""
from __future__ import annotations

import numpy as np

from . import _lib
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
