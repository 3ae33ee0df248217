"""Drop-in for the sampler class of the reference's ``lmc.py`` (lmc.py:29-190)."""
from __future__ import annotations

import numpy as np
from numpy.random import default_rng

from . import _lib
from .mixtures import MixtureSampler


def gaussian_components(target, mus, Sigmas, omegas):
    """Per-component constants, computed with the numpy calls the reference makes on
    every evaluation (lmc.py:40-42): np.linalg.inv, np.linalg.det."""
    n = len(mus)
    if not 1 <= n <= _lib.MAX_COMP:
        raise ValueError(f"1..{_lib.MAX_COMP} mixture components supported, got {n}")
    target.n = n
    for i in range(n):
        mu = np.asarray(mus[i], dtype=np.float64)
        Sg = np.asarray(Sigmas[i], dtype=np.float64)
        if mu.shape != (2,) or Sg.shape != (2, 2):
            raise ValueError("the CUDA path covers the reference's d=2 experiments")
        P = np.linalg.inv(Sg)
        target.mu[i][0], target.mu[i][1] = mu
        for j, v in enumerate(P.ravel()):
            target.prec[i][j] = v
        target.norm[i] = np.sqrt((2 * np.pi) ** 2 * np.abs(np.linalg.det(Sg)))
        target.omega[i] = float(omegas[i])


class LangevinMonteCarlo(MixtureSampler):
    """Gaussian-mixture target with ULA / MALA / PULA / IHPULA / MLA (lmc.py:29-190).

    Same constructor and method signatures as the reference.  ``dtype`` (extra
    keyword) selects the arithmetic of the single-chain calls; the default float64
    matches the reference's numpy path."""

    _valid_algs = ("ula", "mala", "pula", "ihpula", "mla")
    _hess_shift = 0.05   # lmc.py:162

    def __init__(self, mus, Sigmas, omegas, K=1000, seed=0, dtype="float64") -> None:
        self.mus = mus
        self.Sigmas = Sigmas
        self.omegas = omegas
        self.n = K
        self.seed = seed
        self.d = mus[0].shape[0]
        self.dtype = dtype

    def _target(self):
        t = _lib.MixTarget()
        t.kind = _lib.TARGET_GAUSS
        gaussian_components(t, self.mus, self.Sigmas, self.omegas)
        return t

    # ---- reference call surface ------------------------------------------------
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

    def true_samples(self):
        """Exact draws from the mixture (lmc.py:82-90).  Evaluation input, not part of
        the sampler hot path: plain host numpy with the reference's draw order."""
        rng = default_rng(self.seed)
        out = np.empty((self.n, self.d))
        for k in range(self.n):
            i = rng.integers(0, len(self.mus))
            out[k] = rng.multivariate_normal(self.mus[i], self.Sigmas[i])
        return out
