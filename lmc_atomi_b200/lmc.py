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

    # ---- evaluation helpers of the driver (plot grids; host numpy, not on the sampler path) -----------
    def multivariate_gaussian(self, theta, mu, Sigma):
        """N(theta; mu, Sigma), broadcast over leading dims (lmc.py:39-44)."""
        d = np.asarray(theta, dtype=np.float64) - mu
        q = ((d @ np.linalg.inv(Sigma)) * d).sum(axis=-1)
        return np.exp(-q / 2) / np.sqrt((2 * np.pi) ** self.d * np.abs(np.linalg.det(Sigma)))

    def density_gaussian_mixture(self, theta):
        """lmc.py:46-48."""
        return sum(self.omegas[i] * self.multivariate_gaussian(theta, self.mus[i], self.Sigmas[i])
                   for i in range(len(self.mus)))

    def potential_gaussian_mixture(self, theta):
        return -np.log(self.density_gaussian_mixture(theta))

    def true_samples(self):
        """Exact draws from the mixture (lmc.py:82-90).  Evaluation input, not part of
        the sampler hot path: plain host numpy with the reference's draw order."""
        rng = default_rng(self.seed)
        out = np.empty((self.n, self.d))
        for k in range(self.n):
            i = rng.integers(0, len(self.mus))
            out[k] = rng.multivariate_normal(self.mus[i], self.Sigmas[i])
        return out


M_PRECONDITIONER = np.array([[1.0, 0.1], [0.1, 0.5]])   # lmc.py:278, lmc_laplace.py:315, prox_lmc.py:374
BETA_MIRROR = np.array([0.7, 0.3])                      # lmc.py:284, lmc_laplace.py:318, prox_lmc.py:380


def mixture_parameters(n):
    """Means / covariances of the drivers' n-component experiments (lmc.py:205-231, prox_lmc.py:274-300)."""
    mu = [np.array(m) for m in ([0., 0.], [-2., 3.], [2., -3.], [3., 3.], [-2., -2.])]
    Sg = [np.array(S) for S in ([[1., -0.5], [-0.5, 1.]], [[0.5, 0.2], [0.2, 0.7]], [[0.5, 0.1], [0.1, 0.9]],
                                [[0.8, 0.02], [0.02, 0.3]], [[1.2, 0.05], [0.05, 0.8]])]
    idx = {1: [0], 2: [0, 1], 3: [0, 1, 2], 4: [1, 2, 3, 4], 5: [0, 1, 2, 3, 4]}[n]
    return [mu[i] for i in idx], [Sg[i] for i in idx], np.ones(n) / n


def plot_grid(N=300, lo=-5.0, hi=5.0):
    """The (N,N,2) evaluation grid of the drivers (lmc.py:198-201, 238-240)."""
    X, Y = np.meshgrid(np.linspace(lo, hi, N), np.linspace(lo, hi, N))
    return X, Y, np.stack([X, Y], axis=-1)


def lmc_gaussian_mixture(gamma_ula=5e-2, gamma_mala=5e-2, gamma_pula=5e-2, gamma_ihpula=5e-2, gamma_mla=5e-2, n=5,
                         K=5000, seed=0, dtype="float64", wasserstein=False):
    """Headless ``lmc_gaussian_mixture`` (lmc.py:194-196; same keyword arguments = the CLI flags): runs the five
    samplers on the CUDA path and returns what the reference goes on to plot.  Plotting and KDE are outside the
    hot path.  ``wasserstein=True`` also evaluates the 2-Wasserstein curves of lmc.py:386-416 on the
    device (``w2.wasserstein_curves``, Sinkhorn approximation of the exact EMD) into ``out["wasserstein"]``."""
    mus, Sigmas, omegas = mixture_parameters(n)
    X, Y, pos = plot_grid()
    lmc = LangevinMonteCarlo(mus, Sigmas, omegas, K, seed, dtype=dtype)
    out = {"X": X, "Y": Y, "density": lmc.density_gaussian_mixture(pos), "true_samples": lmc.true_samples()}
    out["ula"] = lmc.ula(gamma_ula)
    out["mala"], out["mala_effective"] = lmc.mala(gamma_mala)
    out["pula"] = lmc.pula(gamma_pula, M_PRECONDITIONER)
    out["ihpula"] = lmc.ihpula(gamma_ihpula)
    out["mla"] = lmc.mla(gamma_mla, BETA_MIRROR)
    if wasserstein:
        from . import w2
        out["wasserstein"] = w2.wasserstein_curves(out["true_samples"],
                                                   {a: out[a] for a in ("ula", "mala", "pula", "ihpula", "mla")})
    return out
