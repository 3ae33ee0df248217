"""TEST INFRASTRUCTURE ONLY -- CPU oracle for the mixture samplers.

A numpy float64 restatement of the reference's low-dimensional sampler loops,
written as pure functions of (state, noise) so that the *same* noise can be fed
to the CUDA kernels.  It is pinned against the unmodified reference classes
(executed through ``oracle/refshim.py``) by ``tests/test_oracle_mix.py`` and by
the committed golden vectors in ``tests/golden/`` (made by
``oracle/make_golden.py``): max abs difference 0.0 on every sampler.

Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s cpu_baseline /
``--impl reference`` leg may import this module.  The product path
(``lmc_atomi_b200``) never does.

Reference lines followed (all under /root/reference):
  Gaussian mixture target .......... lmc.py:39-78 (dup prox_lmc.py:42-95)
  smoothed Laplacian mixture ....... lmc_laplace.py:43-95, prox.py:18-19
  Laplacian prior / envelopes ...... prox_lmc.py:56-67,114-118,165-176,192-196,218-242
  sampler loops .................... lmc.py:94-190, lmc_laplace.py:110-216,
                                     prox_lmc.py:99-255
The arithmetic order of every expression mirrors the reference so that results
are bit-identical; the deliberately preserved quirks are listed in DESIGN.md.
"""
from __future__ import annotations

import numpy as np
from numpy.random import default_rng
from scipy.linalg import sqrtm
from scipy.stats import multivariate_normal


# --------------------------------------------------------------------------- noise
def reference_noise(seed: int, K: int, d: int = 2, uniforms: bool = False):
    """The reference's random stream as arrays.

    Every sampler does ``rng = default_rng(seed)``, ``theta0 = rng.standard_normal(d)``
    (lmc.py:96-97; ``rng.normal(0, 1, d)`` in lmc_laplace.py:113 yields the same
    bits), then per step ``rng.multivariate_normal(0, I_d)`` (same bits as
    ``standard_normal(d)`` -- checked in tests/test_oracle_mix.py) and, for the
    Metropolis-adjusted samplers only, one ``rng.random()`` *after* the Gaussian
    draw (lmc.py:124-128).
    """
    rng = default_rng(seed)
    theta0 = rng.standard_normal(d)
    xi = np.empty((K, d))
    u = np.empty(K) if uniforms else None
    if uniforms:
        for k in range(K):
            xi[k] = rng.standard_normal(d)
            u[k] = rng.random()
    else:
        # draws are sequential in the bit stream, so one bulk call is identical
        xi[:] = rng.standard_normal((K, d))
    return theta0, xi, u


def soft(x, thr):
    """prox.py:18-19 (soft threshold)."""
    return np.sign(x) * np.maximum(np.abs(x) - thr, 0)


# --------------------------------------------------------------------------- targets
class GaussMix:
    """lmc.py:39-78.  ``inv``/``det`` are re-evaluated on every call exactly as the
    reference does (that is where its time goes)."""

    def __init__(self, mus, Sigmas, omegas):
        self.mus, self.Sigmas, self.omegas = mus, Sigmas, omegas
        self.d = mus[0].shape[0]

    def comp(self, th, mu, Sg):
        nrm = np.sqrt((2 * np.pi) ** self.d * np.abs(np.linalg.det(Sg)))
        P = np.linalg.inv(Sg)
        quad = np.einsum('...k,kl,...l->...', th - mu, P, th - mu)
        return np.exp(-quad / 2) / nrm

    def density(self, th):
        return sum(self.omegas[i] * self.comp(th, self.mus[i], self.Sigmas[i])
                   for i in range(len(self.mus)))

    def grad_density(self, th):
        return sum(self.omegas[i] * (self.comp(th, self.mus[i], self.Sigmas[i])
                                     * np.linalg.inv(self.Sigmas[i]) @ (self.mus[i] - th))
                   for i in range(len(self.mus)))

    def grad_U(self, th):
        return -self.grad_density(th) / self.density(th)

    def hess_density(self, th):
        out = 0
        for i in range(len(self.mus)):
            P = np.linalg.inv(self.Sigmas[i])
            dd = th - self.mus[i]
            out = out + self.omegas[i] * (self.comp(th, self.mus[i], self.Sigmas[i])
                                          * (P @ np.outer(dd, dd) @ P - P))
        return out

    def hess_U(self, th):
        p = self.density(th)
        gp = self.grad_density(th)
        Hp = self.hess_density(th)
        return np.outer(gp, gp) / p ** 2 - Hp / p

    # density used in the MH ratio
    mh_density = density
    hess_shift = 0.05          # lmc.py:162


class LaplaceMix:
    """lmc_laplace.py:43-95 (Moreau-smoothed Laplacian mixture)."""

    def __init__(self, mus, alphas, omegas, lamda):
        self.mus, self.alphas, self.omegas, self.lamda = mus, alphas, omegas, lamda
        self.d = mus[0].shape[0]

    def prox_c(self, th, thr, mu):
        return mu + soft(th - mu, thr)

    def env(self, th, mu, a):
        p = self.prox_c(th, self.lamda * a, mu)
        return (a * np.linalg.norm(p - mu, ord=1, axis=-1)
                + np.linalg.norm(p - th, ord=2, axis=-1) ** 2 / (2 * self.lamda))

    def comp(self, th, mu, a):
        return (a / 2) ** self.d * np.exp(-self.env(th, mu, a))

    def raw_density(self, th):
        return sum(self.omegas[i] * ((self.alphas[i] / 2) ** self.d
                                     * np.exp(-self.alphas[i] * np.linalg.norm(th - self.mus[i], ord=1, axis=-1)))
                   for i in range(len(self.alphas)))

    def density(self, th):
        return sum(self.omegas[i] * self.comp(th, self.mus[i], self.alphas[i])
                   for i in range(len(self.alphas)))

    def grad_density(self, th):
        return sum(self.omegas[i] * (self.comp(th, self.mus[i], self.alphas[i])
                                     * (self.prox_c(th, self.lamda * self.alphas[i], self.mus[i]) - th)
                                     / self.lamda)
                   for i in range(len(self.alphas)))

    def grad_U(self, th):
        return -self.grad_density(th) / self.density(th)

    def hess_density(self, th):
        # lmc_laplace.py:80-82: note ||theta||_1 *without* mu and the unsmoothed
        # exponent -- reproduced as is.
        out = 0
        for i in range(len(self.alphas)):
            a, mu = self.alphas[i], self.mus[i]
            g = (th - self.prox_c(th, self.lamda * a, mu)) / self.lamda
            out = out + self.omegas[i] * ((a / 2) ** self.d
                                          * np.exp(-a * np.linalg.norm(th, ord=1, axis=-1))
                                          * (-np.eye(self.d) + np.outer(g, g)))
        return out

    def hess_U(self, th):
        p = self.density(th)
        gp = self.grad_density(th)
        Hp = self.hess_density(th)
        return np.outer(gp, gp) / p ** 2 - Hp / p

    mh_density = density       # lmc_laplace.py:130-131 uses the *smoothed* density
    hess_shift = 0.02          # lmc_laplace.py:186


class GaussMixL1(GaussMix):
    """prox_lmc.py:29-95: Gaussian-mixture likelihood with a Laplacian (l1) prior
    of scale ``alpha`` centred at ``mu`` (zero in the driver, prox_lmc.py:304)."""

    def __init__(self, mus, Sigmas, omegas, lamda, alpha, mu):
        super().__init__(mus, Sigmas, omegas)
        self.lamda, self.alpha, self.mu = lamda, alpha, mu

    def prior(self, th):                                   # prox_lmc.py:56-57
        return (self.alpha / 2) ** self.d * np.exp(
            -self.alpha * np.linalg.norm(th - self.mu, ord=1, axis=-1))

    def grad_env(self, th):                                # prox_lmc.py:114-115
        return (th - soft(th, self.lamda * self.alpha)) / self.lamda

    def pprox(self, x, thr, Q, t):                         # prox_lmc.py:165-173
        rho = 1 / np.linalg.norm(Q, ord=2)
        eps = max(min(1, rho) - 1e-5, 1e-9)
        eta = rho - eps
        w = np.zeros_like(x)
        for _ in range(t):
            u = x - Q @ w
            w = w + eta * u - eta * soft(w / eta + u, thr / eta)
        return u

    def grad_fb_env(self, th):                             # prox_lmc.py:192-193
        fwd = th - self.lamda * self.grad_U(th)
        return (np.identity(th.shape[0]) - self.lamda * self.hess_U(th)) @ (
            th - soft(fwd, self.lamda * self.alpha)) / self.lamda

    def bregman_prox(self, th, beta, g):                   # prox_lmc.py:226-233
        lo = beta * np.sinh(np.arcsinh(th / beta) - g)
        hi = beta * np.sinh(np.arcsinh(th / beta) + g)
        mid = np.sqrt(th ** 2 + beta ** 2) - beta
        p = np.where(th > beta * np.sinh(g), lo, mid)
        return np.where(th < beta * np.sinh(-g), hi, p)

    def grad_bm_env(self, th, beta):                       # prox_lmc.py:235-236
        return 1 / self.lamda * (th ** 2 + beta ** 2) ** (-.5) * (
            th - self.bregman_prox(th, beta, self.lamda * self.alpha))

    def mh_density(self, th):                              # prox_lmc.py:139-140
        return self.density(th) * self.prior(th)


# --------------------------------------------------------------------------- steps
# Each step maps (target, theta, xi[, u], params) -> proposal.  The drift helpers
# below are the reference's ``gd_update`` family.
def _drift(tg, th, g):
    return th - g * tg.grad_U(th)


def _q(tg, a, b, g, centre):
    """N(a; centre(b), 2g I) evaluated with scipy's frozen distribution, exactly as
    lmc.py:108-109 / prox_lmc.py:134-135 do."""
    return multivariate_normal(mean=centre(tg, b, g), cov=2 * g).pdf(a)


def _mh_ratio(tg, new, old, g, centre):
    dens = tg.mh_density(new) / tg.mh_density(old)
    return dens * (_q(tg, old, new, g, centre) / _q(tg, new, old, g, centre))


def _myula_centre(tg, th, g):
    return _drift(tg, th, g) + (-g * tg.grad_env(th))


def step_ula(tg, th, xi, g):
    return _drift(tg, th, g) + np.sqrt(2 * g) * xi


def step_pula(tg, th, xi, g, M):
    return th - g * M @ tg.grad_U(th) + np.sqrt(2 * g) * sqrtm(M) @ xi


def step_ihpula(tg, th, xi, g, n_comp):
    H = tg.hess_U(th)
    if n_comp > 1:
        e = np.linalg.eigvals(H)
        M = np.linalg.inv(H + (np.abs(min(e)) + tg.hess_shift) * np.identity(th.shape[0]))
    else:
        M = np.linalg.inv(H)
    return th - g * M @ tg.grad_U(th) + np.sqrt(2 * g) * sqrtm(M) @ xi


def step_mla(tg, th, xi, g, beta):
    dual = np.arcsinh(th / beta) - g * tg.grad_U(th) + np.sqrt(2 * g) * (th ** 2 + beta ** 2) ** (-.25) * xi
    return beta * np.sinh(dual)


def step_pgld(tg, th, xi, g):
    th = soft(th, tg.lamda * tg.alpha)
    return _drift(tg, th, g) + np.sqrt(2 * g) * xi


def step_myula(tg, th, xi, g):
    return _drift(tg, th, g) + (-g * tg.grad_env(th)) + np.sqrt(2 * g) * xi


def step_ppula(tg, th, xi, g, M, Q, t):
    pp = -g * np.linalg.inv(Q) @ (th - tg.pprox(th, tg.lamda, Q, t)) / tg.lamda
    return th - g * M @ tg.grad_U(th) + pp + np.sqrt(2 * g) * sqrtm(M) @ xi


def step_fbula(tg, th, xi, g):
    return th - g * tg.grad_fb_env(th) + np.sqrt(2 * g) * xi


def step_lbmumla(tg, th, xi, g, beta, sigma):
    # prox_lmc.py:251: ``sigma`` lands in the ``beta`` slot of grad_BM_env
    dual = (np.arcsinh(th / beta) + (-g * tg.grad_U(th)) + (-g * tg.grad_bm_env(th, sigma))
            + np.sqrt(2 * g) * (th ** 2 + beta ** 2) ** (-.25) * xi)
    return beta * np.sinh(dual)


UNADJUSTED = {
    "ula": step_ula, "pula": step_pula, "ihpula": step_ihpula, "mla": step_mla,
    "pgld": step_pgld, "myula": step_myula, "ppula": step_ppula, "fbula": step_fbula,
    "lbmumla": step_lbmumla,
}


def run_unadjusted(alg, tg, theta0, xi, g, *args):
    """Trajectory (K, d) of an unadjusted sampler driven by the given noise."""
    step = UNADJUSTED[alg]
    th = np.asarray(theta0, dtype=np.float64)
    out = np.empty((len(xi), th.shape[0]))
    for k in range(len(xi)):
        th = step(tg, th, xi[k], g, *args)
        out[k] = th
    return out


def run_adjusted(alg, tg, theta0, xi, u, g):
    """MALA (lmc.py:118-131, lmc_laplace.py:135-151) / MYMALA (prox_lmc.py:145-158).

    Returns (states (K, d) after every step, accept flags (K,), ratios (K,)); the
    reference's return value is ``states[accept], accept.sum()``.
    ``min(1, p)`` is Python's builtin: a NaN ratio yields 1, i.e. accept.
    """
    centre = _drift if alg == "mala" else _myula_centre
    th = np.asarray(theta0, dtype=np.float64)
    K = len(xi)
    states = np.empty((K, th.shape[0]))
    acc = np.zeros(K, dtype=bool)
    ratios = np.empty(K)
    for k in range(K):
        new = centre(tg, th, g) + np.sqrt(2 * g) * xi[k]
        p = _mh_ratio(tg, new, th, g, centre)
        ratios[k] = p
        if u[k] <= min(1, p):
            th = new
            acc[k] = True
        states[k] = th
    return states, acc, ratios
