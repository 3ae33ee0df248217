"""TEST INFRASTRUCTURE ONLY -- CPU oracle for the image (deconvolution) samplers and the MAP branch.

The per-iteration arithmetic of the reference's image path lives in ``pylops`` (Convolve2D, Gradient) and
``pyproximal`` (L2, L1, L21, TV, AdaptivePrimalDual), which are neither vendored in /root/reference, nor pinned to
a version (README.md:5 ``pip install -U``), nor installed or installable in this image.  The classes below restate
their published behaviour, each rule one small named function.  PARITY PINNED against results the reference itself
holds: the per-sample SNR / PSNR / MSE curves of its committed figures
fig/fig_prox_lmc_deconv_einstein_{MYULA,ULPDA,MAP}_1000_snr_psnr_mse.pdf, digitised by oracle/digitise_figs.py into
tests/golden/deconv_curves.npz (2.2e-7 dB per vertex) and reproduced by this module on fig/einstein.png with the
reference's noise stream to <= 1e-7 dB (tests/test_oracle_pinned.py; the CUDA path: tests/test_pinned_gpu.py).

* the sampler loops and ``L2_ncvx_tv`` follow the reference's own source line by line (algs.py:22-291, 295-474,
  477-587);
* the LSQR solve is executed by ``scipy.sparse.linalg.lsqr`` itself -- the very function the reference calls
  (algs.py:250, pyproximal ``L2.prox``) -- and the blur by ``scipy.signal.convolve/correlate`` (what ``Convolve2D``
  reduces to);
* operator identities (adjoint dot-tests, Moreau identity, prox optimality) are checked in
  tests/test_oracle_deconv.py.

Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s cpu_baseline / ``--impl reference`` leg may import
this module; the product path never does.
"""
from __future__ import annotations

import numpy as np
from numpy.random import default_rng
from scipy.signal import convolve as _sp_convolve, correlate as _sp_correlate
from scipy.sparse.linalg import LinearOperator, lsqr as sp_lsqr


# --------------------------------------------------------------------------- I1 blur
def pad_kernel_like_pylops(h, offset):
    """pylops ``ConvolveND.__init__`` padding rule: the filter centre is
    moved to ``offset`` by zero padding; an even-sized axis gets one extra trailing
    zero so the padded filter is odd."""
    h = np.asarray(h, dtype=np.float64)
    off = 2 * (np.array(h.shape, dtype=int) // 2 - np.array(offset, dtype=int))
    pad = [(0, 0)] * h.ndim
    for ax, nh in enumerate(h.shape):
        if nh % 2 == 0:
            off[ax] -= 1
        if off[ax] != 0:
            pad[ax] = (off[ax] if off[ax] > 0 else 0, -off[ax] if off[ax] < 0 else 0)
    return np.pad(h, pad, mode="constant")


class Convolve2D:
    """``pylops.signalprocessing.Convolve2D(dims, h, offset)`` (prox_lmc_deconv.py:58-69):
    matvec = convolve(x, h_padded, 'same'), rmatvec = correlate(x, h_padded, 'same'),
    zero boundary, FFT method."""

    def __init__(self, dims, h, offset=(0, 0), method="fft"):
        self.dims = tuple(dims)
        self.h_in = np.asarray(h, dtype=np.float64)
        self.offset = tuple(int(o) for o in offset)
        self.h = pad_kernel_like_pylops(self.h_in, self.offset)
        self.method = method
        n = int(np.prod(self.dims))
        self.shape = (n, n)
        self.dtype = np.dtype("float64")
        self.explicit = False

    def matvec(self, x):
        return _sp_convolve(x.reshape(self.dims), self.h, mode="same", method=self.method).ravel()

    def rmatvec(self, x):
        return _sp_correlate(x.reshape(self.dims), self.h, mode="same", method=self.method).ravel()

    def __mul__(self, x):
        x = np.asarray(x, dtype=np.float64)
        y = self.matvec(x.ravel())
        return y.reshape(x.shape) if x.ndim == 2 else y

    __matmul__ = __mul__


def box_window(k, offset=None):
    """Window [lo, hi] (inclusive, relative to the output pixel) that a k x k box filter
    with pylops offset k//2 covers along one axis: forward (matvec) and adjoint."""
    hp = pad_kernel_like_pylops(np.ones((k, k)), (k // 2, k // 2) if offset is None else offset)
    n = hp.shape[0]
    c = (n - 1) // 2            # scipy 'same' centring of an odd filter
    nz = np.nonzero(hp[:, 0])[0]
    fwd = (c - nz.max(), c - nz.min())       # out[i] = sum_a hp[a] x[i + c - a]
    adj = (nz.min() - c, nz.max() - c)       # out[i] = sum_a hp[a] x[i + a - c]
    return fwd, adj


# --------------------------------------------------------------------------- I2 gradient
def grad2d(x):
    """Forward differences, unit sampling, last row/col zero (pylops Gradient
    kind='forward', edge=False; also UNLocBoX ``gradient_op``)."""
    dx = np.zeros_like(x)
    dy = np.zeros_like(x)
    dx[:-1, :] = x[1:, :] - x[:-1, :]
    dy[:, :-1] = x[:, 1:] - x[:, :-1]
    return dx, dy


def grad2d_adj(dx, dy):
    """Adjoint of :func:`grad2d`: out[0]=-v[0]; out[i]=v[i-1]-v[i]; out[last]=v[last-1]."""
    out = np.zeros_like(dx)
    out[:-1, :] -= dx[:-1, :]
    out[1:, :] += dx[:-1, :]
    out[:, :-1] -= dy[:, :-1]
    out[:, 1:] += dy[:, :-1]
    return out


def div2d(dx, dy):
    """UNLocBoX ``div_op`` = -grad^T."""
    return -grad2d_adj(dx, dy)


class Gradient:
    """``pylops.Gradient(dims, sampling=1, edge=False, kind='forward')``
    (prox_lmc_deconv.py:98): (P,) -> (2P,) with the axis-0 block first."""

    def __init__(self, dims):
        self.dims = tuple(dims)
        n = int(np.prod(self.dims))
        self.shape = (2 * n, n)
        self.dtype = np.dtype("float64")

    def matvec(self, x):
        dx, dy = grad2d(x.reshape(self.dims))
        return np.concatenate([dx.ravel(), dy.ravel()])

    def rmatvec(self, v):
        n = self.shape[1]
        return grad2d_adj(v[:n].reshape(self.dims), v[n:].reshape(self.dims)).ravel()


class Identity:
    def __init__(self, n):
        self.shape = (n, n)

    def matvec(self, x):
        return x

    def rmatvec(self, x):
        return x


# --------------------------------------------------------------------------- prox operators
def soft_threshold(x, thresh):
    """pyproximal ``_softthreshold`` (real case)."""
    return np.maximum(np.abs(x) - thresh, 0.) * np.sign(x)


class ProxOperator:
    """Base: ``proxdual`` by the Moreau identity, as pyproximal.ProxOperator does."""

    def proxdual(self, x, tau):
        return x - tau * self.prox(x / tau, 1. / tau)


class L1(ProxOperator):
    """``pyproximal.L1(sigma)`` (prox_lmc_deconv.py:119): anisotropic term."""

    def __init__(self, sigma=1.):
        self.sigma = sigma

    def __call__(self, x):
        return self.sigma * np.sum(np.abs(x))

    def prox(self, x, tau):
        return soft_threshold(x, tau * self.sigma)


class L21(ProxOperator):
    """``pyproximal.L21(ndim, sigma)`` (prox_lmc_deconv.py:116): group soft threshold over
    the ``ndim`` stacked blocks."""

    def __init__(self, ndim, sigma=1.):
        self.ndim, self.sigma = ndim, sigma

    def __call__(self, x):
        x = x.reshape(self.ndim, len(x) // self.ndim)
        return self.sigma * np.sum(np.sqrt(np.sum(x ** 2, axis=0)))

    def prox(self, x, tau):
        x = x.reshape(self.ndim, len(x) // self.ndim)
        aux = np.sqrt(np.sum(x ** 2, axis=0))
        aux = np.vstack([aux] * self.ndim).ravel()
        return (1 - (self.sigma * tau) / np.maximum(aux, self.sigma * tau)) * x.ravel()


TV_EXTRA_PASS = 1     # `while iter <= niter`: niter + 1 passes (pinned, see tv_prox_2d)


def tv_prox_2d(x2d, gamma, niter, rtol, trace=None, sigma=1.):
    """Isotropic TV prox by FISTA on the dual (UNLocBoX ``prox_tv`` as carried by
    pyproximal.TV).  Returns (sol, executed dual updates).

    PINNED against the reference's own figures (oracle/digitise_figs.py ->
    tests/golden/deconv_curves.npz, gate in tests/test_oracle_pinned.py): the per-sample
    SNR/PSNR/MSE curves of the einstein MYULA run agree with this recipe to the 2e-7 dB
    resolution of the plot, and disagree (drift to 1e-5 dB within 25 samples) when the
    objective below leaves out ``sigma``.

    Rules kept:
    * the loop is ``iter = 0; while iter <= niter`` -> at most niter + 1 passes;
    * obj = 1/2 ||x - sol||^2 + gamma * TV.__call__(sol) and ``TV.__call__`` carries the
      operator's own ``sigma``, so the weight of sum|grad sol| is gamma * sigma
      (gamma = sigma * tau);
    * t = (1 + sqrt(4 t_old^2)) / 2 (no "1 +" under the root, so t = t_old + 1/2);
    * the value returned is ``sol`` from the START of the last executed pass; the
      relative-change test uses obj from that same start."""
    x = x2d
    rr = np.zeros_like(x)
    ss = np.zeros_like(x)
    pold = np.zeros_like(x)
    qold = np.zeros_like(x)
    told, prev_obj = 1., 0.
    sol = x
    nupd = 0
    for _ in range(niter + TV_EXTRA_PASS):
        sol = x - gamma * div2d(rr, ss)
        dx, dy = grad2d(sol)
        obj = 0.5 * np.linalg.norm(x - sol) ** 2 + gamma * (sigma * np.sum(np.sqrt(dx ** 2 + dy ** 2)))
        rel_obj = np.abs(obj - prev_obj) / obj if obj > 1e-10 else 2.
        prev_obj = obj
        if trace is not None:
            trace.append(obj)
        if rel_obj < rtol:
            break
        r = rr - dx / (8. * gamma)
        s = ss - dy / (8. * gamma)
        w = np.maximum(1., np.sqrt(r ** 2 + s ** 2))
        p = r / w
        q = s / w
        t = (1. + np.sqrt(4. * told ** 2)) / 2.
        rr = p + (told - 1.) / t * (p - pold)
        ss = q + (told - 1.) / t * (q - qold)
        pold, qold, told = p, q, t
        nupd += 1
    return sol, nupd


class TV(ProxOperator):
    """``pyproximal.TV(dims, sigma, niter=10, rtol=1e-4)`` (prox_lmc_deconv.py:122)."""

    def __init__(self, dims, sigma=1., niter=10, rtol=1e-4):
        self.dims, self.sigma, self.niter, self.rtol = tuple(dims), sigma, niter, rtol
        self.ndim = len(self.dims)
        self.count = 0
        self.updates = []        # executed dual updates per prox call (roofline accounting)

    def __call__(self, x):
        dx, dy = grad2d(x.reshape(self.dims))
        return self.sigma * np.sum(np.sqrt(dx ** 2 + dy ** 2))

    def prox(self, x, tau):
        self.count += 1
        sol, n = tv_prox_2d(x.reshape(self.dims), self.sigma * tau, self.niter, self.rtol, sigma=self.sigma)
        self.updates.append(n)
        return sol.ravel()


class L2(ProxOperator):
    """``pyproximal.L2(Op, b, sigma, niter, warm=True)`` (prox_lmc_deconv.py:101-103;
    structure mirrored in the reference's own copy algs.py:158-162,224-256,283-284)."""

    def __init__(self, Op=None, b=None, sigma=1., niter=10, x0=None, warm=True, kwargs_solver=None, q=None, alpha=1.,
                 qgrad=True):
        self.Op, self.b, self.sigma, self.niter = Op, b, sigma, niter
        self.q, self.alpha, self.qgrad = q, alpha, qgrad      # alpha q^T x (algs.py:186-187, 226-227, 288-289)
        self.x0, self.warm = x0, warm
        self.count = 0
        self.kwargs_solver = {} if kwargs_solver is None else kwargs_solver
        self.OpTb = self.sigma * self.Op.rmatvec(self.b)
        self.lsqr_itn = []       # LSQR iterations per prox call (roofline accounting)

    def __call__(self, x):
        f = (self.sigma / 2.) * (np.linalg.norm(self.Op.matvec(x) - self.b) ** 2)
        if self.q is not None:
            f += self.alpha * np.dot(self.q, x)
        return f

    def grad(self, x):
        g = self.sigma * self.Op.rmatvec(self.Op.matvec(x) - self.b)
        if self.q is not None and self.qgrad:
            g += self.alpha * self.q
        return g

    def _op1(self, tau):
        c = float(tau * self.sigma)
        n = self.Op.shape[1]
        mv = lambda v: v + c * self.Op.rmatvec(self.Op.matvec(v))   # Identity + c * (Op.H * Op)
        return LinearOperator((n, n), matvec=mv, rmatvec=mv, dtype=np.float64)

    def _solve(self, y, tau):
        niter = self.niter if isinstance(self.niter, int) else self.niter(self.count)   # algs.py:203-207
        out = sp_lsqr(self._op1(tau), y, iter_lim=niter, x0=self.x0, **self.kwargs_solver)
        self.lsqr_itn.append(out[2])
        self.last_istop = out[1]
        return out[0]

    def prox(self, x, tau):
        self.count += 1
        y = x + tau * self.OpTb
        if self.q is not None:
            y -= tau * self.alpha * self.q
        x = self._solve(y, tau)
        if self.warm:
            self.x0 = x
        return x


class L2_ncvx_tv(L2):
    """``algs.L2_ncvx_tv`` (algs.py:22-291) for the two configurations the driver uses
    (prox_lmc_deconv.py:106-113): MC-TV (Op2=Gradient, isotropic) and ME-TV (Op2=None,
    isotropic, inner TV(dims, 1., niter, rtol))."""

    def __init__(self, dims, Op=None, Op2=None, b=None, sigma=1., lamda=1., gamma=.5,
                 isotropic=False, niter=10, rtol=1e-4, x0=None, warm=True, q=None, alpha=1., qgrad=True):
        super().__init__(Op=Op, b=b, sigma=sigma, niter=niter, x0=x0, warm=warm, q=q, alpha=alpha, qgrad=qgrad)
        self.dims, self.ndim = tuple(dims), len(dims)
        self.Op2, self.lamda, self.gamma, self.isotropic, self.rtol = Op2, lamda, gamma, isotropic, rtol
        if Op2 is not None:
            self.g_gamma = L1(1.)                                  # algs.py:166
        else:
            if not isotropic:
                raise NotImplementedError("1-D TV variant (algs.py:170) is not used by the driver")
            self.g_gamma = TV(dims, 1., niter, rtol)               # algs.py:169

    def grad_moreau(self, x):
        """algs.py:271-282."""
        if self.Op2 is not None:
            Op2x = self.Op2.matvec(x)
            if self.isotropic:
                e = np.linalg.norm(Op2x.reshape(self.ndim, len(Op2x) // self.ndim), axis=0)
                e = np.where(e != 0, e, 1e-9)
                return self.Op2.rmatvec(np.minimum(1 / self.gamma, np.tile(1 / e, 2)) * Op2x)
            return self.Op2.rmatvec(Op2x - self.g_gamma.prox(Op2x, self.gamma)) / self.gamma
        return (x - self.g_gamma.prox(x, self.gamma)) / self.gamma

    def __call__(self, x):
        """algs.py:173-190 (isotropic branch applies L1 to the pixel norms)."""
        Op2x = self.Op2.matvec(x) if self.Op2 is not None else x
        if self.Op2 is not None and self.isotropic:
            Op2x = Op2x.reshape(self.ndim, len(Op2x) // self.ndim)
            Op2x = np.sqrt(np.sum(Op2x ** 2, axis=0))
        mp = self.g_gamma.prox(Op2x, self.gamma)
        env = self.g_gamma(mp) + np.linalg.norm(Op2x - mp) ** 2 / (2 * self.gamma)
        f = (self.sigma / 2.) * (np.linalg.norm(self.Op.matvec(x) - self.b) ** 2)
        if self.q is not None:
            f += self.alpha * np.dot(self.q, x)
        return f - self.lamda * env

    def grad(self, x):
        """algs.py:270-291."""
        g = self.sigma * self.Op.rmatvec(self.Op.matvec(x) - self.b)
        if self.q is not None and self.qgrad:
            g += self.alpha * self.q
        return g - self.lamda * self.grad_moreau(x)

    def prox(self, x, tau):
        """algs.py:201-267: x += tau*lamda*grad_moreau(x) IN PLACE, then the L2 solve."""
        self.count += 1
        x += tau * self.lamda * self.grad_moreau(x)
        y = x + tau * self.OpTb
        if self.q is not None:
            y -= tau * self.alpha * self.q
        x = self._solve(y, tau)
        if self.warm:
            self.x0 = x
        return x


# --------------------------------------------------------------------------- I7 noise
def image_noise(seed, niter, n):
    """The reference's per-iteration draw ``scipy.stats.multivariate_normal.rvs(size=x.shape,
    random_state=rng)`` with ``rng = default_rng(seed)`` (algs.py:431-433, 561-565) equals
    ``rng.standard_normal(n)`` bit for bit (checked in tests/test_oracle_deconv.py)."""
    rng = default_rng(seed)
    return rng.standard_normal((niter, n))


# --------------------------------------------------------------------------- I8 MYULA
def MoreauYosidaUnadjustedLangevin(proxf, proxg, x0, tau=None, gamma=.1, epsg=1., niter=10, seed=0,
                                   callback=None, show=False, noise=None):
    """algs.py:477-587.  ``noise`` (niter, P) overrides the seeded stream (validation)."""
    x = x0.copy()
    x_samples = []
    rng = default_rng(seed)
    for iiter in range(niter):
        xi = rng.standard_normal(x.shape) if noise is None else noise[iiter]
        # algs.py:569
        x = (1 - tau / gamma) * x - tau * proxf.grad(x) + tau / gamma * proxg.prox(x, epsg * gamma) \
            + np.sqrt(2 * tau) * xi
        x_samples.append(x)
        if callback is not None:
            callback(x)
    return np.array(x_samples)


# --------------------------------------------------------------------------- I9 ULPDA
def UnadjustedLangevinPrimalDual(proxf, proxg, A, x0, tau, mu, y0=None, z=None, theta=1., niter=10,
                                 seed=0, gfirst=True, callback=None, callbacky=False, returny=False,
                                 show=False, noise=None):
    """algs.py:295-474."""
    if isinstance(tau, (int, float)):
        tau = tau * np.ones(niter, dtype=x0.dtype)
    if isinstance(mu, (int, float)):
        mu = mu * np.ones(niter, dtype=x0.dtype)
    x = x0.copy()
    xhat = x.copy()
    y = y0.copy() if y0 is not None else np.zeros(A.shape[0], dtype=x.dtype)
    x_samples, y_samples = [], []
    rng = default_rng(seed)
    for iiter in range(niter):
        xi = rng.standard_normal(x.shape) if noise is None else noise[iiter]
        xold = x.copy()
        if gfirst:
            y = proxg.proxdual(y + mu[iiter] * A.matvec(xhat), mu[iiter])
            ATy = A.rmatvec(y)
            if z is not None:
                ATy += z
            x = proxf.prox(x - tau[iiter] * ATy, tau[iiter]) + np.sqrt(2 * tau[iiter]) * xi
            xhat = x + theta * (x - xold)
        else:
            ATy = A.rmatvec(y)
            if z is not None:
                ATy += z
            x = proxf.prox(x - tau[iiter] * ATy, tau[iiter]) + np.sqrt(2 * tau[iiter]) * xi
            xhat = x + theta * (x - xold)
            y = proxg.proxdual(y + mu[iiter] * A.matvec(xhat), mu[iiter])
        x_samples.append(x)
        if returny:
            y_samples.append(y)
        if callback is not None:
            if callbacky:
                callback(x, y)
            else:
                callback(x)
    if not returny:
        return np.array(x_samples)
    return np.array(x_samples), np.array(y_samples)


# --------------------------------------------------------------------------- MAP (adaptive PDHG)
def AdaptivePrimalDual(proxf, proxg, A, x0, tau, mu, alpha=0.5, eta=0.95, s=1., delta=1.5, z=None, niter=10,
                       tol=1e-10, callback=None, norm_ord=2):
    """``pyproximal.optimization.primaldual.AdaptivePrimalDual`` as the driver calls it (prox_lmc_deconv.py:138-445):
    adaptive PDHG of Goldstein et al.  Restated from the published algorithm (pyproximal is absent) and PINNED
    against the reference's MAP figure (tests/test_oracle_pinned.py): with l2-norm residuals the per-iteration
    SNR / PSNR / MSE of model M1 agree with fig_prox_lmc_deconv_einstein_MAP_1000_snr_psnr_mse to the plot's
    resolution through 100+ iterations and several step-size adaptations; ``norm_ord=1`` (the other plausible
    reading) leaves the curve at the first adaptation.  Returns (x, steps)."""
    x = x0.copy()
    y = np.zeros(A.shape[0], dtype=x.dtype)
    ATy = np.zeros_like(x)
    Ax = np.zeros_like(y)
    steps = []
    for _ in range(niter):
        xold, yold, ATyold, Axold = x, y, ATy, Ax
        x = proxf.prox(x - tau * (ATy + z if z is not None else ATy), tau)
        Ax = A.matvec(x)
        y = proxg.proxdual(y + mu * (2 * Ax - Axold), mu)
        ATy = A.rmatvec(y)
        p = np.linalg.norm((xold - x) / tau - (ATyold - ATy), norm_ord)
        d = np.linalg.norm((yold - y) / mu - (Axold - Ax), norm_ord)
        if p > s * d * delta:
            tau, mu, alpha = tau / (1 - alpha), mu * (1 - alpha), alpha * eta
        elif p < s * d / delta:
            tau, mu, alpha = tau * (1 - alpha), mu / (1 - alpha), alpha * eta
        steps.append((tau, mu, alpha))
        if callback is not None:
            callback(x)
        if max(p, d) < tol:
            break
    return x, np.array(steps)


# --------------------------------------------------------------------------- driver pieces
def signal_noise_ratio(image_true, image_test):
    """prox_lmc_deconv.py:35-36."""
    return 20 * np.log10(np.linalg.norm(image_true) / np.linalg.norm(image_test - image_true))


def psnr(image_true, image_test, data_range=255.):
    """skimage.metrics.peak_signal_noise_ratio for a uint8 ground truth (data_range 255)."""
    err = np.mean((np.asarray(image_true, dtype=np.float64) - np.asarray(image_test, dtype=np.float64)) ** 2)
    return 10 * np.log10(data_range ** 2 / err)


def mse(image_true, image_test):
    return np.mean((np.asarray(image_true, dtype=np.float64) - np.asarray(image_test, dtype=np.float64)) ** 2)


def synthetic_image(ny, nx, seed=1234):
    """Deterministic piecewise-smooth test image in [0, 255] (uint8-valued), used where the
    reference's camera image (skimage, absent) would be: blocks, a disc, a ramp and texture."""
    yy, xx = np.mgrid[0:ny, 0:nx]
    img = 40. + 60. * (xx / max(nx - 1, 1))
    img[(yy > ny // 8) & (yy < ny // 2) & (xx > nx // 6) & (xx < nx // 2)] = 200.
    img[(yy - 0.65 * ny) ** 2 + (xx - 0.7 * nx) ** 2 < (0.18 * min(ny, nx)) ** 2] = 20.
    img[(yy > 0.6 * ny) & (xx < 0.4 * nx)] += 50. * np.sin(xx[(yy > 0.6 * ny) & (xx < 0.4 * nx)] * 0.35)
    rng = default_rng(seed)
    img += rng.normal(0, 2., size=img.shape)
    return np.clip(np.rint(img), 0, 255).astype(np.uint8)


def build_problem(img, sigma=0.75, tau=0.3, seed=0, kernel=5, niter_l2=50, niter_tv=10,
                  gamma_mc=15., gamma_me=15.):
    """Operator construction of prox_lmc_deconv.py:52-125 for one blur size.  Returns a dict
    with the blurred observation, operators and the step sizes of :88-94."""
    img = np.asarray(img)
    ny, nx = img.shape
    rng = default_rng(seed)
    h5 = np.ones((5, 5)) / 25.
    H5 = Convolve2D((ny, nx), h=h5, offset=(2, 2))
    y = H5 * img + rng.normal(0, sigma, size=(ny, nx))            # prox_lmc_deconv.py:59
    hk = np.ones((kernel, kernel))
    hk /= hk.sum()
    Hk = Convolve2D((ny, nx), h=hk, offset=(kernel // 2, kernel // 2))
    L = 1. / sigma ** 2
    Gop = Gradient((ny, nx))
    out = dict(img=img, y=y, H=Hk, Gop=Gop, dims=(ny, nx), tau0=0.95 / L, mu0=1.,
               gamma_myula=1. / L, tau_myula=0.2 / L, x0=np.zeros(ny * nx))
    out["l2"] = lambda: L2(Op=Hk, b=y.ravel(), sigma=1 / sigma ** 2, niter=50, warm=True)
    out["l2_mc"] = lambda: L2_ncvx_tv(dims=(ny, nx), Op=Hk, Op2=Gop, b=y.ravel(), sigma=1 / sigma ** 2,
                                      lamda=tau, gamma=gamma_mc, isotropic=True, niter=niter_l2, warm=True)
    out["l2_me"] = lambda: L2_ncvx_tv(dims=(ny, nx), Op=Hk, b=y.ravel(), sigma=1 / sigma ** 2, lamda=tau,
                                      gamma=gamma_me, isotropic=True, niter=niter_l2, warm=True)
    out["l1iso"] = lambda: L21(ndim=2, sigma=tau)
    out["l1"] = lambda: L1(sigma=tau)
    out["tv"] = lambda: TV(dims=(ny, nx), sigma=tau, niter=niter_tv)
    return out
