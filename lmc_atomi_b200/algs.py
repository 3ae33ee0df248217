"""Drop-in for the reference's ``algs`` module (the image samplers of algs.py:295-587):
same function names, argument order, defaults, seed semantics and return shapes.  The
per-iteration Python loops of the reference are replaced by launches of the persistent
CUDA kernels behind include/lmc_b200.h; the operator objects (``proxf``, ``proxg``, ``A``)
are read for their parameters by duck typing (real pylops / pyproximal objects or the
stand-ins of ``compat``).  Unrecognised operator combinations raise NotImplementedError:
there is no CPU fallback.

Noise: the reference draws ``scipy.stats.multivariate_normal.rvs(size=x.shape,
random_state=default_rng(seed))`` every iteration (algs.py:431-433, 561-565), which is
``rng.standard_normal(P)`` bit for bit; the same stream is generated on the
host and fed to the kernels, so for a given seed the samples follow the reference's
trajectory (``philox=True`` switches to in-kernel counter-based noise instead).
"""
from __future__ import annotations

import time

import numpy as np
from numpy.random import default_rng

from . import _lib, imaging

__all__ = ["MoreauYosidaUnadjustedLangevin", "UnadjustedLangevinPrimalDual", "L2_ncvx_tv", "DeviceMetrics"]


class DeviceMetrics:
    """Collects, on the GPU, what the driver's per-iteration host callback computes
    (prox_lmc_deconv.py:128-133: cost f(x) + g(Kx), ||x - xtrue||, SNR, PSNR, MSE).  Pass an instance as
    ``device_metrics=`` instead of ``callback=``: the sampler then runs its iterations in large launches and
    reduces every sample to four sums on the device; after the run the lists ``cost``, ``err``, ``snr``,
    ``psnr``, ``mse`` hold one value per iteration (f = L2 data term; g = proxg.sigma * TV / l1)."""

    KEYS = ("cost", "err", "snr", "psnr", "mse")

    def __init__(self, xtrue):
        self.xtrue = np.asarray(xtrue, dtype=np.float64).ravel()
        for k in self.KEYS:
            setattr(self, k, [])

    def _extend(self, res):
        for k in self.KEYS:
            getattr(self, k).extend(np.asarray(res[k], dtype=np.float64).ravel().tolist())


CHUNK_BYTES = 1 << 30     # device staging for noise + samples per launch


def _kind(obj):
    return type(obj).__name__


def _require(cond, msg):
    if not cond:
        raise NotImplementedError(msg + " (no CUDA path for this combination and no CPU fallback by design)")


def _blur_of(proxf):
    Op = getattr(proxf, "Op", None)
    _require(Op is not None and hasattr(Op, "h") and hasattr(Op, "dims"),
             "proxf.Op must be a Convolve2D-like operator exposing .h and .dims")
    _require(getattr(proxf, "b", None) is not None, "proxf must carry data b")
    dims = tuple(int(d) for d in Op.dims)
    _require(len(dims) == 2, "2-D images")
    return dims, np.asarray(Op.h, dtype=np.float64)


def _engine(proxf, dtype):
    _require(_kind(proxf) in ("L2", "L2_ncvx_tv"),
             f"proxf of type {_kind(proxf)}; supported: pyproximal.L2, algs.L2_ncvx_tv")
    dims, h = _blur_of(proxf)
    return imaging.DeconvEngine(dims, h, np.asarray(proxf.b, dtype=np.float64), float(proxf.sigma), dtype=dtype)


def _linear_term(proxf, P):
    """alpha * q of pyproximal.L2 / algs.L2_ncvx_tv (algs.py:186-187, 226-227, 288-289) as a float64 (P,) array, or None."""
    q = getattr(proxf, "q", None)
    if q is None:
        return None
    q = np.asarray(q, dtype=np.float64).ravel()
    _require(q.size == P, "q must have one entry per pixel")
    return float(getattr(proxf, "alpha", 1.0)) * q


def _lsqr_limit(proxf, count):
    """``iter_lim`` of the LSQR solve inside ``proxf.prox``: ``niter`` itself, or ``niter(count)`` when it is a callable
    of the prox call count (already incremented for the call, as pyproximal's ``_increment_count`` does; algs.py:203-207)."""
    n = proxf.niter
    return int(n(count)) if callable(n) else int(n)


def _moreau_term(proxf, dims):
    """None for pyproximal.L2; the MC-TV / ME-TV term of an L2_ncvx_tv (algs.py:164-170)."""
    if _kind(proxf) != "L2_ncvx_tv":
        return None
    _require(tuple(int(d) for d in proxf.dims) == tuple(dims), "L2_ncvx_tv dims must equal the image dims")
    _require(getattr(proxf, "densesolver", None) is None, "dense solvers")
    if proxf.Op2 is not None:
        _require(_is_gradient(proxf.Op2, dims), "Op2 must be pylops.Gradient(dims, kind='forward', edge=False)")
        return imaging.MoreauTerm("mc", proxf.lamda, proxf.gamma, bool(proxf.isotropic))
    _require(bool(proxf.isotropic), "ME-TV with the 1-D TV of algs.py:170")
    _require(not callable(proxf.niter), "ME-TV with a callable niter (it also sets the inner TV's iteration count)")
    return imaging.MoreauTerm("me", proxf.lamda, proxf.gamma, True, int(proxf.niter), float(proxf.rtol))


NOISE_CHUNK_ITERS = 16    # iterations per launch when the reference's host noise stream is used (pipelined)


def _chunk(niter, P, itemsize, arrays):
    per_it = max(1, arrays * P * itemsize)
    return int(max(1, min(niter, CHUNK_BYTES // per_it)))


class _NoiseStream:
    """The reference's noise, ``default_rng(seed).standard_normal(P)`` once per iteration (algs.py:431-433,
    561-565), drawn in order by a worker thread one chunk ahead of the GPU: the host RNG (the slowest stage of a
    single-chain run) overlaps the kernels and the copies.  Same stream, same values."""

    def __init__(self, seed, P, chunks):
        import queue
        import threading
        self._q = queue.Queue(maxsize=2)
        rng = default_rng(seed)

        def work():
            for n in chunks:
                self._q.put(rng.standard_normal((1, n, P)))

        self._t = threading.Thread(target=work, daemon=True)
        self._t.start()

    def next(self):
        return self._q.get()


class _SuppliedNoise:
    """Validation mode: caller-supplied N(0,1) draws, (niter, P) host array or device tensor, consumed in order."""

    def __init__(self, noise, niter, P):
        self._z, self._pos = noise, 0
        if tuple(noise.shape) != (niter, P):
            raise ValueError(f"noise must have shape (niter, P) = {(niter, P)}, got {tuple(noise.shape)}")
        self._chunks = None

    def plan(self, chunks):
        self._chunks = list(chunks)
        return self

    def next(self):
        n = self._chunks.pop(0)
        out = self._z[self._pos:self._pos + n]
        self._pos += n
        return out[None] if isinstance(out, np.ndarray) else out.unsqueeze(0)


def _noise_source(noise, philox, seed, niter, P, chunks):
    if philox:
        return None
    if noise is not None:
        return _SuppliedNoise(noise, niter, P).plan(chunks)
    return _NoiseStream(seed, P, chunks)


class _RunningMean:
    """``reduce="mean"``: the posterior mean the driver takes of the returned samples (prox_lmc_deconv.py:474-475),
    accumulated on the device by the kernels (Welford), so the (niter, P) samples never cross PCIe."""

    def __init__(self, eng):
        self.t = eng.torch.zeros((1, eng.P), dtype=eng.tdt, device=eng.dev)

    def result(self):
        return self.t[0].double().cpu().numpy()


def _plan(niter, step):
    return [min(step, niter - d) for d in range(0, niter, step)]


def _show_line(iiter, niter):
    return iiter < 10 or niter - iiter < 10 or (niter >= 10 and iiter % (niter // 10) == 0)


def MoreauYosidaUnadjustedLangevin(proxf, proxg, x0, tau=None, gamma=.1, epsg=1., niter=10, seed=0,
                                   callback=None, show=False, dtype="float64", philox=False, device_metrics=None,
                                   noise=None, reduce=None):
    """MYULA (algs.py:477-587):
    x <- (1 - tau/gamma) x - tau grad f(x) + tau/gamma prox_{gamma epsg g}(x) + sqrt(2 tau) xi.
    Returns the (niter, P) float64 array of samples, as the reference does.

    Extensions (keyword only in spirit; the reference's positional order is unchanged): ``dtype`` arithmetic of
    the kernels; ``philox`` in-kernel noise instead of the reference's host stream; ``noise`` (niter, P) supplied
    N(0,1) draws (validation mode); ``device_metrics`` see DeviceMetrics; ``reduce="mean"`` returns the (P,)
    mean of the samples, accumulated on the device, instead of the samples."""
    _require(tau is not None, "tau=None (the reference itself fails on it at algs.py:569)")
    _require(np.asarray(epsg).size == 1, "vector epsg")
    _require(_kind(proxg) == "TV", f"proxg of type {_kind(proxg)}; supported: pyproximal.TV")
    eng = _engine(proxf, dtype)
    torch = eng.torch
    ncvx = _moreau_term(proxf, eng.dims)
    _require(tuple(int(d) for d in proxg.dims) == eng.dims, "TV dims must equal the image dims")
    x0 = np.asarray(x0, dtype=np.float64)
    P = eng.P
    # alpha q enters grad f when qgrad (algs.py:288-289): a constant extra gradient term of the kernels
    aq = _linear_term(proxf, P) if getattr(proxf, "qgrad", True) else None
    _require(aq is None or ncvx is None, "MYULA with both a non-convex term and a q term")
    fextra = None if aq is None else eng._buf(aq, 1)
    if show:   # banner / header / line formats verbatim from algs.py:545-560, 576-583 (output compatibility)
        tstart = time.time()
        print('Moreau--Yosida Unadjusted Langevin\n'
              '---------------------------------------------------------\n'
              'Proximal operator (f): %s\nProximal operator (g): %s\n'
              'tau = %s\tgamma=%10e\nepsg = %s\tniter = %d\n'
              % (type(proxf), type(proxg), str(tau), gamma, str(epsg), niter))
        print('   Itn       x[0]          f           g     J = f + eps*g')
    _require(reduce in (None, "mean"), f"reduce={reduce!r}")
    x = eng._buf(x0, 1)
    rmean = _RunningMean(eng) if reduce == "mean" else None
    out = np.empty((niter, P), dtype=np.float64) if rmean is None else None
    per_iter = callback is not None or show
    need_samples = rmean is None or per_iter or device_metrics is not None
    step = 1 if per_iter else _chunk(niter, P, x.element_size(), 2)
    if not philox and not per_iter:
        step = min(step, NOISE_CHUNK_ITERS)
    stream = _noise_source(noise, philox, seed, niter, P, _plan(niter, step))
    done = 0
    while done < niter:
        n = min(step, niter - done)
        nz = None if philox else stream.next()
        samples = torch.empty((1, n, P), dtype=eng.tdt, device=eng.dev) if need_samples else None
        keep = eng.myula(x, n, tau, gamma, float(proxg.sigma), int(proxg.niter), float(proxg.rtol),
                         epsg=float(epsg), seed=seed, noise=nz, samples=samples, iter_offset=done, ncvx=ncvx,
                         mean=None if rmean is None else rmean.t, stat_count0=done,
                         fgrad_extra=fextra, fgrad_scale=1.0 if fextra is not None else 0.0)
        if device_metrics is not None:
            device_metrics._extend(eng.sample_metrics(samples[0], device_metrics.xtrue, float(proxg.sigma), True))
        if out is not None:
            out[done:done + n] = samples[0].double().cpu().numpy()   # synchronises the stream
        else:
            torch.cuda.current_stream(eng.dev).synchronize()         # workspace / noise lifetime
        del keep
        if per_iter:
            xh = out[done] if out is not None else samples[0, 0].double().cpu().numpy()
            if callback is not None:
                callback(xh)
            if show and _show_line(done, niter):
                pf, pg = proxf(xh), proxg(xh)
                print('%6g  %12.5e  %10.3e  %10.3e  %10.3e' % (done + 1, xh[0], pf, pg, pf + epsg * pg))
        done += n
    if show:
        print('\nTotal time (s) = %.2f' % (time.time() - tstart))
        print('---------------------------------------------------------\n')
    return out if rmean is None else rmean.result()


def _is_gradient(A, dims):
    if _kind(A) not in ("Gradient", "VStack"):
        return False
    P = dims[0] * dims[1]
    if tuple(getattr(A, "shape", ())) != (2 * P, P):
        return False
    return getattr(A, "kind", "forward") == "forward" and not getattr(A, "edge", False)


def UnadjustedLangevinPrimalDual(proxf, proxg, A, x0, tau, mu, y0=None, z=None, theta=1., niter=10,
                                 seed=0, gfirst=True, callback=None, callbacky=False, returny=False,
                                 show=False, dtype="float64", philox=False, device_metrics=None, noise=None,
                                 reduce=None):
    """ULPDA (algs.py:295-474) for U(x) = f(x) + x^T z + g(Ax) with f = L2 (prox by warm-started
    LSQR), g = L21 / L1 (dual prox = projection), A = forward-difference Gradient.
    Returns (niter, P) samples [and (niter, 2P) dual samples when ``returny``].
    ``dtype``, ``philox``, ``device_metrics``, ``noise``, ``reduce``: as for MoreauYosidaUnadjustedLangevin
    (``reduce="mean"`` returns the (P,) mean of the x samples; not combined with ``returny``)."""
    eng = _engine(proxf, dtype)
    torch = eng.torch
    P = eng.P
    ncvx = _moreau_term(proxf, eng.dims)
    _require(_kind(proxg) in ("L21", "L1"), f"proxg of type {_kind(proxg)}; supported: pyproximal.L21, L1")
    iso = _kind(proxg) == "L21"
    _require(not iso or int(getattr(proxg, "ndim", 2)) == 2, "L21 with ndim != 2")
    _require(getattr(proxg, "g", None) is None, "L1 with a shift vector")
    _require(_is_gradient(A, eng.dims), "A must be pylops.Gradient(dims, kind='forward', edge=False)")
    x0 = np.asarray(x0, dtype=np.float64)
    # y = x + tau OpTb - tau alpha q (algs.py:226-227) with x = x - tau (A^T y + z): alpha q adds to z.  With a non-convex
    # term the Moreau pre-step sits between the two (algs.py:208-223), so the fold would move its evaluation point.
    aq = _linear_term(proxf, P)
    _require(aq is None or ncvx is None, "ULPDA with both a non-convex term and a q term")
    zk = z if aq is None else (aq if z is None else np.asarray(z, dtype=np.float64).ravel() + aq)
    # algs.py:401-408: scalars become constant arrays
    fixedtau, fixedmu = isinstance(tau, (int, float)), isinstance(mu, (int, float))
    tau = tau * np.ones(niter, dtype=x0.dtype) if fixedtau else np.asarray(tau, dtype=np.float64)
    mu = mu * np.ones(niter, dtype=x0.dtype) if fixedmu else np.asarray(mu, dtype=np.float64)
    constant = bool(np.all(tau == tau[0]) and np.all(mu == mu[0])) if niter > 0 else True
    if show:   # banner / header / line formats verbatim from algs.py:410-423, 459-470 (output compatibility)
        tstart = time.time()
        print('Unadjusted Langevin primal-dual: U(x) = f(x) + x^T z + g(Ax)\n'
              '---------------------------------------------------------\n'
              'Proximal operator (f): %s\nProximal operator (g): %s\nLinear operator (A): %s\n'
              'Additional vector (z): %s\ntau = %s\t\tmu = %s\ntheta = %.2f\t\tniter = %d\n' %
              (type(proxf), type(proxg), type(A), None if z is None else 'vector',
               str(tau[0]) if fixedtau else 'Variable', str(mu[0]) if fixedmu else 'Variable', theta, niter))
        print('   Itn       x[0]          f         g o A        z^x      U = f + g o A + z^x')
    x = eng._buf(x0, 1)
    xhat = x.clone()
    y = eng._buf(y0, 1, inner=2) if y0 is not None else torch.zeros((1, 2, P), dtype=eng.tdt, device=eng.dev)
    warm_on = bool(getattr(proxf, "warm", True))
    warm = eng._buf(getattr(proxf, "x0", None) if warm_on else None, 1)
    kw = getattr(proxf, "kwargs_solver", None) or {}
    _require(reduce in (None, "mean") and not (reduce and returny), f"reduce={reduce!r} with returny={returny}")
    rmean = _RunningMean(eng) if reduce == "mean" else None
    out = np.empty((niter, P), dtype=np.float64) if rmean is None else None
    outy = np.empty((niter, 2 * P), dtype=np.float64) if returny else None
    need_y = returny or (callback is not None and callbacky)
    count0 = int(getattr(proxf, "count", 0))
    per_iter = callback is not None or show or not constant or callable(proxf.niter)   # niter(count): one launch each
    need_samples = rmean is None or per_iter or device_metrics is not None
    step = 1 if per_iter else _chunk(niter, P, x.element_size(), 2 + (2 if need_y else 0))
    if not philox and not per_iter:
        step = min(step, NOISE_CHUNK_ITERS)
    stream = _noise_source(noise, philox, seed, niter, P, _plan(niter, step))
    done = 0
    while done < niter:
        n = min(step, niter - done)
        nz = None if philox else stream.next()
        samples = torch.empty((1, n, P), dtype=eng.tdt, device=eng.dev) if need_samples else None
        ys = torch.empty((1, n, 2, P), dtype=eng.tdt, device=eng.dev) if need_y else None
        keep = eng.ulpda(x, xhat, y, warm, n, float(tau[done]), float(mu[done]), float(theta), float(proxg.sigma),
                         iso, gfirst, _lsqr_limit(proxf, count0 + done + 1), warm_start=warm_on, atol=kw.get("atol", 1e-6),
                         btol=kw.get("btol", 1e-6), conlim=kw.get("conlim", 1e8), z=zk, seed=seed, noise=nz,
                         samples=samples, ysamples=ys, iter_offset=done, ncvx=ncvx,
                         mean=None if rmean is None else rmean.t, stat_count0=done)
        if device_metrics is not None:
            device_metrics._extend(eng.sample_metrics(samples[0], device_metrics.xtrue, float(proxg.sigma), iso))
        if out is not None:
            out[done:done + n] = samples[0].double().cpu().numpy()
        else:
            torch.cuda.current_stream(eng.dev).synchronize()
        if need_y:
            yh = ys[0].double().cpu().numpy().reshape(n, 2 * P)
            if returny:
                outy[done:done + n] = yh
        del keep
        if per_iter and (callback is not None or show):
            xh = out[done] if out is not None else samples[0, 0].double().cpu().numpy()
        if callback is not None:
            if callbacky:
                callback(xh, yh[0])
            else:
                callback(xh)
        if show and _show_line(done, niter):
            pf, pg = proxf(xh), proxg(A.matvec(xh))
            pf = 0. if type(pf) == bool else pf
            pg = 0. if type(pg) == bool else pg
            zx = 0. if z is None else float(np.dot(z, xh))
            print('%6g  %12.5e  %10.3e  %10.3e  %10.3e      %10.3e' % (done + 1, xh[0], pf, pg, zx, pf + pg + zx))
        done += n
    if warm_on and niter > 0:
        proxf.x0 = warm[0].double().cpu().numpy()     # pyproximal.L2 keeps the last solution as warm start
    if hasattr(proxf, "count"):
        proxf.count += niter
    if show:
        print('\nTotal time (s) = %.2f' % (time.time() - tstart))
        print('---------------------------------------------------------\n')
    if rmean is not None:
        return rmean.result()
    if not returny:
        return out
    return out, outy


class L2_ncvx_tv:
    """The reference's non-convex data term (algs.py:22-291):
    sigma/2 ||Op x - b||^2 - lamda * env_gamma(TV)(x), MC-TV when ``Op2`` (the Gradient) is given, ME-TV
    otherwise.  Same constructor as algs.py:129-131.  The samplers read its parameters; ``grad``, ``prox`` and
    ``__call__`` evaluate on the GPU with the same unit kernels the samplers use."""

    def __init__(self, dims, Op=None, Op2=None, b=None, q=None, sigma=1., alpha=1., lamda=1., gamma=.5,
                 qgrad=True, isotropic=False, niter=10, rtol=1e-4, x0=None, warm=True, densesolver=None,
                 kwargs_solver=None):
        self.dims, self.ndim = tuple(dims), len(dims)
        self.Op, self.Op2, self.b, self.q = Op, Op2, b, q
        self.sigma, self.alpha, self.lamda, self.gamma = sigma, alpha, lamda, gamma
        self.qgrad, self.isotropic, self.niter, self.rtol = qgrad, isotropic, niter, rtol
        self.x0, self.warm, self.densesolver, self.count = x0, warm, densesolver, 0
        self.kwargs_solver = {} if kwargs_solver is None else kwargs_solver
        self.hasgrad = True
        self._eng = None

    def _engine64(self):
        if self._eng is None:
            self._eng = _engine(self, "float64")
        return self._eng

    def _dev(self, x):
        eng = self._engine64()
        return eng.torch.as_tensor(np.asarray(x, dtype=np.float64)).to(eng.dev).reshape(1, -1).contiguous()

    def grad_moreau(self, x):
        """algs.py:270-282."""
        eng = self._engine64()
        out = imaging.moreau_apply(eng.dims, self._dev(x), _moreau_term(self, eng.dims))
        return out.cpu().numpy().reshape(np.shape(x))

    def grad(self, x):
        """sigma Op^H (Op x - b) - lamda grad_moreau(x)   (algs.py:270-291)."""
        eng = self._engine64()
        xd = self._dev(x)
        r = imaging.blur(eng.spec, xd) - eng.b
        g = (self.sigma * imaging.blur(eng.spec, r, adjoint=True)).cpu().numpy().reshape(np.shape(x))
        aq = _linear_term(self, eng.P)
        if aq is not None and self.qgrad:      # algs.py:288-289
            g = g + aq.reshape(np.shape(x))
        return g - self.lamda * self.grad_moreau(x)

    def prox(self, x, tau):
        """algs.py:201-267: x += tau lamda grad_moreau(x) IN PLACE (as the reference does), then the LSQR solve."""
        eng = self._engine64()
        torch = eng.torch
        self.count += 1
        term = _moreau_term(self, eng.dims)
        arg = imaging.moreau_apply(eng.dims, self._dev(x), term, tau=float(tau))
        if isinstance(x, np.ndarray) and x.dtype == np.float64:
            x[...] = arg.cpu().numpy().reshape(x.shape)
        aq = _linear_term(self, eng.P)
        if aq is not None:      # y = x + tau OpTb - tau alpha q (algs.py:226-227)
            arg = arg - float(tau) * self._dev(aq)
        warm = self._dev(self.x0).clone() if (self.x0 is not None and self.warm) else \
            torch.zeros((1, eng.P), dtype=torch.float64, device=eng.dev)
        kw = self.kwargs_solver
        out = eng.l2_prox(arg, tau, warm, iter_lim=_lsqr_limit(self, self.count), atol=kw.get("atol", 1e-6),
                          btol=kw.get("btol", 1e-6), conlim=kw.get("conlim", 1e8))
        out = out.cpu().numpy().reshape(np.shape(x))
        if self.warm:
            self.x0 = out
        return out

    def __call__(self, x):
        """f(x) = sigma/2 ||Op x - b||^2 - lamda env_gamma(x)   (algs.py:173-190)."""
        eng = self._engine64()
        torch = eng.torch
        xd = self._dev(x)
        r = imaging.blur(eng.spec, xd) - eng.b
        f = float(self.sigma / 2.0 * r.pow(2).sum().item())
        g = float(self.gamma)
        if self.Op2 is not None:
            d = imaging.grad(eng.dims, xd)[0]                    # (2,P)
            v = d.pow(2).sum(dim=0).sqrt() if self.isotropic else d.reshape(-1)
            mp = torch.sign(v) * torch.clamp(v.abs() - g, min=0.0)   # L1(1).prox(v, gamma)
            env = float((mp.abs().sum() + (v - mp).pow(2).sum() / (2 * g)).item())
        else:
            mp = imaging.tv_prox(eng.dims, xd, g, int(self.niter), float(self.rtol))
            dg = imaging.grad(eng.dims, mp)[0]
            env = float((dg.pow(2).sum(dim=0).sqrt().sum() + (xd - mp).pow(2).sum() / (2 * g)).item())
        aq = _linear_term(self, eng.P)
        if aq is not None:      # algs.py:186-187
            f += float(np.dot(aq, np.asarray(x, dtype=np.float64).ravel()))
        return f - self.lamda * env
