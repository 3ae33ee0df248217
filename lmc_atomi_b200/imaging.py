"""Host side of the image-chain path (MYULA / ULPDA deconvolution): descriptors, device
buffers (PyTorch tensors used only as buffers), launches through the C ABI.

``DeconvEngine`` is what both the reference-named functions in ``algs.py`` (one chain,
the reference's numpy noise stream, every sample returned) and the many-chain
production runs (``run_myula`` / ``run_ulpda`` with in-kernel Philox noise and running
posterior statistics) are built on.
"""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import _lib


# --------------------------------------------------------------------------- operators
def pad_filter(h, offset):
    """Zero padding that ``pylops.signalprocessing.ConvolveND`` applies to the filter so
    that its centre sits at ``offset`` and the padded filter is odd."""
    h = np.asarray(h, dtype=np.float64)
    off = 2 * (np.array(h.shape, dtype=int) // 2 - np.array(offset, dtype=int))
    pad = [(0, 0)] * h.ndim
    for ax, nh in enumerate(h.shape):
        if nh % 2 == 0:
            off[ax] -= 1
        if off[ax] != 0:
            pad[ax] = (int(off[ax]) if off[ax] > 0 else 0, int(-off[ax]) if off[ax] < 0 else 0)
    return np.pad(h, pad, mode="constant")


def separable_taps(h_padded):
    """Forward taps of ``convolve(x, h_padded, mode='same')`` per axis for a separable
    (rank-1) filter: returns (wy, lo_y, wx, lo_x) with out[i] = sum_t w[t] x[i + lo + t]."""
    hp = np.asarray(h_padded, dtype=np.float64)
    if hp.ndim != 2:
        raise NotImplementedError("the CUDA blur covers 2-D filters")
    rows, cols, tot = hp.sum(axis=1), hp.sum(axis=0), hp.sum()
    if tot != 0:
        ay, ax = rows / tot, cols
    else:
        u, s, vt = np.linalg.svd(hp)
        ay, ax = u[:, 0] * s[0], vt[0]
    if not np.allclose(np.outer(ay, ax), hp, rtol=1e-12, atol=1e-14 * max(1.0, np.abs(hp).max())):
        raise NotImplementedError("the CUDA blur covers separable (rank-1) filters; this one is not")

    def axis_taps(a):
        n = len(a)
        c = (n - 1) // 2   # scipy 'same' centring: out[i] = sum_a h[a] x[i + c - a]
        nz = np.nonzero(a)[0]
        if len(nz) == 0:
            return np.zeros(1), 0
        lo, hi = c - nz.max(), c - nz.min()
        w = np.array([a[c - t] for t in range(lo, hi + 1)])
        return w, int(lo)

    wy, lo_y = axis_taps(ay)
    wx, lo_x = axis_taps(ax)
    for w, lo in ((wy, lo_y), (wx, lo_x)):
        if len(w) > 7 or lo > 0 or lo < -(len(w) - 1):
            raise NotImplementedError("blur window outside the kernel's range (<= 7 taps containing the pixel)")
    return wy, lo_y, wx, lo_x


def _dt(dtype):
    return _lib.F64 if dtype == "float64" else _lib.F32


def _tdt(torch, dtype):
    return torch.float64 if dtype == "float64" else torch.float32


class BlurSpec:
    """Geometry + filter of the data term (what ``Convolve2D`` carries)."""

    def __init__(self, dims, h_padded):
        self.dims = (int(dims[0]), int(dims[1]))
        self.wy, self.lo_y, self.wx, self.lo_x = separable_taps(h_padded)

    def fill(self, pb: _lib.ImgProblem):
        pb.ny, pb.nx = self.dims
        pb.ntap_y, pb.ntap_x = len(self.wy), len(self.wx)
        pb.lo_y, pb.lo_x = self.lo_y, self.lo_x
        for t, v in enumerate(self.wy):
            pb.wy[t] = float(v)
        for t, v in enumerate(self.wx):
            pb.wx[t] = float(v)


def _stream(torch, dev):
    return C.c_void_p(torch.cuda.current_stream(dev).cuda_stream)


# --------------------------------------------------------------------------- unit operators
def blur(spec: BlurSpec, x, adjoint=False):
    """Convolve2D matvec / rmatvec on device tensors (C,P) or (P,)."""
    torch = _lib.require_cuda()
    lib = _lib.load()
    x2 = x.reshape(-1, spec.dims[0] * spec.dims[1]).contiguous()
    out = torch.empty_like(x2)
    pb = _lib.ImgProblem()
    spec.fill(pb)
    with torch.cuda.device(x2.device):
        rc = lib.lmc_img_blur(C.byref(pb), _dt("float64" if x2.dtype == torch.float64 else "float32"),
                              1 if adjoint else 0, x2.shape[0], x2.data_ptr(), out.data_ptr(),
                              _stream(torch, x2.device))
    _lib.check(rc, "lmc_img_blur")
    return out.reshape(x.shape)


def grad(dims, x, adjoint=False):
    """pylops.Gradient matvec ((C,P)->(C,2,P)) / rmatvec on device tensors."""
    torch = _lib.require_cuda()
    lib = _lib.load()
    P = dims[0] * dims[1]
    if not adjoint:
        x2 = x.reshape(-1, P).contiguous()
        out = torch.empty((x2.shape[0], 2, P), dtype=x2.dtype, device=x2.device)
    else:
        x2 = x.reshape(-1, 2, P).contiguous()
        out = torch.empty((x2.shape[0], P), dtype=x2.dtype, device=x2.device)
    with torch.cuda.device(x2.device):
        rc = lib.lmc_img_grad(_dt("float64" if x2.dtype == torch.float64 else "float32"), dims[0], dims[1],
                              1 if adjoint else 0, x2.shape[0], x2.data_ptr(), out.data_ptr(),
                              _stream(torch, x2.device))
    _lib.check(rc, "lmc_img_grad")
    return out


def proxdual(v, sigma, isotropic=True):
    """L21 / L1 dual prox (projection), in place on a (2, n) device tensor."""
    torch = _lib.require_cuda()
    lib = _lib.load()
    assert v.is_contiguous() and v.shape[0] == 2
    with torch.cuda.device(v.device):
        rc = lib.lmc_img_proxdual(_dt("float64" if v.dtype == torch.float64 else "float32"), 1 if isotropic else 0,
                                  float(sigma), v.shape[1], v.data_ptr(), _stream(torch, v.device))
    _lib.check(rc, "lmc_img_proxdual")
    return v


def tv_prox(dims, x, gamma, niter=10, rtol=1e-4, counters=None, sigma=1.0):
    """pyproximal.TV(dims, sigma, niter, rtol).prox with sigma*tau = gamma on device tensors (C,P) or (P,).
    ``sigma`` enters only the objective of the early-exit test (gamma * sigma * sum |grad sol|, see lmc_b200.h)."""
    torch = _lib.require_cuda()
    lib = _lib.load()
    P = dims[0] * dims[1]
    x2 = x.reshape(-1, P).contiguous()
    out = torch.empty_like(x2)
    dt = "float64" if x2.dtype == torch.float64 else "float32"
    nb = lib.lmc_img_myula_workspace_bytes(_dt(dt), dims[0], dims[1], x2.shape[0])
    ws = torch.empty(nb, dtype=torch.uint8, device=x2.device)
    with torch.cuda.device(x2.device):
        rc = lib.lmc_img_tv_prox(_dt(dt), dims[0], dims[1], float(gamma), float(sigma), int(niter), float(rtol), x2.shape[0],
                                 x2.data_ptr(), out.data_ptr(),
                                 counters.data_ptr() if counters is not None else None, ws.data_ptr(), nb,
                                 _stream(torch, x2.device))
    _lib.check(rc, "lmc_img_tv_prox")
    torch.cuda.current_stream(x2.device).synchronize()   # ws lifetime
    return out.reshape(x.shape)



class MoreauTerm:
    """The concave part ``-lamda * env_gamma(TV)`` of the reference's ``algs.L2_ncvx_tv`` (algs.py:22-291):
    ``kind`` "mc" = MC-TV (Op2 = Gradient; isotropic or not), "me" = ME-TV (inner TV(dims, 1, niter, rtol))."""

    def __init__(self, kind, lamda, gamma, isotropic=True, niter=50, rtol=1e-4):
        if kind not in ("mc", "me"):
            raise ValueError("kind must be 'mc' or 'me'")
        if kind == "me" and not isotropic:
            raise NotImplementedError("ME-TV with the 1-D TV of algs.py:170 has no CUDA path (the driver never uses it)")
        self.kind, self.lamda, self.gamma = kind, float(lamda), float(gamma)
        self.isotropic, self.niter, self.rtol = bool(isotropic), int(niter), float(rtol)

    def descriptor(self):
        """The C-ABI form (lmc_img_ncvx) handed to the sampler kernels for in-kernel evaluation."""
        d = _lib.ImgNcvx()
        d.kind, d.isotropic = (1 if self.kind == "mc" else 2), int(self.isotropic)
        d.lamda, d.gamma, d.niter, d.rtol = self.lamda, self.gamma, self.niter, self.rtol
        return d


def moreau_apply(dims, x, term: MoreauTerm, tau=None, counters=None):
    """``tau is None``: grad_moreau(x) (algs.py:270-282).  Otherwise the in-place pre-step of
    ``L2_ncvx_tv.prox``: x + tau*lamda*grad_moreau(x), with the reference's grouping of the scalars
    (algs.py:208-223).  Device tensors (C,P); returns a new tensor."""
    torch = _lib.require_cuda()
    lib = _lib.load()
    P = dims[0] * dims[1]
    x2 = x.reshape(-1, P).contiguous()
    out = torch.empty_like(x2)
    dt = _dt("float64" if x2.dtype == torch.float64 else "float32")
    g = term.gamma
    if term.kind == "mc":
        if tau is None:
            base, scale, post = 0.0, 1.0, (1.0 if term.isotropic else g)
        else:
            base, post = 1.0, 1.0
            scale = tau * term.lamda if term.isotropic else tau * term.lamda / g
        with torch.cuda.device(x2.device):
            rc = lib.lmc_img_mctv_grad(dt, dims[0], dims[1], g, 1 if term.isotropic else 0, base, scale, post,
                                       x2.shape[0], x2.data_ptr(), out.data_ptr(), _stream(torch, x2.device))
        _lib.check(rc, "lmc_img_mctv_grad")
    else:
        p = tv_prox(dims, x2, g, term.niter, term.rtol, counters)
        if tau is None:
            base, scale, den = 0.0, 0.0, g
        else:
            base, scale, den = 1.0, tau * term.lamda / g, 0.0
        with torch.cuda.device(x2.device):
            rc = lib.lmc_img_metv_combine(dt, x2.numel(), base, scale, den, x2.data_ptr(), p.data_ptr(),
                                          out.data_ptr(), _stream(torch, x2.device))
        _lib.check(rc, "lmc_img_metv_combine")
        torch.cuda.current_stream(x2.device).synchronize()   # p lifetime
    return out.reshape(x.shape)


def pd_arg(dims, x, y, tau, z=None):
    """x - tau * (G^T y + z): the argument ULPDA hands to proxf.prox (algs.py:436-446)."""
    torch = _lib.require_cuda()
    lib = _lib.load()
    out = torch.empty_like(x)
    dt = _dt("float64" if x.dtype == torch.float64 else "float32")
    with torch.cuda.device(x.device):
        rc = lib.lmc_img_pd_arg(dt, dims[0], dims[1], float(tau), x.shape[0], x.data_ptr(), y.data_ptr(),
                                z.data_ptr() if z is not None else None, out.data_ptr(), _stream(torch, x.device))
    _lib.check(rc, "lmc_img_pd_arg")
    return out


def philox_image_noise(dims, C_, N, seed, chain_offset=0, iter_offset=0, dtype="float32", device=None):
    """The N(0,1) draws the samplers generate in-kernel, as a (C,N,P) device tensor."""
    torch = _lib.require_cuda()
    lib = _lib.load()
    dev = torch.device(device) if device is not None else torch.device("cuda", torch.cuda.current_device())
    out = torch.empty((C_, N, dims[0] * dims[1]), dtype=_tdt(torch, dtype), device=dev)
    with torch.cuda.device(dev):
        rc = lib.lmc_img_noise(_dt(dtype), dims[0], dims[1], C_, N, seed & (2 ** 64 - 1), chain_offset, iter_offset,
                               out.data_ptr(), _stream(torch, dev))
    _lib.check(rc, "lmc_img_noise")
    return out


# --------------------------------------------------------------------------- sampler engine
class ImageStats:
    """Result of a many-chain run: device tensors (C,P) + executed inner-iteration counters."""

    def __init__(self, x, mean, m2, count, counters, dims):
        self.x, self.mean_dev, self.m2_dev, self.count, self.dims = x, mean, m2, count, dims
        self.counters = counters      # dict: tv_sweeps / lsqr_iterations / lsqr_solves

    @property
    def mean(self):
        """Posterior mean pooled over chains, (ny, nx) float64 ndarray."""
        return self.mean_dev.double().mean(dim=0).reshape(self.dims).cpu().numpy()

    def pooled_sums(self):
        """(count, sum x, sum x^2) over all chains and samples as device tensors (for all-reduce)."""
        m = self.mean_dev.double()
        n = float(self.count)
        s1 = (m * n).sum(dim=0)
        s2 = (self.m2_dev.double() + n * m * m).sum(dim=0) if self.m2_dev is not None else None
        return n * m.shape[0], s1, s2


class DeconvEngine:
    """Device-resident deconvolution problem + launches of the two image samplers.

    Parameters mirror what the reference builds in prox_lmc_deconv.py:55-125: the blur
    filter (padded as pylops does), the observation ``b`` and the L2 weight ``sigma_l2``."""

    def __init__(self, dims, h_padded, b, sigma_l2, dtype="float64", device=None):
        torch = _lib.require_cuda()
        self.torch = torch
        self.lib = _lib.load()
        self.dims = (int(dims[0]), int(dims[1]))
        self.P = self.dims[0] * self.dims[1]
        self.dtype = dtype
        self.tdt = _tdt(torch, dtype)
        self.dev = torch.device(device) if device is not None else torch.device("cuda", torch.cuda.current_device())
        self.spec = BlurSpec(self.dims, h_padded)
        self.sigma_l2 = float(sigma_l2)
        b = torch.as_tensor(np.asarray(b, dtype=np.float64) if not torch.is_tensor(b) else b)
        self.b = b.reshape(-1, self.P).to(self.dev, self.tdt).contiguous()
        self.counters = torch.zeros(8, dtype=torch.int64, device=self.dev)

    # ------------------------------------------------------------------ descriptors
    def _problem(self):
        pb = _lib.ImgProblem()
        self.spec.fill(pb)
        pb.sigma_l2 = self.sigma_l2
        pb.d_b = self.b.data_ptr()
        pb.b_chain_stride = 0 if self.b.shape[0] == 1 else self.P
        return pb

    def _buf(self, t, Cn, inner=None):
        """(Cn, [inner,] P) contiguous device tensor from user data (broadcast one image to all chains)."""
        torch = self.torch
        shape = (Cn, self.P) if inner is None else (Cn, inner, self.P)
        if t is None:
            return torch.zeros(shape, dtype=self.tdt, device=self.dev)
        t = torch.as_tensor(t).to(self.dev, self.tdt)
        t = t.reshape((-1,) + shape[1:])
        if t.shape[0] == 1 and Cn > 1:
            t = t.expand(shape)
        return t.contiguous().clone()

    def _io(self, Cn, N, noise, samples, mean, m2, burnin, stat_count0, iter_offset, ws):
        io = _lib.ImgIo()
        io.d_noise = noise.data_ptr() if noise is not None else None
        io.d_samples = samples.data_ptr() if samples is not None else None
        io.d_mean = mean.data_ptr() if mean is not None else None
        io.d_m2 = m2.data_ptr() if m2 is not None else None
        io.d_counters = self.counters.data_ptr()
        io.d_workspace = ws.data_ptr()
        io.workspace_bytes = ws.numel()
        io.burnin, io.stat_count0, io.iter_offset = int(burnin), int(stat_count0), int(iter_offset)
        return io

    def _noise(self, noise, Cn, N):
        if noise is None:
            return None
        t = self.torch.as_tensor(noise).to(self.dev, self.tdt).reshape(Cn, N, self.P).contiguous()
        return t

    def read_counters(self):
        c = self.counters.cpu().numpy()
        return {"tv_sweeps": int(c[0]), "lsqr_iterations": int(c[1]), "lsqr_solves": int(c[2])}

    # ------------------------------------------------------------------ MYULA
    def myula_band_supported(self):
        """True when the row-band MYULA kernel (csrc/img_myula_band.inl) covers this problem."""
        pb = self._problem()
        return bool(self.lib.lmc_img_myula_band_supported(C.byref(pb), _dt(self.dtype)))

    def onchip_supported(self):
        """True when the on-chip cluster kernel (csrc/img_myula_onchip.cu) covers this problem."""
        pb = self._problem()
        return bool(self.lib.lmc_img_myula_onchip_supported(C.byref(pb), _dt(self.dtype)))

    def myula(self, x, N, tau, gamma, tv_sigma, tv_niter=10, tv_rtol=1e-4, epsg=1.0, seed=0, chain_offset=0,
              noise=None, samples=None, mean=None, m2=None, burnin=0, stat_count0=0, iter_offset=0,
              engine="auto", ncvx=None, fgrad_extra=None, fgrad_scale=0.0):
        """Advances the (C,P) device tensor ``x`` in place by N MYULA iterations (algs.py:564-583).

        ``fgrad_extra`` / ``fgrad_scale``: a constant (C,P) term added to grad f in every iteration, scaled
        (alpha q of pyproximal.L2 / L2_ncvx_tv, algs.py:288-289); band / tile kernels, not combined with ``ncvx``.

        ``ncvx``: a MoreauTerm makes the data term the reference's L2_ncvx_tv (MC-TV / ME-TV): the band / tile
        kernels evaluate grad_moreau(x) in-kernel every iteration (MC-TV: a 7-point stencil in the gradient phase;
        ME-TV: an inner TV prox with the kernel's own TV machinery), grad f = sigma H^T(Hx - b) - lamda grad_moreau(x)
        (algs.py:283-291), so the N iterations stay one launch."""
        if ncvx is not None and engine == "onchip":
            raise NotImplementedError("the on-chip MYULA kernel has no non-convex data term; use engine='auto'")
        return self._myula_launch(x, N, tau, gamma, tv_sigma, tv_niter, tv_rtol, epsg, seed, chain_offset, noise,
                                  samples, mean, m2, burnin, stat_count0, iter_offset, engine, ncvx=ncvx,
                                  fgrad_extra=fgrad_extra, fgrad_scale=fgrad_scale)

    def _myula_launch(self, x, N, tau, gamma, tv_sigma, tv_niter=10, tv_rtol=1e-4, epsg=1.0, seed=0, chain_offset=0,
                      noise=None, samples=None, mean=None, m2=None, burnin=0, stat_count0=0, iter_offset=0,
                      engine="auto", fgrad_extra=None, fgrad_scale=0.0, ncvx=None):
        """One launch: N iterations of all chains.

        ``engine``: "onchip" = cluster kernel with the chain state in shared memory (fp32, images up to
        256 x 256), "band" = row-band kernel through HBM (H5 blur, even width; any height, fp32/fp64),
        "stream" = 32x32-tile kernel through HBM (any separable blur and size), "auto" = on-chip, else band,
        else tile.  All are CUDA; there is no CPU path."""
        torch, lib = self.torch, self.lib
        Cn = x.shape[0]
        pr = _lib.ImgMyulaParams()
        pr.tau, pr.gamma, pr.epsg = float(tau), float(gamma), float(epsg)
        pr.tv_sigma, pr.tv_rtol, pr.tv_niter = float(tv_sigma), float(tv_rtol), int(tv_niter)
        if engine not in ("auto", "onchip", "band", "stream"):
            raise ValueError("engine must be 'auto', 'onchip', 'band' or 'stream'")
        onchip = engine == "onchip" or (engine == "auto" and self.onchip_supported() and fgrad_extra is None
                                        and ncvx is None)
        band = engine == "band" or (engine == "auto" and not onchip and self.myula_band_supported())
        if onchip:
            ws = torch.empty(256, dtype=torch.uint8, device=self.dev)
        else:
            nb = lib.lmc_img_myula_workspace_bytes(_dt(self.dtype), self.dims[0], self.dims[1], Cn)
            ws = torch.empty(nb, dtype=torch.uint8, device=self.dev)
        nz = self._noise(noise, Cn, N)
        io = self._io(Cn, N, nz, samples, mean, m2, burnin, stat_count0, iter_offset, ws)
        io.d_x = x.data_ptr()
        if fgrad_extra is not None:
            io.d_fgrad_extra, io.fgrad_scale = fgrad_extra.data_ptr(), float(fgrad_scale)
        nc = ncvx.descriptor() if ncvx is not None else None
        if nc is not None:
            io.ncvx = C.pointer(nc)
        pb = self._problem()
        fn, name = ((lib.lmc_img_myula_run_onchip, "lmc_img_myula_run_onchip") if onchip
                    else (lib.lmc_img_myula_run_band, "lmc_img_myula_run_band") if band
                    else (lib.lmc_img_myula_run, "lmc_img_myula_run"))
        with torch.cuda.device(self.dev):
            rc = fn(C.byref(pb), C.byref(pr), _dt(self.dtype), Cn, int(N), seed & (2 ** 64 - 1),
                    int(chain_offset), C.byref(io), _stream(torch, self.dev))
        _lib.check(rc, name)
        return ws, nz   # keep alive until the caller synchronises

    # ------------------------------------------------------------------ ULPDA
    def ulpda(self, x, xhat, y, warm, N, tau, mu, theta=1.0, g_sigma=1.0, isotropic=True, gfirst=True,
              lsqr_iter_lim=50, warm_start=True, atol=1e-6, btol=1e-6, conlim=1e8, z=None, seed=0,
              chain_offset=0, noise=None, samples=None, ysamples=None, mean=None, m2=None, burnin=0,
              stat_count0=0, iter_offset=0, engine="auto", group=0, ncvx=None):
        """Advances (x, xhat, y, warm) in place by N ULPDA iterations (algs.py:432-467).

        ``ncvx``: a MoreauTerm makes proxf the reference's L2_ncvx_tv: every iteration the kernel forms the primal
        argument x - tau (G^T y + z) and adds tau*lamda*grad_moreau of it (algs.py:208-223) in-kernel (MC-TV: a
        7-point stencil phase; ME-TV: the inner TV prox by the tile TV pipeline), so the N iterations stay one
        launch; both gfirst orders."""
        return self._ulpda_launch(x, xhat, y, warm, N, tau, mu, theta, g_sigma, isotropic, gfirst, lsqr_iter_lim,
                                  warm_start, atol, btol, conlim, z, seed, chain_offset, noise, samples, ysamples,
                                  mean, m2, burnin, stat_count0, iter_offset, engine, group, ncvx=ncvx)

    def _ulpda_launch(self, x, xhat, y, warm, N, tau, mu, theta=1.0, g_sigma=1.0, isotropic=True, gfirst=True,
                      lsqr_iter_lim=50, warm_start=True, atol=1e-6, btol=1e-6, conlim=1e8, z=None, seed=0,
                      chain_offset=0, noise=None, samples=None, ysamples=None, mean=None, m2=None, burnin=0,
                      stat_count0=0, iter_offset=0, engine="auto", group=0, prox_arg=None, ncvx=None):
        """One launch: N iterations of all chains.

        ``engine``: "band" = row-band kernel (fused two-pass blur, chains advanced in L2-sized groups of
        ``group`` chains, 0 = automatic), "tile" = 32x32-tile kernel (any separable filter), "auto" = band
        when it applies.  Both are CUDA; there is no CPU path."""
        torch, lib = self.torch, self.lib
        Cn = x.shape[0]
        pr = _lib.ImgUlpdaParams()
        pr.tau, pr.mu, pr.theta, pr.g_sigma = float(tau), float(mu), float(theta), float(g_sigma)
        pr.g_isotropic, pr.gfirst = int(bool(isotropic)), int(bool(gfirst))
        pr.lsqr_iter_lim, pr.warm = int(lsqr_iter_lim), int(bool(warm_start))
        pr.lsqr_atol, pr.lsqr_btol, pr.lsqr_conlim = float(atol), float(btol), float(conlim)
        nc = ncvx.descriptor() if ncvx is not None else None
        nb = (lib.lmc_img_ulpda_workspace_bytes(_dt(self.dtype), self.dims[0], self.dims[1], Cn) if nc is None else
              lib.lmc_img_ulpda_workspace_bytes_ncvx(_dt(self.dtype), self.dims[0], self.dims[1], Cn, nc.kind))
        ws = torch.empty(nb, dtype=torch.uint8, device=self.dev)
        nz = self._noise(noise, Cn, N)
        io = self._io(Cn, N, nz, samples, mean, m2, burnin, stat_count0, iter_offset, ws)
        if nc is not None:
            io.ncvx = C.pointer(nc)
        io.d_x, io.d_xhat, io.d_y, io.d_warm = x.data_ptr(), xhat.data_ptr(), y.data_ptr(), warm.data_ptr()
        zt = None
        if z is not None:
            zt = torch.as_tensor(z).to(self.dev, self.tdt).reshape(self.P).contiguous()
            io.d_z = zt.data_ptr()
        io.d_ysamples = ysamples.data_ptr() if ysamples is not None else None
        if prox_arg is not None:
            io.d_prox_arg = prox_arg.data_ptr()
        pb = self._problem()
        if engine not in ("auto", "band", "tile"):
            raise ValueError("engine must be 'auto', 'band' or 'tile'")
        band = engine == "band" or (engine == "auto" and self.band_supported())
        with torch.cuda.device(self.dev):
            if band:
                rc = lib.lmc_img_ulpda_run_band(C.byref(pb), C.byref(pr), _dt(self.dtype), Cn, int(N),
                                                seed & (2 ** 64 - 1), int(chain_offset), int(group), C.byref(io),
                                                _stream(torch, self.dev))
            else:
                rc = lib.lmc_img_ulpda_run(C.byref(pb), C.byref(pr), _dt(self.dtype), Cn, int(N),
                                           seed & (2 ** 64 - 1), int(chain_offset), C.byref(io),
                                           _stream(torch, self.dev))
        _lib.check(rc, "lmc_img_ulpda_run_band" if band else "lmc_img_ulpda_run")
        return ws, nz, zt

    def band_supported(self):
        """True when the row-band ULPDA kernel (csrc/img_ulpda_band.inl) covers this problem."""
        pb = self._problem()
        return bool(self.lib.lmc_img_ulpda_band_supported(C.byref(pb), _dt(self.dtype)))

    # ------------------------------------------------------------------ L2 prox (LSQR)
    def l2_prox(self, x, tau, warm, iter_lim=50, atol=1e-6, btol=1e-6, conlim=1e8):
        """pyproximal.L2.prox(x, tau) for (C,P) device tensors; ``warm`` (C,P) is the LSQR start and is
        overwritten with the solution (the warm start of the next call)."""
        torch, lib = self.torch, self.lib
        x2 = x.reshape(-1, self.P).contiguous()
        Cn = x2.shape[0]
        out = torch.empty_like(x2)
        nb = lib.lmc_img_ulpda_workspace_bytes(_dt(self.dtype), self.dims[0], self.dims[1], Cn)
        ws = torch.empty(nb, dtype=torch.uint8, device=self.dev)
        pb = self._problem()
        with torch.cuda.device(self.dev):
            rc = lib.lmc_img_l2_prox(C.byref(pb), _dt(self.dtype), float(tau), int(iter_lim), float(atol),
                                     float(btol), float(conlim), Cn, x2.data_ptr(), warm.data_ptr(), out.data_ptr(),
                                     self.counters.data_ptr(), ws.data_ptr(), nb, _stream(torch, self.dev))
        _lib.check(rc, "lmc_img_l2_prox")
        torch.cuda.current_stream(self.dev).synchronize()
        return out.reshape(x.shape)

    # ------------------------------------------------------------------ per-sample metrics on the device
    def sample_metrics(self, x, xtrue=None, g_sigma=1.0, isotropic=True, images_per_chain=None):
        """Cost / error / SNR / PSNR / MSE of every image of ``x`` ((M,P) or (C,N,P) device tensor) without
        copying the images to the host: what the reference's per-iteration callback computes
        (prox_lmc_deconv.py:35-36,128-133) for f = L2 and g = g_sigma * TV (isotropic) or g_sigma * ||G x||_1.
        Returns a dict of float64 numpy arrays shaped like the leading dims of ``x``."""
        torch, lib = self.torch, self.lib
        lead = tuple(x.shape[:-1])
        if images_per_chain is None:
            images_per_chain = lead[1] if len(lead) == 2 else 1
        x2 = x.reshape(-1, self.P).contiguous()
        M = x2.shape[0]
        xt = None
        if xtrue is not None:
            xt = torch.as_tensor(xtrue).to(self.dev, self.tdt).reshape(self.P).contiguous()
        out = torch.empty((M, 4), dtype=torch.float64, device=self.dev)
        nb = lib.lmc_img_metrics_workspace_bytes(M)
        ws = torch.empty(nb, dtype=torch.uint8, device=self.dev)
        pb = self._problem()
        with torch.cuda.device(self.dev):
            rc = lib.lmc_img_sample_metrics(C.byref(pb), _dt(self.dtype), 1 if isotropic else 0, M,
                                            int(images_per_chain), x2.data_ptr(),
                                            xt.data_ptr() if xt is not None else None, out.data_ptr(), ws.data_ptr(),
                                            nb, _stream(torch, self.dev))
        _lib.check(rc, "lmc_img_sample_metrics")
        s = out.cpu().numpy().reshape(lead + (4,))     # synchronises; (M,4) doubles only
        res = {"f": self.sigma_l2 / 2.0 * s[..., 0], "g": float(g_sigma) * s[..., 1]}
        res["cost"] = res["f"] + res["g"]
        if xtrue is not None:
            with np.errstate(divide="ignore"):
                res["err"] = np.sqrt(s[..., 2])
                res["snr"] = 10.0 * np.log10(s[..., 3] / s[..., 2])
                res["mse"] = s[..., 2] / self.P
                res["psnr"] = 10.0 * np.log10(255.0 ** 2 / res["mse"])
        return res

    # ------------------------------------------------------------------ many-chain production runs
    def run_myula(self, chains, N, tau, gamma, tv_sigma, tv_niter=10, tv_rtol=1e-4, seed=0, chain_offset=0,
                  burnin=0, x0=None, second_moment=True, sync=True, engine="auto") -> ImageStats:
        torch = self.torch
        x = self._buf(x0, chains)
        mean = torch.zeros((chains, self.P), dtype=self.tdt, device=self.dev)
        m2 = torch.zeros_like(mean) if second_moment else None
        keep = self.myula(x, N, tau, gamma, tv_sigma, tv_niter, tv_rtol, seed=seed, chain_offset=chain_offset,
                          mean=mean, m2=m2, burnin=burnin, engine=engine)
        if sync:
            torch.cuda.current_stream(self.dev).synchronize()
            keep = None
        st = ImageStats(x, mean, m2, max(N - burnin, 0), None, self.dims)
        st._keep = keep
        return st

    def run_ulpda(self, chains, N, tau, mu, theta=1.0, g_sigma=1.0, isotropic=True, gfirst=False, lsqr_iter_lim=50,
                  seed=0, chain_offset=0, burnin=0, x0=None, second_moment=True, sync=True, engine="auto",
                  group=0) -> ImageStats:
        torch = self.torch
        x = self._buf(x0, chains)
        xhat = x.clone()
        y = torch.zeros((chains, 2, self.P), dtype=self.tdt, device=self.dev)
        warm = torch.zeros((chains, self.P), dtype=self.tdt, device=self.dev)
        mean = torch.zeros((chains, self.P), dtype=self.tdt, device=self.dev)
        m2 = torch.zeros_like(mean) if second_moment else None
        keep = self.ulpda(x, xhat, y, warm, N, tau, mu, theta, g_sigma, isotropic, gfirst, lsqr_iter_lim,
                          seed=seed, chain_offset=chain_offset, mean=mean, m2=m2, burnin=burnin, engine=engine,
                          group=group)
        if sync:
            torch.cuda.current_stream(self.dev).synchronize()
            keep = None
        st = ImageStats(x, mean, m2, max(N - burnin, 0), None, self.dims)
        st._keep = (keep, xhat, y, warm)
        return st
