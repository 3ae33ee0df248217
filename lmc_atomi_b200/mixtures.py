"""Host side of the mixture-chain path: descriptors, launches, result objects.

The three reference-named sampler classes (``lmc.LangevinMonteCarlo``,
``lmc_laplace.LangevinMonteCarloLaplacian``, ``prox_lmc.ProximalLangevinMonteCarlo``)
derive from :class:`MixtureSampler`.  Two ways to run a sampler:

* the reference call surface -- ``sampler.ula(gamma)`` etc. -- runs ONE chain of K
  steps on the GPU in fp64 with the reference's own numpy random stream
  (``default_rng(seed)``; lmc.py:96-100), so the returned array is comparable
  sample-by-sample with the reference's;
* ``sampler.sample_chains(alg, gamma, chains=C, ...)`` runs C independent chains with
  in-kernel Philox noise and returns reduced statistics (:class:`ChainStats`).

PyTorch is used only to own device buffers and streams.
"""
from __future__ import annotations

import ctypes as C

import numpy as np
from numpy.random import default_rng

from . import _lib

HIST_BINS = 100          # lmc.py:295 (hist2d bins=100)
HIST_RANGE = (-5.0, 5.0)  # lmc.py:199-200 plotting grid

_ADJUSTED = ("mala", "mymala")


def reference_noise(seed, K, d=2, uniforms=False):
    """Arrays holding the reference's random stream: theta0, xi (K,d)[, u (K,)].

    ``default_rng(seed).standard_normal(d)`` for the start (lmc.py:96-97), then per step
    one d-vector of N(0,1) (``rng.multivariate_normal(0, I)`` yields the same bits) and,
    for MALA/MYMALA, one ``rng.random()`` after it (lmc.py:124-128).
    """
    rng = default_rng(seed)
    theta0 = rng.standard_normal(d)
    if not uniforms:
        return theta0, rng.standard_normal((K, d)), None
    xi = np.empty((K, d))
    u = np.empty(K)
    for k in range(K):
        xi[k] = rng.standard_normal(d)
        u[k] = rng.random()
    return theta0, xi, u


class ChainStats:
    """Reduced output of a many-chain run (sums are over emitted samples).

    The statistics stay on the device (``moments_dev`` (6,) float64, ``hist_dev``
    (bins,bins) int64) until one of the host properties is read, so a caller can
    all-reduce them across GPUs first (see ``dist.allreduce_stats_``)."""

    def __init__(self, alg, chains, steps, dtype, moments_dev, hist_dev, theta, naccept=None,
                 hist_range=HIST_RANGE, keepalive=None):
        self.alg, self.chains, self.steps, self.dtype = alg, chains, steps, dtype
        self.moments_dev, self.hist_dev = moments_dev, hist_dev
        self.theta, self.naccept = theta, naccept      # torch tensors (C,2), (C,) int32
        self.hist_range = hist_range
        self.extra = {"moments": moments_dev, "hist": hist_dev}
        self._keepalive = keepalive
        self._host = None

    def _m(self):
        if self._host is None:
            self._host = self.moments_dev.cpu().numpy()   # synchronises
            self._keepalive = None
        return self._host

    @property
    def count(self):
        return float(self._m()[0])

    @property
    def sum1(self):
        return self._m()[1:3].copy()

    @property
    def sum2(self):
        m = self._m()
        return np.array([[m[3], m[4]], [m[4], m[5]]])

    @property
    def hist(self):
        return None if self.hist_dev is None else self.hist_dev.cpu().numpy()

    @property
    def mean(self):
        return self.sum1 / self.count

    @property
    def second_moment(self):
        return self.sum2 / self.count

    @property
    def cov(self):
        m = self.mean
        return self.second_moment - np.outer(m, m)

    @property
    def accept_rate(self):
        if self.naccept is None:
            return 1.0
        return float(self.naccept.double().mean().item()) / self.steps

    def merge_(self, other: "ChainStats"):
        """Adds another shard's statistics (same device)."""
        self.moments_dev += other.moments_dev.to(self.moments_dev.device)
        if self.hist_dev is not None and other.hist_dev is not None:
            self.hist_dev += other.hist_dev.to(self.hist_dev.device)
        self.chains += other.chains
        self._host = None
        return self


def _mat4(a):
    a = np.asarray(a, dtype=np.float64).reshape(2, 2)
    return (C.c_double * 4)(*a.ravel())


class MixtureSampler:
    """Common engine.  Subclasses provide ``_target()`` and the reference-named methods."""

    _valid_algs: tuple = ()
    _hess_shift = 0.05

    # ------------------------------------------------------------------ descriptors
    def _target(self) -> _lib.MixTarget:
        raise NotImplementedError

    def _step(self, alg, gamma, M=None, Q=None, t=100, beta=None, sigma=None):
        from scipy.linalg import sqrtm
        if alg not in self._valid_algs:
            raise ValueError(f"{type(self).__name__} has no sampler {alg!r}")
        s = _lib.MixStep()
        s.alg = _lib.ALGS[alg]
        s.gamma = float(gamma)
        s.t = int(t)
        s.hess_shift = self._hess_shift
        if alg in ("pula", "ppula"):
            if M is None:
                raise ValueError(f"{alg} needs the preconditioner M")
            M = np.asarray(M, dtype=np.float64)
            s.M = _mat4(M)
            s.sqrtM = _mat4(np.real(sqrtm(M)))      # lmc.py:145
        if alg == "ppula":
            if Q is None:
                raise ValueError("ppula needs the metric Q")
            Q = np.asarray(Q, dtype=np.float64)
            if int(t) < 1:
                raise ValueError("ppula needs t >= 1 (the reference reads `u` after the loop)")
            s.Q = _mat4(Q)
            s.Qinv = _mat4(np.linalg.inv(Q))        # prox_lmc.py:176
            rho = 1 / np.linalg.norm(Q, ord=2)      # prox_lmc.py:166-168
            eps = max(min(1, rho) - 1e-5, 1e-9)
            s.eta = rho - eps
        if alg in ("mla", "lbmumla"):
            if beta is None:
                raise ValueError(f"{alg} needs beta")
            s.beta = (C.c_double * 2)(*np.asarray(beta, dtype=np.float64))
        if alg == "lbmumla":
            if sigma is None:
                raise ValueError("lbmumla needs sigma")
            s.sigma = (C.c_double * 2)(*np.asarray(sigma, dtype=np.float64))
        return s

    # ------------------------------------------------------------------ launches
    def _launch(self, alg, step, dtype, Cn, K, seed, chain_offset, theta, *, noise=None, unif=None,
                traj=None, accept=None, naccept=None, moments=None, hist=None, init_from_rng=False,
                hist_every=0, burnin=0, hist_bins=HIST_BINS, hist_range=HIST_RANGE):
        torch = _lib.require_cuda()
        lib = _lib.load()
        dev = theta.device
        ws_bytes = lib.lmc_mix_workspace_bytes(Cn)
        # one workspace per (device, stream), grown on demand and reused by later launches: launches on one stream are
        # ordered, so the next launch may overwrite it (statistics are reduced into the caller's buffers by then)
        key = (dev, torch.cuda.current_stream(dev).cuda_stream)
        cache = self.__dict__.setdefault("_ws_cache", {})
        ws = cache.get(key)
        if ws is None or ws.numel() < ws_bytes:
            ws = cache[key] = torch.empty(ws_bytes, dtype=torch.uint8, device=dev)
        io = _lib.MixIo()
        io.d_theta = theta.data_ptr()
        io.d_noise = noise.data_ptr() if noise is not None else None
        io.d_uniforms = unif.data_ptr() if unif is not None else None
        io.d_traj = traj.data_ptr() if traj is not None else None
        io.d_accept = accept.data_ptr() if accept is not None else None
        io.d_naccept = naccept.data_ptr() if naccept is not None else None
        io.d_moments = moments.data_ptr() if moments is not None else None
        io.d_hist = hist.data_ptr() if hist is not None else None
        io.d_workspace = ws.data_ptr()
        io.workspace_bytes = ws_bytes
        io.init_from_rng = 1 if init_from_rng else 0
        io.hist_bins = hist_bins
        io.hist_lo, io.hist_hi = hist_range
        io.hist_every = hist_every
        io.burnin = burnin
        tgt = self._target()
        with torch.cuda.device(dev):
            stream = torch.cuda.current_stream(dev).cuda_stream
            rc = lib.lmc_mix_run(C.byref(tgt), C.byref(step), _lib.F64 if dtype == "float64" else _lib.F32,
                                 Cn, K, seed & (2**64 - 1), chain_offset, C.byref(io), C.c_void_p(stream))
        _lib.check(rc, f"lmc_mix_run[{alg}]")
        return ws  # keep alive until the caller synchronises

    def _device(self, device=None):
        torch = _lib.require_cuda()
        if device is None:
            return torch.device("cuda", torch.cuda.current_device())
        return torch.device(device)

    def run_with_noise(self, alg, gamma, theta0, xi, u=None, dtype="float64", device=None, **params):
        """Validation mode: C chains driven by caller-supplied noise.

        theta0 (C,2), xi (C,K,2), u (C,K) for MALA/MYMALA.  Returns
        (traj (C,K,2) float64 ndarray, accept (C,K) bool ndarray or None)."""
        torch = _lib.require_cuda()
        dev = self._device(device)
        tdt = torch.float64 if dtype == "float64" else torch.float32
        theta0 = np.atleast_2d(np.asarray(theta0, dtype=np.float64))
        xi = np.asarray(xi, dtype=np.float64)
        if xi.ndim == 2:
            xi = xi[None]
        Cn, K = xi.shape[0], xi.shape[1]
        adj = alg in _ADJUSTED
        if adj and u is None:
            raise ValueError(f"{alg} needs uniforms")
        step = self._step(alg, gamma, **params)
        th = torch.from_numpy(np.ascontiguousarray(theta0)).to(dev, tdt)
        nz = torch.from_numpy(np.ascontiguousarray(xi)).to(dev, tdt)
        un = torch.from_numpy(np.ascontiguousarray(np.asarray(u, dtype=np.float64).reshape(Cn, K))).to(dev, tdt) if adj else None
        traj = torch.empty((Cn, K, 2), dtype=tdt, device=dev)
        acc = torch.empty((Cn, K), dtype=torch.uint8, device=dev) if adj else None
        ws = self._launch(alg, step, dtype, Cn, K, 0, 0, th, noise=nz, unif=un, traj=traj, accept=acc)
        out = traj.double().cpu().numpy()
        a = acc.cpu().numpy().astype(bool) if adj else None
        del ws
        return out, a

    def _reference_call(self, alg, gamma, **params):
        """The reference call surface: one chain, K = self.n steps, seed = self.seed,
        the reference's numpy random stream, fp64 arithmetic."""
        adj = alg in _ADJUSTED
        th0, xi, u = reference_noise(self.seed, self.n, self.d, uniforms=adj)
        traj, acc = self.run_with_noise(alg, gamma, th0[None], xi[None], None if u is None else u[None],
                                        dtype=getattr(self, "dtype", "float64"), **params)
        if adj:
            sel = traj[0][acc[0]]
            return sel, int(acc[0].sum())
        return traj[0]

    def sample_chains(self, alg, gamma, chains, K=None, dtype="float32", seed=None, chain_offset=0,
                      burnin=0, hist=True, hist_every=0, theta0=None, device=None, _sync=True, **params) -> ChainStats:
        """Production mode: ``chains`` independent chains x K steps in one persistent launch.

        Chain c draws from Philox4x32-10 keyed by ``seed`` with counter words
        (step, stream, c + chain_offset): results are independent of the sharding."""
        torch = _lib.require_cuda()
        dev = self._device(device)
        K = self.n if K is None else int(K)
        seed = self.seed if seed is None else int(seed)
        tdt = torch.float64 if dtype == "float64" else torch.float32
        step = self._step(alg, gamma, **params)
        adj = alg in _ADJUSTED
        if theta0 is None:
            th = torch.empty((chains, 2), dtype=tdt, device=dev)
        else:
            th = torch.as_tensor(theta0).to(dev, tdt).contiguous().clone()
        mom = torch.zeros(6, dtype=torch.float64, device=dev)
        hs = torch.zeros((HIST_BINS, HIST_BINS), dtype=torch.int64, device=dev) if hist else None
        nacc = torch.zeros(chains, dtype=torch.int32, device=dev) if adj else None
        ws = self._launch(alg, step, dtype, chains, K, seed, chain_offset, th, naccept=nacc, moments=mom,
                          hist=hs, init_from_rng=theta0 is None, hist_every=hist_every, burnin=burnin)
        st = ChainStats(alg=alg, chains=chains, steps=K, dtype=dtype, moments_dev=mom, hist_dev=hs, theta=th,
                        naccept=nacc, keepalive=ws)
        if _sync:
            st._m()
        return st
