"""Parameter-carrying stand-ins for the third-party objects the reference's image driver
hands to ``algs`` (prox_lmc_deconv.py:55-125): ``pylops.signalprocessing.Convolve2D``,
``pylops.Gradient``, ``pylops.Identity``, ``pyproximal.L2 / L1 / L21 / TV``.

Neither library exists in this image, so a headless run of the driver needs
objects with the same constructor arguments and attribute names.  They hold parameters; the
few methods the driver's ``callback`` / ``show`` paths call (``matvec``, ``rmatvec``,
``__call__``, ``prox``, ``grad``, ``proxdual``) are served by the CUDA unit operators of
include/lmc_b200.h (host array in, host array out) -- there is no numpy implementation here.
``algs`` reads real pylops/pyproximal objects by the same attribute names (duck typing).
"""
from __future__ import annotations

import numpy as np

from . import _lib, imaging


def _dev(a, shape=None):
    torch = _lib.require_cuda()
    t = torch.as_tensor(np.ascontiguousarray(a, dtype=np.float64)).cuda()
    return t if shape is None else t.reshape(shape)


class _LinOp:
    def __mul__(self, x):
        return self.matvec(x)

    __matmul__ = __mul__

    @property
    def H(self):
        return _Adjoint(self)


class _Adjoint(_LinOp):
    def __init__(self, op):
        self.op = op
        self.shape = op.shape[::-1]

    def matvec(self, x):
        return self.op.rmatvec(x)

    def rmatvec(self, x):
        return self.op.matvec(x)


class Convolve2D(_LinOp):
    """pylops.signalprocessing.Convolve2D(dims, h, offset) -- SURVEY 8a I1."""
    explicit = False

    def __init__(self, dims, h, offset=(0, 0), axes=(-2, -1), method="fft", dtype="float64", name="C"):
        self.dims = tuple(int(d) for d in dims)
        self.offset = tuple(int(o) for o in offset)
        self.h = imaging.pad_filter(h, self.offset)
        n = self.dims[0] * self.dims[1]
        self.shape = (n, n)
        self.dtype = np.dtype(dtype)
        self._spec = imaging.BlurSpec(self.dims, self.h)

    def _apply(self, x, adjoint):
        x = np.asarray(x)
        out = imaging.blur(self._spec, _dev(x, (1, -1)), adjoint=adjoint).cpu().numpy()
        return out.reshape(x.shape)

    def matvec(self, x):
        return self._apply(x, False)

    def rmatvec(self, x):
        return self._apply(x, True)


class Gradient(_LinOp):
    """pylops.Gradient(dims, sampling=1, edge=False, kind='forward') -- SURVEY 8a I2."""
    explicit = False

    def __init__(self, dims, sampling=1, edge=False, kind="forward", dtype="float64", name="G"):
        if kind != "forward" or edge or np.any(np.asarray(sampling, dtype=float) != 1.0):
            raise NotImplementedError("the CUDA path covers Gradient(kind='forward', edge=False, sampling=1)")
        self.dims = tuple(int(d) for d in dims)
        self.sampling, self.edge, self.kind = sampling, edge, kind
        n = self.dims[0] * self.dims[1]
        self.shape = (2 * n, n)
        self.dtype = np.dtype(dtype)

    def matvec(self, x):
        return imaging.grad(self.dims, _dev(x, (1, -1))).cpu().numpy().ravel()

    def rmatvec(self, v):
        return imaging.grad(self.dims, _dev(v, (1, 2, -1)), adjoint=True).cpu().numpy().ravel()


class Identity(_LinOp):
    explicit = False

    def __init__(self, N, M=None, dtype="float64"):
        self.shape = (int(N) if M is None else int(M), int(N))
        self.dtype = np.dtype(dtype)

    def matvec(self, x):
        return x

    rmatvec = matvec


# --------------------------------------------------------------------------- proximal operators
class ProxOperator:
    def __init__(self, Op=None, hasgrad=False):
        self.Op = Op
        self.hasgrad = hasgrad

    def proxdual(self, x, tau):
        """Moreau identity (pyproximal ProxOperator._proxdual_moreau)."""
        return x - tau * self.prox(x / tau, 1.0 / tau)


class L1(ProxOperator):
    """pyproximal.L1(sigma): anisotropic norm; dual prox = clip."""

    def __init__(self, sigma=1.0, g=None):
        super().__init__(None, False)
        if g is not None:
            raise NotImplementedError("L1 with a shift vector g is not on the reference's path")
        self.sigma, self.g = sigma, g

    def __call__(self, x):
        return float(self.sigma * _dev(x).abs().sum().item())

    def proxdual(self, x, tau):
        v = _dev(x, (2, -1)).contiguous()
        return imaging.proxdual(v, self.sigma, isotropic=False).cpu().numpy().reshape(np.shape(x))

    def prox(self, x, tau):
        x = np.asarray(x, dtype=np.float64)
        return x - tau * self.proxdual(x / tau, 1.0 / tau)


class L21(ProxOperator):
    """pyproximal.L21(ndim, sigma): sum of pixel-wise 2-norms; dual prox = projection."""

    def __init__(self, ndim, sigma=1.0):
        super().__init__(None, False)
        if int(ndim) != 2:
            raise NotImplementedError("the CUDA path covers L21(ndim=2)")
        self.ndim, self.sigma = int(ndim), sigma

    def __call__(self, x):
        v = _dev(x, (self.ndim, -1))
        return float(self.sigma * v.pow(2).sum(dim=0).sqrt().sum().item())

    def proxdual(self, x, tau):
        v = _dev(x, (2, -1)).contiguous()
        return imaging.proxdual(v, self.sigma, isotropic=True).cpu().numpy().reshape(np.shape(x))

    def prox(self, x, tau):
        x = np.asarray(x, dtype=np.float64)
        return x - tau * self.proxdual(x / tau, 1.0 / tau)


class TV(ProxOperator):
    """pyproximal.TV(dims, sigma, niter, rtol): isotropic TV, prox by dual FISTA."""

    def __init__(self, dims, sigma=1.0, niter=10, rtol=1e-4, **kwargs):
        super().__init__(None, False)
        self.dims = tuple(int(d) for d in dims)
        if len(self.dims) != 2:
            raise NotImplementedError("the CUDA path covers 2-D TV")
        self.ndim, self.sigma, self.niter, self.rtol = 2, sigma, int(niter), rtol

    def __call__(self, x):
        g = imaging.grad(self.dims, _dev(x, (1, -1)))[0]
        return float(self.sigma * g.pow(2).sum(dim=0).sqrt().sum().item())

    def prox(self, x, tau):
        x = np.asarray(x)
        out = imaging.tv_prox(self.dims, _dev(x, (1, -1)), self.sigma * tau, self.niter, self.rtol, sigma=self.sigma)
        return out.cpu().numpy().reshape(x.shape)


class L2(ProxOperator):
    """pyproximal.L2(Op, b, sigma, niter, warm): sigma/2 ||Op x - b||^2."""

    def __init__(self, Op=None, b=None, q=None, sigma=1.0, alpha=1.0, qgrad=True, niter=10, x0=None,
                 warm=True, densesolver=None, kwargs_solver=None):
        super().__init__(Op, True)
        if densesolver is not None:
            raise NotImplementedError("L2 with a dense solver (explicit operators) has no CUDA path")
        if Op is None or b is None:
            raise NotImplementedError("the CUDA path covers L2 with an operator and data")
        self.b, self.q, self.sigma, self.alpha, self.qgrad = b, q, sigma, alpha, qgrad
        self.niter, self.x0, self.warm, self.count = (niter if callable(niter) else int(niter)), x0, warm, 0
        self.kwargs_solver = {} if kwargs_solver is None else kwargs_solver
        self._eng = None

    def _engine(self):
        if self._eng is None:
            self._eng = imaging.DeconvEngine(self.Op.dims, self.Op.h, np.asarray(self.b), self.sigma, dtype="float64")
        return self._eng

    def __call__(self, x):
        eng = self._engine()
        r = imaging.blur(eng.spec, _dev(x, (1, -1))) - eng.b
        f = float(self.sigma / 2.0 * r.pow(2).sum().item())
        if self.q is not None:      # pyproximal.L2: + alpha q^T x
            f += float(self.alpha * np.dot(np.asarray(self.q, dtype=np.float64).ravel(), np.asarray(x, dtype=np.float64).ravel()))
        return f

    def grad(self, x):
        eng = self._engine()
        r = imaging.blur(eng.spec, _dev(x, (1, -1))) - eng.b
        g = (self.sigma * imaging.blur(eng.spec, r, adjoint=True)).cpu().numpy().reshape(np.shape(x))
        if self.q is not None and self.qgrad:
            g = g + self.alpha * np.asarray(self.q, dtype=np.float64).reshape(np.shape(x))
        return g

    def prox(self, x, tau):
        eng = self._engine()
        torch = eng.torch
        self.count += 1      # pyproximal's _increment_count; a callable niter sees the incremented count
        lim = int(self.niter(self.count)) if callable(self.niter) else self.niter
        warm = _dev(self.x0, (1, -1)).clone() if (self.x0 is not None and self.warm) else \
            torch.zeros((1, eng.P), dtype=torch.float64, device=eng.dev)
        kw = self.kwargs_solver
        xin = _dev(x, (1, -1))
        if self.q is not None:      # y = x + tau OpTb - tau alpha q
            xin = xin - float(tau) * self.alpha * _dev(np.asarray(self.q, dtype=np.float64), (1, -1))
        out = eng.l2_prox(xin, tau, warm, iter_lim=lim, atol=kw.get("atol", 1e-6),
                          btol=kw.get("btol", 1e-6), conlim=kw.get("conlim", 1e8))
        out = out.cpu().numpy().reshape(np.shape(x))
        if self.warm:
            self.x0 = out
        return out


# pyproximal.optimization.primaldual.AdaptivePrimalDual (MAP branch of the driver, prox_lmc_deconv.py:138-445)
import types as _types

from . import primaldual as _primaldual

optimization = _types.SimpleNamespace(primaldual=_types.SimpleNamespace(
    AdaptivePrimalDual=_primaldual.AdaptivePrimalDual))


# ---------------------------------------------------------------------------------------------------------------
# Running the reference's own driver unchanged: `prox_lmc_deconv.py` does `import pylops, pyproximal, algs` plus
# skimage / fire / plotting imports (prox_lmc_deconv.py:4-32).  install() registers this package's stand-ins under
# those names, so `python -c "from lmc_atomi_b200 import compat; compat.install(); import prox_lmc_deconv"` binds the
# unmodified driver to the CUDA path.  Plot / CLI libraries become inert (their calls do nothing).
def _skimage_modules():
    """skimage.io.imread / data.camera / metrics.{peak_signal_noise_ratio, mean_squared_error} as the driver uses
    them (prox_lmc_deconv.py:25-27, 45-48, 128-133)."""
    sk = _types.ModuleType("skimage")

    def imread(path):
        from PIL import Image
        return np.asarray(Image.open(path))

    def camera():
        raise NotImplementedError("skimage.data.camera() (the 512x512 cameraman image) is not available offline; "
                                  "use --image=einstein")

    def mean_squared_error(a, b):
        return float(np.mean((np.asarray(a, dtype=np.float64) - np.asarray(b, dtype=np.float64)) ** 2))

    def peak_signal_noise_ratio(image_true, image_test, data_range=None):
        if data_range is None:       # skimage: the dtype range of image_true (uint8 -> 255) when it is non-negative
            t = np.asarray(image_true)
            data_range = 255. if t.dtype == np.uint8 else float(t.max() - t.min())
        return float(10 * np.log10(data_range ** 2 / mean_squared_error(image_true, image_test)))

    sk.io = _types.ModuleType("skimage.io")
    sk.io.imread = imread
    sk.data = _types.ModuleType("skimage.data")
    sk.data.camera = camera
    sk.metrics = _types.ModuleType("skimage.metrics")
    sk.metrics.peak_signal_noise_ratio = peak_signal_noise_ratio
    sk.metrics.mean_squared_error = mean_squared_error
    return {"skimage": sk, "skimage.io": sk.io, "skimage.data": sk.data, "skimage.metrics": sk.metrics}


class _InertModule(_types.ModuleType):
    """fire / matplotlib / seaborn / scienceplots / fastprogress: every attribute is a do-nothing callable."""

    class _Obj:
        def __call__(self, *a, **k):
            return _InertModule._Obj()

        def __getattr__(self, name):
            if name.startswith("__"):
                raise AttributeError(name)
            return _InertModule._Obj()

        def __getitem__(self, k):
            return _InertModule._Obj()

        def __iter__(self):
            return iter((_InertModule._Obj(), _InertModule._Obj()))

    def __getattr__(self, name):
        if name.startswith("__"):
            raise AttributeError(name)
        return _InertModule._Obj()


def install(force=False):
    """Registers `algs`, `pylops`, `pyproximal`, `skimage` (and inert plot / CLI modules) in sys.modules unless the
    real library is importable (``force`` replaces it anyway).  Returns the list of names it registered."""
    import importlib
    import sys
    from . import algs as _algs
    this = sys.modules[__name__]
    pl = _types.ModuleType("pylops")
    pl.Gradient, pl.Identity = Gradient, Identity
    pl.signalprocessing = _types.ModuleType("pylops.signalprocessing")
    pl.signalprocessing.Convolve2D = Convolve2D
    pp = _types.ModuleType("pyproximal")
    for name in ("L1", "L2", "L21", "TV", "ProxOperator"):
        setattr(pp, name, getattr(this, name))
    pp.optimization = _types.ModuleType("pyproximal.optimization")
    pp.optimization.primaldual = _types.ModuleType("pyproximal.optimization.primaldual")
    pp.optimization.primaldual.AdaptivePrimalDual = _primaldual.AdaptivePrimalDual
    mods = {"algs": _algs, "pylops": pl, "pylops.signalprocessing": pl.signalprocessing, "pyproximal": pp,
            "pyproximal.optimization": pp.optimization, "pyproximal.optimization.primaldual": pp.optimization.primaldual}
    mods.update(_skimage_modules())
    for name in ("fire", "matplotlib", "matplotlib.pyplot", "matplotlib.cm", "seaborn", "scienceplots", "fastprogress"):
        mods[name] = _InertModule(name)
    mods["matplotlib"].pyplot, mods["matplotlib"].cm = mods["matplotlib.pyplot"], mods["matplotlib.cm"]
    done = []
    for name, mod in mods.items():
        root = name.split(".")[0]
        if not force and name != "algs":
            try:
                importlib.import_module(root)
                if not isinstance(sys.modules.get(root), (_InertModule,)) and root not in done and \
                        sys.modules[root] is not mods.get(root):
                    continue       # the real library is installed: leave it
            except ImportError:
                pass
        sys.modules[name] = mod
        done.append(name)
    return done
