"""2-Wasserstein evaluation of the mixture experiments on the GPU.

The reference measures its samplers with ``ot.emd2(a, b, ot.dist(Z1, Z[:k+1]))**.5`` between K true samples and
chain prefixes (lmc.py:386-416, lmc_laplace.py:426-451), a CPU network simplex.  Here the same transport problem
is solved on the device by ``lmc_w2_sinkhorn`` (log-domain Sinkhorn, epsilon scaling, fp64): an approximation
whose distance to the exact EMD is controlled by ``eps`` (default 1e-3 in squared-distance units): the GPU tests
hold it within 1e-3 of the exact linear programme on point sets up to 300 x 300 and, at the drivers' K = 5000,
within 5.6e-4 of the exact assignment (equal-size uniform sets: scipy ``linear_sum_assignment``).  Only the final
eps stage decides ``converged``; intermediate stages of the eps-scaling that reach ``max_iters`` are warm starts.
No CPU path: without the CUDA library / a device the calls raise.
"""
from __future__ import annotations

import ctypes as C
import warnings

import numpy as np

from . import _lib

EPS_DEFAULT = 1e-3


class SinkhornW2:
    """Workspace + device copies of a fixed set of ``true_samples``; ``distance(chain)`` evaluates W2 to it."""

    def __init__(self, true_samples, eps=EPS_DEFAULT, eps_scale=0.5, tol=1e-3, omega=1.0, max_iters=20000,
                 check_every=16, device="cuda:0"):
        self.torch = _lib.require_cuda()
        self.lib = _lib.load()
        self.device = self.torch.device(device)
        self.x = self._points(true_samples)
        self.eps, self.eps_scale, self.tol, self.omega = float(eps), float(eps_scale), float(tol), float(omega)
        self.max_iters, self.check_every = int(max_iters), int(check_every)
        self._ws = None
        self.last = None
        self.history = []          # one dict per distance() call (iterations, convergence)

    def _points(self, a):
        t = self.torch.as_tensor(np.ascontiguousarray(np.asarray(a, dtype=np.float64))) if not self.torch.is_tensor(a) else a
        if t.dim() != 2 or t.shape[0] < 1 or not 1 <= t.shape[1] <= 4:
            raise ValueError("point sets must be (n, d) with n >= 1 and d <= 4")
        return t.to(device=self.device, dtype=self.torch.float64).contiguous()

    def distance(self, chain, warm_start=False, eps_start=None):
        """W2 between the true samples and ``chain`` (m, d), uniform weights on both (``wass**.5`` of the reference).
        ``warm_start`` reuses the true-sample potentials of the previous call and starts at ``eps_start``."""
        y = self._points(chain)
        if y.shape[1] != self.x.shape[1]:
            raise ValueError("dimension mismatch")
        n, m = self.x.shape[0], y.shape[0]
        need = self.lib.lmc_w2_workspace_bytes(n, m)
        if self._ws is None or self._ws.numel() * 8 < need:
            if warm_start and self._ws is not None:
                keep = self._ws[:n].clone()
            else:
                keep, warm_start = None, False
            self._ws = self.torch.zeros((need + 7) // 8 + 2 * max(n, m), dtype=self.torch.float64, device=self.device)
            if keep is not None:
                self._ws[:n] = keep
        elif warm_start and self.last is None:
            warm_start = False
        res = _lib.W2Result()
        with self.torch.cuda.device(self.device):
            st = self.torch.cuda.current_stream().cuda_stream
            _lib.check(self.lib.lmc_w2_sinkhorn(self.x.data_ptr(), n, y.data_ptr(), m, self.x.shape[1], self.eps,
                                                float(eps_start) if (warm_start and eps_start) else (4 * self.eps if warm_start else 0.0),
                                                self.eps_scale, self.tol, self.omega, self.max_iters, self.check_every,
                                                1 if warm_start else 0, self._ws.data_ptr(), self._ws.numel() * 8,
                                                C.byref(res), st), "lmc_w2_sinkhorn")
        if not res.converged:
            warnings.warn("lmc_w2_sinkhorn stopped at max_iters=%d with potential change %.3g > tol*eps at the final "
                          "eps: the column marginals are not matched to tolerance (mass %.6f), so the value is "
                          "neither an upper nor a lower bound of the exact EMD" % (self.max_iters, res.err, res.mass))
        self.last = {"cost": res.cost, "w2": res.w2, "eps": res.eps, "err": res.err, "mass": res.mass,
                     "iters": int(res.iters), "converged": bool(res.converged)}
        self.history.append(self.last)
        return res.w2

    def curve(self, chain, interval=100, warm_start=False):
        """lmc.py:396-413: W2 to the prefixes ``chain[:k+1]`` for k = 1, 1+interval, ... < K (K = number of true
        samples), as the reference's ``wass_*_list``.  Every prefix is a cold start by default: measured at K = 5000,
        re-using the previous potentials did not lower the iteration count and twice ran into ``max_iters``."""
        y = self._points(chain)
        K = self.x.shape[0]
        out = []
        for k in range(1, min(K, y.shape[0])):
            if (k - 1) % interval == 0:
                out.append(self.distance(y[:k + 1], warm_start=warm_start and len(out) > 1))
        return out


def wasserstein2(true_samples, chain, eps=EPS_DEFAULT, **kw):
    """``ot.emd2(a, b, ot.dist(true_samples, chain))**.5`` with uniform a, b, evaluated on the GPU."""
    return SinkhornW2(true_samples, eps=eps, **kw).distance(chain)


def wasserstein2_curve(true_samples, chain, interval=100, eps=EPS_DEFAULT, **kw):
    return SinkhornW2(true_samples, eps=eps, **kw).curve(chain, interval=interval)


def wasserstein_curves(true_samples, chains, interval=100, eps=EPS_DEFAULT, **kw):
    """The evaluation loop of the mixture drivers (lmc.py:386-416, lmc_laplace.py:426-451): for every chain of the
    dict ``chains`` the list of W2 distances between the true samples and the prefixes ``Z[:k+1]``,
    k = 1, 1+interval, ... < K.  A chain shorter than K (MALA keeps accepted steps only) stops at its own length,
    as the ``if k < len(Z3)`` guard of lmc.py:408 does."""
    ev = SinkhornW2(true_samples, eps=eps, **kw)
    out = {}
    for name, z in chains.items():
        ev.last = None                       # potentials of another chain are not a warm start
        out[name] = ev.curve(z, interval=interval)
    return out
