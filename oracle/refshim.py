"""TEST INFRASTRUCTURE ONLY -- never imported by the product path.

Import shim that lets the *unmodified* reference sampler classes be executed in
the build container (where /root/reference is mounted read-only).  The reference
driver scripts import plotting / CLI / optimal-transport libraries at module top
(lmc.py:5-21, lmc_laplace.py:5-29, prox_lmc.py:5-26, prox.py:6) that are absent
from this image; none of them is touched by the sampler methods themselves, so
they are replaced by inert stubs (``fastprogress.progress_bar`` := identity).

This module is used by ``oracle/make_golden.py`` to produce the committed golden
vectors in ``tests/golden/`` and by the (container-only) tests that pin the
numpy/C restatements against the real reference.  It cannot travel to the GPU
box (no /root/reference there); nothing in ``-m gpu`` tests, ``smoke()`` or
``bench.py`` imports it.
"""
from __future__ import annotations

import contextlib
import io
import os
import sys
import types

REFERENCE_ROOT = os.environ.get("LMC_REFERENCE_ROOT", "/root/reference")

_STUBS = [
    "fire", "fastprogress", "ot", "ot.plot", "matplotlib", "matplotlib.pyplot",
    "matplotlib.cm", "seaborn", "scienceplots", "pylops",
]


class _Inert(types.ModuleType):
    """Module whose every attribute is an inert callable/namespace."""

    def __getattr__(self, name):  # pragma: no cover - trivial
        if name.startswith("__"):
            raise AttributeError(name)
        obj = _InertObj()
        setattr(self, name, obj)
        return obj


class _InertObj:
    def __call__(self, *a, **k):
        return _InertObj()

    def __getattr__(self, name):
        if name.startswith("__"):
            raise AttributeError(name)
        return _InertObj()

    def __getitem__(self, k):
        return _InertObj()

    def __iter__(self):
        return iter(())


def reference_available() -> bool:
    return os.path.isfile(os.path.join(REFERENCE_ROOT, "lmc.py"))


def install_stubs() -> None:
    for name in _STUBS:
        if name in sys.modules and not isinstance(sys.modules[name], _Inert):
            continue  # a real library is present: use it
        try:
            if name not in sys.modules:
                __import__(name)
                continue
        except Exception:
            pass
        mod = _Inert(name)
        mod.__path__ = []  # behave like a package so "import a.b" works
        sys.modules[name] = mod
    fp = sys.modules["fastprogress"]
    if isinstance(fp, _Inert):
        fp.progress_bar = lambda it, *a, **k: it
    mpl = sys.modules["matplotlib"]
    if isinstance(mpl, _Inert):
        mpl.pyplot = sys.modules["matplotlib.pyplot"]
        mpl.cm = sys.modules["matplotlib.cm"]
    ot = sys.modules["ot"]
    if isinstance(ot, _Inert):
        ot.plot = sys.modules["ot.plot"]


def load_reference():
    """Return the reference modules (lmc, lmc_laplace, prox_lmc, prox)."""
    if not reference_available():
        raise RuntimeError(
            f"reference tree not found at {REFERENCE_ROOT}; golden vectors in "
            "tests/golden/ are the travelling form of the reference outputs")
    install_stubs()
    if REFERENCE_ROOT not in sys.path:
        sys.path.insert(0, REFERENCE_ROOT)
    import importlib
    mods = {}
    for name in ("prox", "lmc", "lmc_laplace", "prox_lmc"):
        mods[name] = importlib.import_module(name)
    return mods


@contextlib.contextmanager
def quiet():
    """The reference samplers print a banner per call; swallow it."""
    buf = io.StringIO()
    with contextlib.redirect_stdout(buf):
        yield


# ---- the reference drivers' own parameter tables (the spec for inputs) -------------
# lmc.py:205-235, lmc_laplace.py:230-251, prox_lmc.py:274-307
import numpy as np  # noqa: E402


def gaussian_mixture_params(n=5):
    mu = [np.array([0., 0.]), np.array([-2., 3.]), np.array([2., -3.]),
          np.array([3., 3.]), np.array([-2., -2.])]
    Sg = [np.array([[1., -0.5], [-0.5, 1.]]), np.array([[0.5, 0.2], [0.2, 0.7]]),
          np.array([[0.5, 0.1], [0.1, 0.9]]), np.array([[0.8, 0.02], [0.02, 0.3]]),
          np.array([[1.2, 0.05], [0.05, 0.8]])]
    idx = {1: [0], 2: [0, 1], 3: [0, 1, 2], 4: [1, 2, 3, 4], 5: [0, 1, 2, 3, 4]}[n]
    return [mu[i] for i in idx], [Sg[i] for i in idx], np.ones(n) / n


def laplacian_mixture_params(n=5, alpha=0.5):
    mus, _, om = gaussian_mixture_params(n)
    return mus, np.ones(n) * alpha, om
