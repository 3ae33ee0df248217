"""north_star: "prox_lmc_deconv.py ... run unchanged against it".  In the build container (reference tree mounted, no
GPU) the UNMODIFIED driver is imported with `algs`, `pylops`, `pyproximal`, `skimage` bound to this package's
stand-ins (compat.install()), and every attribute the driver's source takes from those modules is resolved, and
every call it makes into `algs` / AdaptivePrimalDual is checked against the signatures of the replacements."""
import ast
import importlib
import inspect
import os
import sys

import pytest

pytestmark = pytest.mark.refonly
REF = "/root/reference"


@pytest.fixture(scope="module")
def driver():
    from lmc_atomi_b200 import compat
    saved = dict(sys.modules)
    compat.install(force=True)
    sys.path.insert(0, REF)
    try:
        sys.modules.pop("prox_lmc_deconv", None)
        mod = importlib.import_module("prox_lmc_deconv")      # the reference's own file, unmodified
        assert os.path.samefile(mod.__file__, os.path.join(REF, "prox_lmc_deconv.py"))
        yield mod
    finally:
        sys.path.remove(REF)
        for k in list(sys.modules):
            if k not in saved:
                del sys.modules[k]
        sys.modules.update(saved)


def _chains(tree, roots):
    """dotted attribute chains rooted at one of `roots` in the driver's source, e.g. pyproximal.optimization.primaldual.X"""
    out = set()
    for node in ast.walk(tree):
        if isinstance(node, ast.Attribute):
            parts, cur = [], node
            while isinstance(cur, ast.Attribute):
                parts.append(cur.attr)
                cur = cur.value
            if isinstance(cur, ast.Name) and cur.id in roots:
                out.add((cur.id,) + tuple(reversed(parts)))
    return out


def test_unmodified_driver_imports_and_binds_to_the_cuda_modules(driver):
    import lmc_atomi_b200.algs as ours
    assert driver.algs is ours
    assert driver.pyproximal.TV.__module__ == "lmc_atomi_b200.compat"
    tree = ast.parse(open(os.path.join(REF, "prox_lmc_deconv.py")).read())
    for chain in sorted(_chains(tree, {"algs", "pylops", "pyproximal", "io", "data"})):
        obj = getattr(driver, chain[0])
        for attr in chain[1:]:
            assert hasattr(obj, attr), ".".join(chain)
            obj = getattr(obj, attr)


def test_every_call_of_the_driver_binds_to_our_signatures(driver):
    """Each call `algs.X(...)`, `pyproximal.X(...)`, `pylops.X(...)` in the driver is bound (inspect.Signature.bind)
    to the replacement with the driver's own positional / keyword layout."""
    tree = ast.parse(open(os.path.join(REF, "prox_lmc_deconv.py")).read())
    checked = 0
    for node in ast.walk(tree):
        if not isinstance(node, ast.Call) or not isinstance(node.func, ast.Attribute):
            continue
        parts, cur = [], node.func
        while isinstance(cur, ast.Attribute):
            parts.append(cur.attr)
            cur = cur.value
        if not (isinstance(cur, ast.Name) and cur.id in ("algs", "pylops", "pyproximal")):
            continue
        obj = getattr(driver, cur.id)
        for attr in reversed(parts):
            obj = getattr(obj, attr)
        sig = inspect.signature(obj)
        args = [object()] * len(node.args)
        kwargs = {k.arg: object() for k in node.keywords if k.arg is not None}
        sig.bind(*args, **kwargs)          # raises TypeError when the driver's call does not fit
        checked += 1
    assert checked >= 40       # 9 sampler + 9 MAP calls x 2 branches, the operator constructors


def test_headless_main_keeps_the_driver_keywords(driver):
    from lmc_atomi_b200.prox_lmc_deconv import prox_lmc_deconv as ours
    ref = inspect.signature(driver.prox_lmc_deconv).parameters
    mine = inspect.signature(ours).parameters
    for name, p in ref.items():
        assert name in mine and mine[name].default == p.default, name
    assert list(mine)[:len(ref)] == list(ref)
